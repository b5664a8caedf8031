"""Networks and losses -- drop-in for the reference's ``stransfer/network.py``.

Same classes, methods, argument meaning and state-dict keys as the reference
(file:line citations are into /root/reference/stransfer/network.py); every numeric
op runs on the hand-written sm_100a kernels of libstransfer_b200.so through
``torch.autograd.Function`` wrappers.  There is no CPU / cuDNN / cuBLAS fallback:
tensors must live on a CUDA device and the library must load, or the call raises.
"""
import os
import shutil

import torch
import torch.nn as nn
import torch.optim as optim
from tqdm import tqdm

from . import c_logging, constants, ops, vgg_loss
from ._lib import StransferError

LOGGER = c_logging.get_logger()


# --------------------------------------------------------------------------
# helpers (reference network.py:25-76)
# --------------------------------------------------------------------------
def get_tensorboard_writer(path: str):
    """network.py:25-35 -- tensorboardX is replaced by torch.utils.tensorboard when
    available, else by a no-op writer (observability sidecar, out of scope)."""
    shutil.rmtree(path, ignore_errors=True)
    try:
        from torch.utils.tensorboard import SummaryWriter
        return SummaryWriter(path)
    except Exception:                                     # pragma: no cover
        class _Null:
            def add_scalar(self, *a, **k): pass
            def add_image(self, *a, **k): pass
        return _Null()


def adaptive_torch_load(weights_path: str):
    """network.py:38-50"""
    return torch.load(weights_path, map_location='cuda' if constants.DEVICE.type == 'cuda' else 'cpu')


def _load_latest_model_weigths(model_name: str, style_name: str, models_path='data/models/'):
    """network.py:53-76 (same lexicographic 'latest' rule and AssertionError)."""
    models_path = os.path.join(constants.PROJECT_ROOT_PATH, models_path)
    try:
        latest_weight_name = sorted([x for x in os.listdir(models_path)
                                     if x.startswith(model_name) and style_name in x])[-1]
    except (IndexError, FileNotFoundError):
        LOGGER.critical('There are no weights for the specified model name (%s) '
                        'and style (%s). In the specified path: %s', model_name, style_name, models_path)
        raise AssertionError('There are no weights for the specified model name and style.')
    return adaptive_torch_load(models_path + latest_weight_name)


# --------------------------------------------------------------------------
# stand-alone loss modules (reference network.py:79-201)
# --------------------------------------------------------------------------
class _GramMSE(torch.autograd.Function):
    """loss = mean((gram(x) - T)^2) for a NCHW fp32 feature map."""

    @staticmethod
    def forward(ctx, x, target):
        f = ops.from_nchw(x)
        G = ops.gram(f)
        loss = torch.zeros(1, dtype=torch.float32, device=x.device)
        ctx.diff = ops.style_mse_fwd(G, target, loss)
        ctx.f, ctx.C = f, x.shape[1]
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        f = ctx.f
        _buf, wp = ops.style_seed(ctx.diff, f.H * f.W, g.reshape(1).contiguous().float())
        out, _, _ = ops.conv2d(f, None, w_ptrs=wp, w_img_rows=f.C)
        return ops.to_nchw(out, ctx.C), None


class _MSE(torch.autograd.Function):
    """mean((x - t)^2) (relu=False) or the FeatureReconstructionLoss form."""

    @staticmethod
    def forward(ctx, x, target_split, feature_form):
        f = ops.from_nchw(x)
        loss = torch.zeros(1, dtype=torch.float32, device=x.device)
        mse = None
        if feature_form:
            mse = torch.zeros(1, dtype=torch.float32, device=x.device)
            ops.mse_fwd(f, target_split, mse)
            ops.mse_fwd(f, target_split, loss, feature_form=True)
        else:
            ops.mse_fwd(f, target_split, loss)
        ctx.f, ctx.t, ctx.mse, ctx.C = f, target_split, mse, x.shape[1]
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        n = float(ctx.f.buf[0].numel())
        scale = 2.0 / n if ctx.mse is None else 4.0 / (n * n)
        out = ops.mse_bwd(ctx.f, ctx.t, g.reshape(1).contiguous().float(), scale, mul=ctx.mse)
        return ops.to_nchw(out, ctx.C), None, None


def _check_channels(t):
    if t.shape[1] % 8 != 0:
        raise StransferError('loss modules need a channel count that is a multiple of 8')


class StyleLoss(nn.Module):
    """network.py:79-131"""

    def __init__(self, target: torch.Tensor):
        super().__init__()
        self.set_target(target)

    def gram_matrix(self, input: torch.Tensor) -> torch.Tensor:
        """network.py:92-108: bmm(F, F^T) / (C*H*W) -> [B, C, C]"""
        ops._require_cuda(input, 'StyleLoss.gram_matrix')
        _check_channels(input)
        return ops.gram(ops.from_nchw(input.detach()))

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        """network.py:110-123: stores the loss in ``self.loss`` and returns the input"""
        ops._require_cuda(input, 'StyleLoss.forward')
        _check_channels(input)
        self.loss = _GramMSE.apply(input.contiguous().float(), self.target)
        return input

    def set_target(self, target: torch.Tensor):
        """network.py:125-131"""
        self.target = self.gram_matrix(target).detach()


class ContentLoss(nn.Module):
    """network.py:134-164"""
    _feature_form = False

    def __init__(self, target: torch.Tensor):
        super().__init__()
        self.set_target(target)

    def set_target(self, target):
        """Accepts a NCHW tensor (reference API) or an internal Split buffer."""
        if isinstance(target, ops.Split):
            self._target_split, self._target = target, None
        else:
            ops._require_cuda(target, type(self).__name__ + '.set_target')
            _check_channels(target)
            self._target = target.detach()
            self._target_split = ops.from_nchw(self._target)

    @property
    def target(self):
        if self._target is None:
            self._target = ops.to_nchw(self._target_split)
        return self._target

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        ops._require_cuda(input, type(self).__name__ + '.forward')
        self.loss = _MSE.apply(input.contiguous().float(), self._target_split, self._feature_form)
        return input


class FeatureReconstructionLoss(ContentLoss):
    """network.py:167-201: (mse)^2 / (B*C*H*W)"""
    _feature_form = True


# --------------------------------------------------------------------------
# VGG-19 loss network (reference network.py:204-458)
# --------------------------------------------------------------------------
_VGG_CFG_E = [64, 64, 'M', 128, 128, 'M', 256, 256, 256, 256, 'M',
              512, 512, 512, 512, 'M', 512, 512, 512, 512, 'M']
# torchvision vgg19().features indices of the five convs that are ever executed
_VGG_CONV_IDX = [0, 2, 5, 7, 10]


def load_vgg19_prefix_weights(seed=None):
    """[(weight, bias)] x5 for conv1_1, conv1_2, conv2_1, conv2_2, conv3_1 on DEVICE.

    The reference uses ``torchvision.models.vgg19(pretrained=True)`` (network.py:246).
    With ``seed`` given, or when the pretrained checkpoint cannot be obtained
    (offline), a seeded random-init VGG-19 (torchvision's kaiming-normal fan_out
    init) is used instead and a warning is logged."""
    import math
    if seed is None:
        try:
            import torchvision
            feats = torchvision.models.vgg19(weights='IMAGENET1K_V1').features
            return [(feats[i].weight.detach().to(constants.DEVICE), feats[i].bias.detach().to(constants.DEVICE))
                    for i in _VGG_CONV_IDX]
        except Exception as exc:                       # no network / no torchvision
            LOGGER.warning('pretrained VGG-19 unavailable (%s); using seeded random init', type(exc).__name__)
            seed = 1
    g = torch.Generator().manual_seed(seed)
    out = []
    for cin, cout, _ in vgg_loss.PREFIX:
        w = torch.randn(cout, cin, 3, 3, generator=g) * math.sqrt(2.0 / (cout * 9))
        out.append((w.to(constants.DEVICE), torch.zeros(cout, device=constants.DEVICE)))
    return out


class StyleNetwork(nn.Module):
    """network.py:204-458.  ``vgg_weights`` (extra, optional) supplies the five
    (weight, bias) pairs explicitly -- tests and benchmarks pass seeded ones."""

    content_layers = ['Conv2d_4']
    style_layers = ['Conv2d_1', 'Conv2d_2', 'Conv2d_3', 'Conv2d_4', 'Conv2d_5']
    feature_loss_layers = ['ReLU_4']

    # layer name -> (piece index, tap index, relu applied) for the supported taps
    _SUPPORTED = {'Conv2d_1': 0, 'Conv2d_2': 1, 'Conv2d_3': 2, 'Conv2d_4': 3, 'Conv2d_5': 4}

    def __init__(self, style_image: torch.Tensor, content_image: torch.Tensor = None, vgg_weights=None):
        super().__init__()
        if (self.content_layers != ['Conv2d_4'] or self.feature_loss_layers != ['ReLU_4'] or
                self.style_layers != ['Conv2d_%d' % i for i in range(1, 6)]):
            raise NotImplementedError('the B200 loss network implements the reference default tap set '
                                      '(style Conv2d_1..5, content Conv2d_4, feature ReLU_4)')
        ops._require_cuda(style_image, 'StyleNetwork')
        dev = style_image.device
        if content_image is None:
            content_image = torch.zeros([1, 3, 256, 256], device=dev)      # network.py:241-243
        weights = vgg_weights if vgg_weights is not None else load_vgg19_prefix_weights()
        weights = [(w.to(dev), b.to(dev)) for w, b in weights]
        self.vgg = vgg_loss.VGGPrefix(weights)
        self.compute_feature_loss = True

        # the same 7 pieces the reference cuts (network.py:257-314), kept as plain
        # containers for introspection; compute goes through self.vgg
        self.net_pieces = self._build_pieces(weights)

        # targets: network.py:280-307
        with torch.no_grad():
            staps = self.vgg.taps(style_image)
            ctaps = self.vgg.taps(content_image, upto=4)
        self.style_losses = []
        for i, tap in enumerate(staps):
            node = StyleLoss.__new__(StyleLoss)
            nn.Module.__init__(node)
            node.target = ops.gram(tap)
            self.style_losses.append([node, [0, 1, 2, 3, 5][i]])
        self.content_losses = [[ContentLoss(ctaps[3]), 3]]
        self.feature_losses = [[FeatureReconstructionLoss(self._relu_split(ctaps[3])), 4]]
        self._content_key = None

    @staticmethod
    def _relu_split(tap):
        """relu of a Split (the ReLU_4 feature target).  Off the hot path (targets only)."""
        out = tap.like()
        rest = tap.float().clamp_(min=0)
        for p in range(out.planes):
            plane = rest.to(torch.bfloat16)
            out.buf[p].copy_(plane)
            rest = rest - plane.float()
        return out

    @staticmethod
    def _build_pieces(weights):
        """Pieces 0..5 hold the layers the losses execute; piece 6 (conv3_2 .. pool5,
        never run by the reference) is left as an empty placeholder."""
        layers, i, cin = [], 0, 3
        for v in _VGG_CFG_E[:7]:
            if v == 'M':
                layers.append(('MaxPool2d_%d' % i, nn.MaxPool2d(2, 2)))
                continue
            i += 1
            conv = nn.Conv2d(cin, v, 3, padding=1)
            conv.weight.data, conv.bias.data = weights[i - 1]
            layers.append(('Conv2d_%d' % i, conv))
            if i < 5:
                layers.append(('ReLU_%d' % i, nn.ReLU(inplace=False)))
            cin = v
        cuts = {'Conv2d_1', 'Conv2d_2', 'Conv2d_3', 'Conv2d_4', 'ReLU_4', 'Conv2d_5'}
        pieces = [nn.Sequential()]
        for name, layer in layers:
            pieces[-1].add_module(name, layer)
            if name in cuts:
                pieces.append(nn.Sequential())
        return pieces

    def run_through_pieces(self, input_g: torch.Tensor, until=-1) -> torch.Tensor:
        """network.py:316-340.  Pieces 0..5 (everything a loss node reads) run on the
        kernels; the unused tail (piece 6, conv3_2..pool5) is not implemented."""
        ops._require_cuda(input_g, 'StyleNetwork.run_through_pieces')
        if until == -1 or until > 5:
            raise NotImplementedError('pieces beyond Conv2d_5 are never executed by the reference losses')
        tap_idx = {0: 0, 1: 1, 2: 2, 3: 3, 4: 3, 5: 4}[until]
        with torch.no_grad():
            tap = self.vgg.taps(input_g, upto=tap_idx + 1)[tap_idx]
            out = ops.to_nchw(tap)
        return out.clamp_(min=0) if until == 4 else out

    def get_total_current_content_loss(self, weight=1) -> torch.Tensor:
        """network.py:342-348"""
        return weight * torch.stack([x[0].loss for x in self.content_losses]).sum()

    def get_total_current_feature_loss(self, weight=1) -> torch.Tensor:
        """network.py:350-356"""
        return weight * torch.stack([x[0].loss for x in self.feature_losses]).sum()

    def get_total_current_style_loss(self, weight=1) -> torch.Tensor:
        """network.py:358-364"""
        return weight * torch.stack([x[0].loss for x in self.style_losses]).sum()

    def _set_content(self, content_image):
        # the reference recomputes the content/feature targets on every call
        # (network.py:378-382); identical tensors (same storage, same version) are cached
        key = (content_image.data_ptr(), content_image._version, tuple(content_image.shape))
        if key != self._content_key:
            with torch.no_grad():
                tap = self.vgg.taps(content_image.detach(), upto=4)[3]
            self.content_losses[0][0].set_target(tap)
            self.feature_losses[0][0].set_target(self._relu_split(tap) if self.compute_feature_loss else tap)
            self._content_key = key

    def forward(self, input_image: torch.Tensor, content_image=None, style_image=None) -> None:
        """network.py:366-401: evaluates every loss node; results live in ``node.loss``."""
        ops._require_cuda(input_image, 'StyleNetwork.forward')
        if content_image is not None:
            self._set_content(content_image)
        if style_image is not None:
            # bug-compatible with network.py:390-394: the *content* image sets the style targets
            with torch.no_grad():
                for (node, _), tap in zip(self.style_losses, self.vgg.taps(content_image)):
                    node.target = ops.gram(tap)
        losses = vgg_loss.vgg_losses(input_image, self.vgg, [n[0].target for n in self.style_losses],
                                     self.content_losses[0][0]._target_split, self.compute_feature_loss)
        for i, (node, _) in enumerate(self.style_losses):
            node.loss = losses[i]
        self.content_losses[0][0].loss = losses[5]
        self.feature_losses[0][0].loss = losses[6]

    def get_content_optimizer(self, input_img, optt=optim.Adam):
        """network.py:403-409"""
        return optt([input_img.requires_grad_()])

    def train_gatys(self, style_image: torch.Tensor, content_image: torch.Tensor,
                    steps=550, style_weight=100_000, content_weight=1) -> torch.Tensor:
        """network.py:411-458.  Same L-BFGS loop; the per-closure ``LOGGER.info`` of the
        loss (a device sync every closure, network.py:453) is emitted every 50th closure."""
        assert isinstance(style_image, torch.Tensor), 'Images need to be already loaded'
        assert isinstance(content_image, torch.Tensor), 'Images need to be already loaded'
        input_image = content_image.clone()
        optimizer = self.get_content_optimizer(input_image, optt=optim.LBFGS)
        self.compute_feature_loss = False           # never read by this loop
        n_closure = [0]
        for step in tqdm(range(steps)):
            def closure():
                optimizer.zero_grad()
                self(input_image, content_image)
                style_loss = self.get_total_current_style_loss(weight=style_weight)
                content_loss = self.get_total_current_content_loss(weight=content_weight)
                total_loss = style_loss + content_loss
                total_loss.backward()
                n_closure[0] += 1
                if n_closure[0] % 50 == 0:
                    LOGGER.info('Loss: %s', total_loss)
                return total_loss
            optimizer.step(closure)
        self.compute_feature_loss = True
        return input_image
