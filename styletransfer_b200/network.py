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

from . import c_logging, constants, dist as st_dist, ops, optim as st_optim, transform, vgg_loss
from ._lib import StransferError

LOGGER = c_logging.get_logger()


# --------------------------------------------------------------------------
# helpers (reference network.py:25-76)
# --------------------------------------------------------------------------
def get_tensorboard_writer(path: str):
    """network.py:25-35 -- tensorboardX is replaced by torch.utils.tensorboard when
    available, else by a no-op writer (observability sidecar, out of scope)."""
    class _Null:
        def add_scalar(self, *a, **k): pass
        def add_image(self, *a, **k): pass
    if path is None:                                      # ranks other than 0 of a data-parallel run
        return _Null()
    shutil.rmtree(path, ignore_errors=True)
    try:
        from torch.utils.tensorboard import SummaryWriter
        return SummaryWriter(path)
    except Exception:                                     # pragma: no cover
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


class GraphedClosure:
    """A closure (forward + backward that writes the ``.grad`` of its leaves) captured ONCE into a
    CUDA graph and replayed afterwards: the ~50 kernel launches of a Gatys closure then cost one
    ``cudaGraphLaunch`` instead of ~50 Python->ctypes round trips.
    ``fn`` must leave the leaves' ``.grad`` as None before its backward so that the gradient
    buffers are created inside the graph's memory pool and rewritten in place by every replay;
    everything ``fn`` reads (image, targets, weights) must live at fixed addresses."""

    def __init__(self, fn, leaves=(), warmup=3):
        self.fn = fn                             # keeps every tensor the closure captured (baked into the graph) alive
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(warmup):              # first-launch setup (attributes, workspaces) stays outside
                fn()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        from ._lib import lib
        n0 = lib().st_launch_count()
        self.graph = torch.cuda.CUDAGraph()
        # same stream as the warm-up: the leaves' AccumulateGrad nodes were created on it
        with torch.cuda.graph(self.graph, stream=side):
            self.out = fn()
        self.launches = int(lib().st_launch_count() - n0)     # kernels of ours per replay
        # the scratch buffers the captured kernels were given are baked into the graph as raw addresses: keep them
        # alive even when a later, larger request replaces them in the grow-only cache (ops.workspace)
        self._workspaces = list(ops._ws_cache.values())
        self.leaves = [(leaf, leaf.grad) for leaf in leaves]   # gradient buffers inside the graph's pool

    def __call__(self):
        for leaf, grad in self.leaves:
            if leaf.grad is not grad:            # an eager backward or zero_grad() replaced it
                leaf.grad = grad
        self.graph.replay()
        return self.out


def cuda_graphs_enabled():
    return os.environ.get('STRANSFER_CUDA_GRAPH', '1') != '0'


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
        _keep, wp = ops.style_seed(ctx.diff[0], ctx.diff[1], f.H * f.W, g.reshape(1).contiguous().float(), f.fmt)
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
def load_vgg19_prefix_weights(seed=None, n_convs=16):
    """[(weight, bias)] for the first ``n_convs`` convs of torchvision's vgg19().features on DEVICE (all 16 by
    default; the reference's default loss layers execute the first five).

    The reference uses ``torchvision.models.vgg19(pretrained=True)`` (network.py:246) and fails when the
    checkpoint cannot be obtained; so does this function (``StransferError``).  A seeded random-init VGG-19
    (torchvision's kaiming-normal fan_out init) is an explicit opt-in only: the ``seed`` argument, or the
    ``STRANSFER_VGG_SEED`` environment variable (tests, benchmarks and offline smoke runs)."""
    import math
    if seed is None and os.environ.get('STRANSFER_VGG_SEED'):
        seed = int(os.environ['STRANSFER_VGG_SEED'])
        LOGGER.warning('STRANSFER_VGG_SEED=%d: the loss network is a seeded RANDOM-INIT VGG-19, not the pretrained one', seed)
    if seed is None:
        try:
            import torchvision
            feats = torchvision.models.vgg19(weights='IMAGENET1K_V1').features
        except Exception as exc:                       # no network / no torchvision
            raise StransferError('the pretrained VGG-19 weights could not be loaded (%s: %s); pass vgg_weights= / seed=, '
                                 'or set STRANSFER_VGG_SEED to opt in to a seeded random-init loss network'
                                 % (type(exc).__name__, exc)) from exc
        convs = [m for m in feats if isinstance(m, nn.Conv2d)][:n_convs]
        return [(m.weight.detach().to(constants.DEVICE), m.bias.detach().to(constants.DEVICE)) for m in convs]
    g = torch.Generator().manual_seed(seed)
    out = []
    for cin, cout, _ in vgg_loss.BLOCKS[:n_convs]:
        w = torch.randn(cout, cin, 3, 3, generator=g) * math.sqrt(2.0 / (cout * 9))
        out.append((w.to(constants.DEVICE), torch.zeros(cout, device=constants.DEVICE)))
    return out


class StyleNetwork(nn.Module):
    """network.py:204-458.  ``vgg_weights`` (extra, optional) supplies the (weight, bias) pairs of the first n VGG-19
    convs explicitly -- tests and benchmarks pass seeded ones; n must cover the deepest loss layer.

    As in the reference the loss layers are CLASS attributes a user may edit (a subclass, or assignment before
    construction): any ``Conv2d_k`` / ``ReLU_k`` / ``MaxPool2d_k`` name of ``vgg19().features`` (k = 1..16 counts the
    convolutions, network.py:257-270).  The network is cut into pieces after every loss layer exactly like the
    reference's builder (network.py:257-314); the compute runs through ONE fused pass (vgg_loss.py)."""

    content_layers = ['Conv2d_4']
    style_layers = ['Conv2d_1', 'Conv2d_2', 'Conv2d_3', 'Conv2d_4', 'Conv2d_5']
    feature_loss_layers = ['ReLU_4']

    def __init__(self, style_image: torch.Tensor, content_image: torch.Tensor = None, vgg_weights=None):
        super().__init__()
        ops._require_cuda(style_image, 'StyleNetwork')
        dev = style_image.device
        if content_image is None:
            content_image = torch.zeros([1, 3, 256, 256], device=dev)      # network.py:241-243
        weights = vgg_weights if vgg_weights is not None else load_vgg19_prefix_weights()
        weights = [(w.to(dev), b.to(dev)) for w, b in weights]
        self.vgg = vgg_loss.VGGPrefix(weights)
        self.compute_feature_loss = True

        # the pieces the reference cuts (network.py:257-314): a new piece starts after every loss layer; kept as
        # plain containers for introspection (and run_through_pieces' indexing), compute goes through self.vgg
        self.net_pieces, piece_of = self._build_pieces(weights)
        self._piece_last = {}                            # piece index -> last layer point inside pieces[: index + 1]
        last = None
        for i, piece in enumerate(self.net_pieces):
            names = [n for n, _ in piece.named_children()]
            if names:
                last = vgg_loss.parse_layer(names[-1])
            self._piece_last[i] = last

        def points(names, what):
            pts = []
            for name in names:
                pt = vgg_loss.parse_layer(name)
                if name not in piece_of:
                    raise StransferError(f'{what} layer {name} needs VGG conv {pt[1]} but only {len(weights)} conv weights were given')
                pts.append((name, pt))
            # the reference appends a layer's node when it reaches the layer: network order, not list order
            return sorted(pts, key=lambda t: vgg_loss.layer_names().index(t[0]))
        style_pts = points(self.style_layers, 'style')
        content_pts = points(self.content_layers, 'content')
        feature_pts = points(self.feature_loss_layers, 'feature')

        # targets: network.py:280-307 (style nodes from the style image, content / feature nodes from the content image)
        self._spec = vgg_loss.LossSpec([[pt, None] for _, pt in style_pts], [[pt, None] for _, pt in content_pts],
                                       [[pt, None] for _, pt in feature_pts])
        self.style_losses, self.content_losses, self.feature_losses = [], [], []
        for name, _ in style_pts:
            node = StyleLoss.__new__(StyleLoss)
            nn.Module.__init__(node)
            self.style_losses.append([node, piece_of[name]])
        for name, _ in content_pts:
            node = ContentLoss.__new__(ContentLoss)
            nn.Module.__init__(node)
            self.content_losses.append([node, piece_of[name]])
        for name, _ in feature_pts:
            node = FeatureReconstructionLoss.__new__(FeatureReconstructionLoss)
            nn.Module.__init__(node)
            self.feature_losses.append([node, piece_of[name]])
        self._set_style_targets(style_image)
        self._content_key = None
        self._set_content(content_image)

    @staticmethod
    def _relu_split(tap):
        """relu of a Split (the value of a ReLU_k layer from the conv output).  Off the hot path (targets only)."""
        return tap.like().set_from_float(tap.float().clamp_(min=0))

    def _build_pieces(self, weights):
        """Generic over the three layer lists (network.py:257-314): every layer of the available VGG blocks goes into
        the current piece under the reference's name; after a layer that carries a loss node a new piece starts.
        Returns (pieces, {loss layer name: index of its piece})."""
        cuts = set(self.content_layers) | set(self.style_layers) | set(self.feature_loss_layers)
        pieces, piece_of = [nn.Sequential()], {}
        for k, (cin, cout, pool) in enumerate(vgg_loss.BLOCKS[:len(weights)], start=1):
            conv = nn.Conv2d(cin, cout, 3, padding=1)
            conv.weight.data, conv.bias.data = weights[k - 1]
            layers = [(f'Conv2d_{k}', conv), (f'ReLU_{k}', nn.ReLU(inplace=False))]
            if pool:
                layers.append((f'MaxPool2d_{k}', nn.MaxPool2d(2, 2)))
            for name, layer in layers:
                pieces[-1].add_module(name, layer)
                if name in cuts:
                    piece_of[name] = len(pieces) - 1
                    pieces.append(nn.Sequential())
        return pieces, piece_of

    def run_through_pieces(self, input_g: torch.Tensor, until=-1) -> torch.Tensor:
        """network.py:316-340: the input through pieces[: until + 1] (all pieces for -1), on the kernels."""
        ops._require_cuda(input_g, 'StyleNetwork.run_through_pieces')
        n = len(self.net_pieces)
        idx = n - 1 if until == -1 else min(until, n - 1)
        point = self._piece_last[idx]
        if point is None:
            return input_g
        with torch.no_grad():
            return self.vgg.value_at(input_g, point)

    def _set_style_targets(self, image):
        """Gram targets of the style nodes from ``image`` (network.py:297-301)."""
        pts = self._spec.points(style=True, mse=False)
        if not pts:
            return
        with torch.no_grad():
            vals, _ = self.vgg.run(image.detach(), pts, keep_taps=False)
            for (node, _), entry in zip(self.style_losses, self._spec.style):
                node.target = ops.gram(vals[entry[0]])
                entry[1] = node.target

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
        """The reference recomputes the content / feature targets on every call (network.py:378-382).  The targets
        of THE SAME tensor object at the same autograd version are reused (the Gatys loop passes one fixed content
        image to every closure); the cache holds a strong reference to that tensor, so its storage cannot be
        recycled for another image while the key is alive, and the loops that feed new batches reset it."""
        key = self._content_key
        if key is not None and key[0] is content_image and key[1] == content_image._version and key[2] == self.compute_feature_loss:
            return
        pts = self._spec.points(with_feature=self.compute_feature_loss, style=False, mse=True)
        if pts:
            with torch.no_grad():
                vals, taps = self.vgg.run(content_image.detach(), pts, keep_taps=False)
            groups = [(self.content_losses, self._spec.content)]
            if self.compute_feature_loss:
                groups.append((self.feature_losses, self._spec.feature))
            for nodes, entries in groups:
                for (node, _), entry in zip(nodes, entries):
                    pt = entry[0]
                    src, relu = vgg_loss.mse_operand(pt, vals, taps)
                    entry[1] = src                           # what the fused kernels compare against (pre-ReLU for ReLU_k)
                    # the node's own target is the VALUE of its layer (reference semantics of node.target / node(input))
                    node.set_target(self._relu_split(src) if relu else src)
        self._content_key = (content_image, content_image._version, self.compute_feature_loss)

    def forward(self, input_image: torch.Tensor, content_image=None, style_image=None) -> None:
        """network.py:366-401: evaluates every loss node; results live in ``node.loss``."""
        ops._require_cuda(input_image, 'StyleNetwork.forward')
        if content_image is not None:
            self._set_content(content_image)
        elif self._content_key is not None and self._content_key[2] != self.compute_feature_loss and self.compute_feature_loss:
            self._set_content(self._content_key[0])          # the feature targets were skipped so far
        if style_image is not None:
            # bug-compatible with network.py:390-394: the *content* image sets the style targets
            self._set_style_targets(content_image)
        losses = vgg_loss.vgg_losses(input_image, self.vgg, self._spec, self.compute_feature_loss)
        self._loss_vector = losses                       # [style nodes, content nodes, feature nodes]
        nodes = self.style_losses + self.content_losses + self.feature_losses
        for i, (node, _) in enumerate(nodes):
            node.loss = losses[i]

    def loss_weights(self, style_weight, content_weight, feature_weight=0.0):
        """Per-entry weights of the loss vector (device tensor)."""
        w = ([float(style_weight)] * len(self.style_losses) + [float(content_weight)] * len(self.content_losses) +
             [float(feature_weight)] * len(self.feature_losses))
        return torch.tensor(w, dtype=torch.float32, device=self.vgg.device)

    def gatys_closure(self, input_image, content_image, style_weight=100_000, content_weight=1, graph=None):
        """The body of train_gatys' closure (network.py:438-451): zero_grad -> forward ->
        style_weight * style + content_weight * content -> backward into ``input_image.grad``.
        Returns a callable; with CUDA graphs enabled (default) it replays a captured graph."""
        input_image.requires_grad_()

        # style_weight * sum(style losses) + content_weight * content loss as one dot product with the
        # loss vector (an elementwise product and a sum instead of the stack / sum / mul / add chain of the getters)
        weights = self.loss_weights(style_weight, content_weight)

        def body():
            input_image.grad = None
            self(input_image, content_image)
            # total = <losses, weights> as one dot product; the weights go into the backward as the gradient of
            # the loss vector itself (no mul / sum nodes for autograd to walk back through)
            losses = self._loss_vector
            total_loss = ops.loss_total(losses.detach(), weights)
            torch.autograd.backward(losses, grad_tensors=weights)
            return total_loss
        if graph is None:
            graph = cuda_graphs_enabled()
        if not graph:
            return body
        # drop the autograd graphs kept alive by earlier ``node.loss`` tensors: their AccumulateGrad
        # node for the image lives on another stream and would invalidate the capture
        for node, _ in self.style_losses + self.content_losses + self.feature_losses:
            node.loss = None
        self._loss_vector = None
        return GraphedClosure(body, leaves=[input_image])

    def get_content_optimizer(self, input_img, optt=optim.Adam):
        """network.py:403-409.  ``optim.LBFGS`` (what train_gatys asks for) is served by the
        device-resident ``optim.DeviceLBFGS`` -- same defaults and stopping rules, three launches
        per iteration and no host round trip; ``STRANSFER_DEVICE_LBFGS=0`` keeps torch's own."""
        dense = input_img.is_cuda and input_img.is_contiguous() and input_img.dtype == torch.float32
        if optt is optim.LBFGS and dense and os.environ.get('STRANSFER_DEVICE_LBFGS', '1') != '0':
            return st_optim.DeviceLBFGS([input_img.requires_grad_()])
        return optt([input_img.requires_grad_()])

    def train_gatys(self, style_image: torch.Tensor, content_image: torch.Tensor,
                    steps=550, style_weight=100_000, content_weight=1) -> torch.Tensor:
        """network.py:411-458.  Same L-BFGS loop; the per-closure ``LOGGER.info`` of the
        loss (a device sync every closure, network.py:453) is emitted every 50th closure."""
        assert isinstance(style_image, torch.Tensor), 'Images need to be already loaded'
        assert isinstance(content_image, torch.Tensor), 'Images need to be already loaded'
        # start from the content image (network.py:429); a loader's permuted HWC->CHW view becomes a dense NCHW tensor
        input_image = content_image.detach().clone(memory_format=torch.contiguous_format).float()
        optimizer = self.get_content_optimizer(input_image, optt=optim.LBFGS)
        self._content_key = None                    # targets of THIS content image, whatever was cached before
        self.compute_feature_loss = False           # never read by this loop
        n_closure = [0]
        evaluate = self.gatys_closure(input_image, content_image, style_weight, content_weight)
        for step in tqdm(range(steps)):
            def closure():
                total_loss = evaluate()
                n_closure[0] += 1
                if n_closure[0] % 50 == 0:
                    LOGGER.info('Loss: %s', total_loss)
                return total_loss
            optimizer.step(closure)
        # diagnostics for callers that time the loop (bench.py): closures evaluated, curvature pairs held at the end
        self.last_train_stats = {'closures': n_closure[0], 'steps': steps}
        if isinstance(optimizer, st_optim.DeviceLBFGS):
            self.last_train_stats['history_pairs'] = int(optimizer.info()['k'])
        self.compute_feature_loss = True
        return input_image


# --------------------------------------------------------------------------
# transform-net building blocks (reference network.py:461-611)
# --------------------------------------------------------------------------
class _ConvFn(torch.autograd.Function):
    """Stand-alone nn.Conv2d replacement (NCHW fp32 boundary) on the conv kernels."""

    @staticmethod
    def forward(ctx, x, weight, bias, stride):
        xs = ops.from_nchw(x, cpad=ops.pad8(x.shape[1]))
        pw = ops.PackedWeight(weight, 0, cols_pad=xs.C)
        _, _, y = ops.conv2d(xs, pw, bias, stride=stride, want_out=False, out_nchw_channels=weight.shape[0])
        ctx.xs, ctx.weight, ctx.stride, ctx.cin = xs, weight, stride, x.shape[1]
        return y

    @staticmethod
    def backward(ctx, gy):
        xs, w, stride = ctx.xs, ctx.weight, ctx.stride
        cout, cin, k, _ = w.shape
        g = ops.from_nchw(gy.contiguous().float(), cpad=ops.pad8(cout), planes=ops.bwd_planes())
        dw, db = ops.conv2d_wgrad(g, xs, k, stride, k // 2, 0, cout, cin)
        gx = None
        if ctx.needs_input_grad[0]:
            pw = ops.PackedWeight(w, 1, rows_pad=xs.C, cols_pad=g.C)
            _, _, gx = ops.conv2d(g, pw, stride=stride, transposed=1, out_hw=(xs.H, xs.W), want_out=False,
                                  out_nchw_channels=cin)
        return gx, dw, db, None


class Conv2d(nn.Module):
    """Parameter container + stand-alone forward with nn.Conv2d's names, shapes and default
    init (kaiming-uniform a=sqrt(5)); zero padding of k//2 (the reference's effective
    arithmetic).  Inside ImageTransformNet the fused path is used instead."""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1, padding=None, padding_mode='zeros'):
        super().__init__()
        import math
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.stride = kernel_size, stride
        self.padding = kernel_size // 2 if padding is None else padding
        assert self.padding == kernel_size // 2, 'the kernels implement "same" padding (k // 2)'
        self.padding_mode = 'zeros'          # 'reflection' ran as zeros in the reference (torch 1.1.0)
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels, kernel_size, kernel_size))
        self.bias = nn.Parameter(torch.empty(out_channels))
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        bound = 1.0 / math.sqrt(in_channels * kernel_size * kernel_size)
        nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x):
        ops._require_cuda(x, 'Conv2d.forward')
        return _ConvFn.apply(x.contiguous().float(), self.weight, self.bias, self.stride)


class _InstanceNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        xs = ops.from_nchw(x)
        ys, stats = ops.instnorm_fwd(xs, gamma, beta, eps=eps)
        ctx.xs, ctx.stats, ctx.gamma, ctx.beta, ctx.C = xs, stats, gamma, beta, x.shape[1]
        return ops.to_nchw(ys, x.shape[1])

    @staticmethod
    def backward(ctx, gy):
        g = ops.from_nchw(gy.contiguous().float(), planes=ops.bwd_planes())
        dx, dg, db = ops.instnorm_bwd(ctx.xs, g, ctx.stats, ctx.gamma, ctx.beta)
        return ops.to_nchw(dx, ctx.C), dg, db, None


class InstanceNorm2d(nn.Module):
    """nn.InstanceNorm2d(affine=True, track_running_stats=False) (network.py:474, :483)."""

    def __init__(self, num_features, eps=1e-5, affine=True):
        super().__init__()
        assert affine, 'the reference only uses affine instance norms'
        self.num_features, self.eps = num_features, eps
        self.weight = nn.Parameter(torch.ones(num_features))
        self.bias = nn.Parameter(torch.zeros(num_features))

    def forward(self, x):
        ops._require_cuda(x, 'InstanceNorm2d.forward')
        return _InstanceNormFn.apply(x.contiguous().float(), self.weight, self.bias, self.eps)


class ResidualBlock(nn.Module):
    """network.py:461-506: conv -> IN -> ReLU -> conv -> (+= x) -> IN, no trailing ReLU."""

    def __init__(self, in_channels, out_channels, kernel_size=3, stride=1):
        super().__init__()
        self.conv1 = Conv2d(in_channels, out_channels, kernel_size, stride)
        self.insn1 = InstanceNorm2d(out_channels, affine=True)
        self.relu = nn.ReLU()
        self.conv2 = Conv2d(out_channels, out_channels, kernel_size, stride)
        self.insn2 = InstanceNorm2d(out_channels, affine=True)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        residual = x
        out = self.relu(self.insn1(self.conv1(x)))
        out = self.conv2(out)
        out = out + residual
        return self.insn2(out)


# --------------------------------------------------------------------------
# total-variation / temporal losses as autograd functions (network.py:621-641, :885-903)
# --------------------------------------------------------------------------
class _TVFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, factor):
        ctx.x, ctx.factor = x, factor
        return ops.tv_fwd(x, factor)

    @staticmethod
    def backward(ctx, g):
        return ops.tv_bwd(ctx.x, ctx.factor, g.reshape(1).contiguous().float()), None


class _TemporalFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, old_content, old_stylized, cur_content, cur_stylized, tw, flow=None):
        if flow is None:
            out3 = ops.temporal_fwd(old_content, old_stylized, cur_content, cur_stylized, tw)
        else:
            out3 = ops.temporal_warp_fwd(old_content, old_stylized, cur_content, cur_stylized, flow, tw)
        ctx.os, ctx.cs, ctx.tw, ctx.out3, ctx.flow = old_stylized, cur_stylized, tw, out3, flow
        return out3[0]

    @staticmethod
    def backward(ctx, g):
        gw = g.reshape(1).contiguous().float()
        if ctx.flow is None:
            gs = ops.temporal_bwd(ctx.os, ctx.cs, ctx.tw, ctx.out3, gw)
        else:
            gs = ops.temporal_warp_bwd(ctx.os, ctx.cs, ctx.flow, ctx.tw, ctx.out3, gw)
        return None, None, None, gs, None, None


# --------------------------------------------------------------------------
# ImageTransformNet (reference network.py:509-832)
# --------------------------------------------------------------------------
class ImageTransformNet(nn.Sequential):
    """network.py:509-832.  Same 23 modules / 62 state-dict tensors as the reference; the
    forward runs the fused split-bf16 pipeline of ``transform.TransformNetFn``."""

    def __init__(self, style_image: torch.Tensor, batch_size=4):
        super().__init__(
            Conv2d(3, 32, 9, 1), InstanceNorm2d(32, affine=True), nn.ReLU(),
            Conv2d(32, 64, 3, 2), InstanceNorm2d(64, affine=True), nn.ReLU(),
            Conv2d(64, 128, 3, 2), InstanceNorm2d(128, affine=True), nn.ReLU(),
            ResidualBlock(128, 128, 3), ResidualBlock(128, 128, 3), ResidualBlock(128, 128, 3),
            ResidualBlock(128, 128, 3), ResidualBlock(128, 128, 3),
            nn.Upsample(mode='nearest', scale_factor=2),
            Conv2d(128, 64, 3, 1), InstanceNorm2d(64, affine=True), nn.ReLU(),
            nn.Upsample(mode='nearest', scale_factor=2),
            Conv2d(64, 32, 3, 1), InstanceNorm2d(32, affine=True), nn.ReLU(),
            Conv2d(32, 3, 9, 1),
        )
        assert isinstance(style_image, torch.Tensor), 'Style image need to be already loaded'
        self.style_image = style_image
        self.batch_size = batch_size
        self._pack_cache = transform._PackCache()

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        ops._require_cuda(input, type(self).__name__ + '.forward')
        sd = dict(self.named_parameters())
        params = [sd[n] for n in transform.PARAM_NAMES]
        if self._pack_cache.stale():               # packed copies owned by the optimizer, weights changed elsewhere
            self._pack_cache.repack(sd)
        x = input.contiguous().float()
        if not torch.is_grad_enabled():
            return self._infer(x, params)
        return transform.TransformNetFn.apply(x, None, self._pack_cache, *params)

    INFER_PIXELS_PER_CHUNK = 1 << 23      # 4 frames of 1080x1920 (32-channel tensors stay below 2^31 elements)

    def _infer(self, x, params, x2=None):
        """Forward without a graph (process_image / process_video / static_test,
        network.py:798-832, :1117-1134, :772-796): the batch goes through the fused pipeline in
        chunks of whole frames, every intermediate is released once consumed."""
        N, _, H, W = x.shape
        per = max(1, self.INFER_PIXELS_PER_CHUNK // (H * W))
        if N <= per:
            return transform.run_forward(x, self._pack_cache, params, keep=False, x2=x2)[0]
        out = torch.empty(N, 3, H, W, dtype=torch.float32, device=x.device)
        for i in range(0, N, per):
            out[i:i + per] = transform.run_forward(x[i:i + per], self._pack_cache, params, keep=False,
                                                   x2=None if x2 is None else x2[i:i + per])[0]
        return out

    def stylize_host_batch(self, frames: torch.Tensor, out: torch.Tensor = None) -> torch.Tensor:
        """Streaming inference for convert-image / convert-video style workloads (network.py:798-832;
        SURVEY section 8f rank 3): ``frames`` [N, C, H, W] fp32 in (pinned) HOST memory -> stylised frames in
        (pinned) host memory.  Chunks of whole frames are double-buffered: a copy stream uploads chunk
        i+1 while the kernels run chunk i and a second copy stream downloads the result of chunk i-1."""
        if frames.device.type != 'cpu' or frames.dtype != torch.float32 or frames.dim() != 4:
            raise StransferError('stylize_host_batch: an fp32 [N, C, H, W] host tensor is required')
        dev = next(self.parameters()).device
        if dev.type != 'cuda':
            raise StransferError('stylize_host_batch: the network is not on a CUDA device (no CPU fallback)')
        N, C, H, W = frames.shape
        per = min(N, max(1, self.INFER_PIXELS_PER_CHUNK // (H * W)))
        if out is None:
            out = torch.empty((N, 3, H, W), dtype=torch.float32).pin_memory()
        sd = dict(self.named_parameters())
        params = [sd[n] for n in transform.PARAM_NAMES]
        cur = torch.cuda.current_stream(dev)
        up, down = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        stage_in = [torch.empty((per, C, H, W), dtype=torch.float32, device=dev) for _ in range(2)]
        stage_out = [torch.empty((per, 3, H, W), dtype=torch.float32, device=dev) for _ in range(2)]
        uploaded, consumed, produced, drained = ([torch.cuda.Event() for _ in range(2)] for _ in range(4))
        for ev in consumed + drained:
            ev.record(cur)
        starts = list(range(0, N, per))

        def upload(ci):
            slot, i = ci & 1, starts[ci]
            n = min(per, N - i)
            with torch.cuda.stream(up):
                up.wait_event(consumed[slot])
                stage_in[slot][:n].copy_(frames[i:i + n], non_blocking=True)
                uploaded[slot].record(up)
        upload(0)
        with torch.no_grad():
            for ci, i in enumerate(starts):
                slot, n = ci & 1, min(per, N - i)
                if ci + 1 < len(starts):
                    upload(ci + 1)
                cur.wait_event(uploaded[slot])
                cur.wait_event(drained[slot])
                stage_out[slot][:n].copy_(transform.run_forward(stage_in[slot][:n], self._pack_cache, params, keep=False)[0])
                consumed[slot].record(cur)
                produced[slot].record(cur)
                with torch.cuda.stream(down):
                    down.wait_event(produced[slot])
                    out[i:i + n].copy_(stage_out[slot][:n], non_blocking=True)
                    drained[slot].record(down)
        cur.wait_stream(down)
        cur.wait_stream(up)
        return out

    def get_total_variation_regularization_loss(self, transformed_image: torch.Tensor,
                                                regularization_factor=1e-6) -> torch.Tensor:
        """network.py:621-641"""
        ops._require_cuda(transformed_image, 'get_total_variation_regularization_loss')
        return _TVFn.apply(transformed_image.contiguous().float(), float(regularization_factor))

    def get_optimizer(self, optimizer=optim.Adam):
        """network.py:643-649.  ``optim.Adam`` is served by ``optim.FusedAdam`` (same update rule,
        one launch over the flat parameter / gradient / moment buffers; its ``.bucket`` is the flat
        gradient buffer the training closures write); ``STRANSFER_FUSED_ADAM=0`` keeps torch's own."""
        params = list(self.parameters())
        if optimizer is optim.Adam and params[0].is_cuda and os.environ.get('STRANSFER_FUSED_ADAM', '1') != '0':
            return st_optim.FusedAdam(params)
        return optimizer(params)

    # -- the closure body of static_train (network.py:690-731), factored out so that the
    #    benchmark and the parity tests run exactly what the training loop runs
    def closure_loss(self, batch, loss_network, style_weight=100_000, content_weight=1, tv_scale=1.0):
        """``tv_scale``: world size under data parallelism (dist.tv_scale()): the TV term is a
        sum over the batch while the other terms are means, see dist.py."""
        transformed_image = self(batch)
        loss_network(transformed_image, content_image=batch)
        style_loss = loss_network.get_total_current_style_loss(weight=style_weight)
        content_loss = loss_network.get_total_current_content_loss(weight=content_weight)
        regularization_loss = self.get_total_variation_regularization_loss(transformed_image)
        if tv_scale != 1.0:
            regularization_loss = regularization_loss * tv_scale
        return style_loss + content_loss + regularization_loss, transformed_image

    def _backward_into_bucket(self, total_loss, grad_tensors=None):
        """backward() with the parameters' (zeroed, bucket-view) ``.grad`` tensors as gradient sinks: the
        backward kernels write dW / db / dgamma / dbeta straight into the flat bucket instead of handing 62
        tensors to autograd's accumulate kernels (transform._PackCache.grad_sink).  ``total_loss`` may be a list of
        tensors with their ``grad_tensors`` (the loss vector and the extra terms of ``_weighted_total``)."""
        if os.environ.get('STRANSFER_GRAD_SINKS', '1') != '0':
            self._pack_cache.grad_sink = {n: p.grad for n, p in self.named_parameters() if p.grad is not None}
        try:
            if grad_tensors is None:
                total_loss.backward()
            else:
                torch.autograd.backward(total_loss, grad_tensors=grad_tensors)
        finally:
            self._pack_cache.grad_sink = None

    def _weighted_total(self, loss_network, weights, extras):
        """The training closures' total loss without the getters' stack / sum / mul / add chain (network.py:704-722):
        ``total = <loss vector, weights> + sum(extra * scale)`` in one launch, and the backward seeded with the weights
        themselves -- the same arithmetic as style_weight * style + content_weight * content + TV (+ temporal), no ATen
        launches between the loss kernels and the backward.  extras: [(0-dim loss tensor, scale, device scalar of that
        scale)].  Returns (total, tensors to differentiate, their gradient seeds)."""
        losses = loss_network._loss_vector
        total = ops.loss_total(losses.detach(), weights, [(t.detach(), sc) for t, sc, _ in extras])
        return total, [losses] + [t for t, _, _ in extras], [weights] + [g for _, _, g in extras]


    def _step_closure(self, body, bucket, optimizer, graph, loss_network=None):
        """Turns the body of a training closure (zero the bucket -> forward -> backward into the bucket) into what the
        loops hand to ``optimizer.step``.

        eager:  body -> all-reduce.
        graph, no fused optimizer: [graph: weight packing + body] -> all-reduce (average) ; the optimizer follows.
        graph + ``optim.FusedAdam``: ONE graph replay per step holds body -> all-reduce (sum, NCCL is capturable) ->
        Adam kernel; the 1/world of the average is the kernel's ``grad_scale`` and the kernel rewrites the packed
        weight copies itself (no packing launches, no separate scaling pass: SURVEY section 8f rank 2).  The returned
        callable carries ``includes_optimizer`` so that ``optimizer.step(closure)`` replays it without a second
        update.  ``STRANSFER_GRAPH_STEP=0`` keeps the optimizer outside the graph."""
        world = st_dist.world_size()
        fused = isinstance(optimizer, st_optim.FusedAdam) and optimizer.bucket is bucket
        if fused:
            optimizer.grad_scale = 1.0 / world

        def reduce():
            bucket.all_reduce(average=not fused)
        if not graph:
            def eager():
                total_loss = body()
                reduce()
                return total_loss
            return eager
        cache = self._pack_cache
        whole_step = fused and os.environ.get('STRANSFER_GRAPH_STEP', '1') != '0'
        if fused:
            # one eager evaluation creates every packed weight copy the closure reads; from here on the optimizer's
            # kernel keeps them up to date and the captured graph contains no packing kernels
            cache.frozen = False
            cache.entries.clear()
            body()
            if loss_network is not None:
                # drop the autograd graph that evaluation left behind (node.loss): its AccumulateGrad nodes live on
                # this stream and would invalidate the capture on the side stream
                for node, _ in loss_network.style_losses + loss_network.content_losses + loss_network.feature_losses:
                    node.loss = None
                loss_network._loss_vector = None
            optimizer.set_packs(cache, dict(self.named_parameters()))
            optimizer._refresh_active(optimizer.param_groups[0]['params'])
        else:
            cache.always_repack = True          # the packing kernels are part of every replay
        try:
            if whole_step:
                def step_body():
                    total_loss = body()
                    reduce()
                    optimizer.launch()
                    return total_loss
                # capturing must not move the weights: the warm-up runs execute the Adam kernel, so the optimizer
                # state is saved and restored around the capture (and the packed copies rewritten from the weights)
                state = [t.clone() for t in (optimizer.flat_param, optimizer.exp_avg, optimizer.exp_avg_sq, optimizer.steps)]
                graphed = GraphedClosure(step_body, leaves=[p for p in self.parameters() if p.requires_grad], warmup=2)
                for dst, src in zip((optimizer.flat_param, optimizer.exp_avg, optimizer.exp_avg_sq, optimizer.steps), state):
                    dst.copy_(src)
                cache.repack(dict(self.named_parameters()))
            else:
                graphed = GraphedClosure(body, leaves=[p for p in self.parameters() if p.requires_grad])
        finally:
            cache.always_repack = False
            if not fused:
                cache.entries.clear()           # packed copies made during capture belong to the graph
        named = dict(self.named_parameters())

        def replay():
            if cache.stale():                   # weights changed behind the optimizer's back (load_state_dict ...)
                cache.repack(named)
            total_loss = graphed()
            if not whole_step:
                reduce()
            return total_loss
        replay.launches = graphed.launches
        if whole_step:
            replay.includes_optimizer = optimizer
        return replay

    def training_closure(self, batch, loss_network, bucket, style_weight=100_000, content_weight=1, graph=None, optimizer=None):
        """The closure ``static_train`` hands to ``optimizer.step`` (network.py:690-731) for a batch
        held in the fixed tensor ``batch`` (the loop copies every new batch into it): zero the
        gradient bucket -> transform net -> loss network -> backward (-> all-reduce).  With CUDA
        graphs enabled (default) everything up to the all-reduce is captured once and replayed, so a
        step costs one graph launch plus the collective and the optimizer instead of ~290 launches
        from Python.  Everything the graph reads lives at fixed addresses: the batch, the fp32 master
        weights (updated in place by the optimizer; their packed copies are rebuilt inside the graph)
        and the flat gradient bucket."""
        world = st_dist.world_size()

        weights = loss_network.loss_weights(style_weight, content_weight)
        tv_seed = torch.full((), float(world), dtype=torch.float32, device=batch.device)

        def body():
            loss_network._content_key = None             # the content image is the batch: new targets every step
            bucket.rebind()
            bucket.zero()
            if os.environ.get('STRANSFER_FAST_TOTAL', '1') == '0':
                total_loss, _ = self.closure_loss(batch, loss_network, style_weight, content_weight, tv_scale=float(world))
                self._backward_into_bucket(total_loss)
                return total_loss
            # closure_loss without the getters: same terms, same weights, one launch for the total
            transformed_image = self(batch)
            loss_network(transformed_image, content_image=batch)
            tv = self.get_total_variation_regularization_loss(transformed_image)
            total_loss, terms, seeds = self._weighted_total(loss_network, weights, [(tv, float(world), tv_seed)])
            self._backward_into_bucket(terms, seeds)
            return total_loss
        if graph is None:
            graph = cuda_graphs_enabled()
        if graph:
            for node, _ in loss_network.style_losses + loss_network.content_losses + loss_network.feature_losses:
                node.loss = None
            loss_network._loss_vector = None
        return self._step_closure(body, bucket, optimizer, graph, loss_network)

    def static_train(self, style_name='nsp', epochs=50, style_weight=100_000, content_weight=1,
                     train_loader=None, test_loader=None, max_iterations=None):
        """network.py:651-770.  ``train_loader`` / ``test_loader`` (extra, optional) replace the
        COCO loaders -- the dataset download is host glue outside the hot path."""
        from . import dataset
        # rank 0 alone owns the log directory, the test-loss log and the checkpoints (every rank holds the same weights)
        rank0 = st_dist.rank() == 0
        tb_writer = get_tensorboard_writer(f'runs/fast-image-style-transfer-still-image_{style_name}' if rank0 else None)
        loss_network = StyleNetwork(self.style_image, torch.rand([1, 3, 256, 256]).to(constants.DEVICE))
        loss_network.compute_feature_loss = False        # the training closure never reads it
        optimizer = self.get_optimizer(optimizer=optim.Adam)
        LOGGER.info('Training network with "%s" optimizer', type(optimizer))
        # data parallel (launched under torchrun): loader batches are GLOBAL batches, sharded
        # over the ranks; one flat-bucket all-reduce per closure (dist.py)
        world = st_dist.world_size()
        # one flat gradient buffer (also the all-reduce bucket); FusedAdam reads it in place
        bucket = getattr(optimizer, 'bucket', None) or st_dist.GradBucket(self.parameters())
        train_buf, closure = None, None
        iteration = 0
        if train_loader is None:
            test_loader, train_loader = dataset.get_coco_loader(test_split=0.10, test_limit=20,
                                                                batch_size=self.batch_size)
        for epoch in range(epochs):
            LOGGER.info('Starting epoch %d', epoch)
            epoch_checkpoint_name = os.path.join(constants.PROJECT_ROOT_PATH,
                                                 f'data/models/fast_st_{style_name}_epoch{epoch}.pth')
            if os.path.isfile(epoch_checkpoint_name):
                self.load_state_dict(adaptive_torch_load(epoch_checkpoint_name))
                continue
            for batch in tqdm(train_loader):
                batch = batch.squeeze(1).to(constants.DEVICE)
                if world > 1:
                    batch = st_dist.shard(batch)
                # the closure (a replayed CUDA graph by default) reads the batch from a fixed buffer
                if train_buf is None or train_buf.shape != batch.shape:
                    train_buf = batch.detach().clone().contiguous().float()
                    closure = self.training_closure(train_buf, loss_network, bucket, style_weight, content_weight,
                                                    optimizer=optimizer)
                else:
                    train_buf.copy_(batch)

                if iteration % 150 == 0 and test_loader is not None and rank0:
                    average_test_loss = self.static_test(test_loader, loss_network)
                    tb_writer.add_scalar('data/fst_test_loss', average_test_loss, iteration)
                # the reference evaluates an extra closure every 20th iteration to log its loss (network.py:733-741);
                # that value is the loss optimizer.step(closure) returns for the same weights and batch
                total_loss = optimizer.step(closure)
                if iteration % 20 == 0:
                    tb_writer.add_scalar('data/fst_train_loss', total_loss, iteration)
                    LOGGER.info('Batch Loss: %.8f', total_loss)
                iteration += 1
                if max_iterations is not None and iteration >= max_iterations:
                    return
            if rank0:
                os.makedirs(os.path.dirname(epoch_checkpoint_name), exist_ok=True)
                torch.save(self.state_dict(), epoch_checkpoint_name)
            st_dist.barrier()

    def static_test(self, test_loader, loss_network, style_weight=100_000, feature_weight=1):
        """network.py:772-796"""
        total_test_loss = []
        prev = loss_network.compute_feature_loss
        loss_network.compute_feature_loss = True
        with torch.no_grad():
            for test_batch in test_loader:
                test_batch = test_batch.squeeze(1).to(constants.DEVICE)
                loss_network._content_key = None     # every test batch sets its own targets (network.py:786)
                transformed_image = torch.clamp(self(test_batch), min=0, max=255)
                loss_network(transformed_image, content_image=test_batch)
                style_loss = style_weight * loss_network.get_total_current_style_loss()
                feature_loss = feature_weight * loss_network.get_total_current_feature_loss()
                total_test_loss.append((style_loss + feature_loss).item())
        loss_network.compute_feature_loss = prev
        average_test_loss = torch.mean(torch.Tensor(total_test_loss))
        LOGGER.info('Average test loss: %.8f', average_test_loss)
        return average_test_loss

    def process_image(self, image_path: str, style_name='nsp', out_dir='results/') -> None:
        """network.py:798-832"""
        from . import img_utils
        self.load_state_dict(_load_latest_model_weigths(model_name='fast_st', style_name=style_name))
        input_image = img_utils.image_loader(os.path.join(constants.PROJECT_ROOT_PATH, image_path))
        with torch.no_grad():
            transformed_image = self(input_image)
        out_dir = os.path.join(constants.PROJECT_ROOT_PATH, out_dir)
        os.makedirs(out_dir, exist_ok=True)
        out_file = os.path.join(out_dir, f'converted_fast_st_{style_name}.png')
        img_utils.imshow(transformed_image, path=out_file)


# --------------------------------------------------------------------------
# VideoTransformNet (reference network.py:835-1158)
# --------------------------------------------------------------------------
class VideoTransformNet(ImageTransformNet):
    """network.py:835-1158: 6-channel first conv (frame_t ++ stylised_{t-1}), optional warm
    start from a fast_st state dict, temporal loss."""

    def __init__(self, style_image: torch.Tensor, batch_size=4, fast_transfer_dict=None):
        super().__init__(style_image, batch_size)
        self[0] = Conv2d(6, 32, 9, 1)
        if fast_transfer_dict is not None:
            if isinstance(fast_transfer_dict, str):
                fast_transfer_dict = adaptive_torch_load(fast_transfer_dict)
            del fast_transfer_dict['0.weight']
            del fast_transfer_dict['0.bias']
            m_sd = self.state_dict().copy()
            m_sd.update(fast_transfer_dict)
            self.load_state_dict(m_sd)
            self.has_external_weights = True
        else:
            self.has_external_weights = False

    def forward_pair(self, frame: torch.Tensor, previous: torch.Tensor) -> torch.Tensor:
        """``self(torch.cat([frame, previous], dim=1))`` (network.py:984-986, :1123-1125) without the concatenation:
        the first layer's loader reads its channels 0..2 from ``frame`` and 3..5 from ``previous``."""
        ops._require_cuda(frame, 'VideoTransformNet.forward_pair')
        sd = dict(self.named_parameters())
        params = [sd[n] for n in transform.PARAM_NAMES]
        a, b = frame.contiguous().float(), previous.contiguous().float()
        if tuple(a.shape) != tuple(b.shape) or a.shape[1] != 3:
            return self(torch.cat([a, b], dim=1))
        if not torch.is_grad_enabled():
            return self._infer(a, params, x2=b)
        return transform.TransformNetFn.apply(a, b, self._pack_cache, *params)

    def get_temporal_loss(self, old_content, old_stylized, current_content, current_stylized,
                          temporal_weight=1, flow=None) -> torch.Tensor:
        """network.py:885-903: tw * ||cur_sty - old_sty|| / (||cur_con - old_con|| + 1); the
        gradient flows into ``current_stylized`` only.  ``flow`` (extension, no reference code: SURVEY F9)
        [B, 2, H, W] in pixels warps the two previous frames bilinearly before the comparison."""
        ops._require_cuda(current_stylized, 'get_temporal_loss')
        c = lambda t: t.detach().contiguous().float()
        if flow is not None:
            assert tuple(flow.shape) == (current_stylized.shape[0], 2) + tuple(current_stylized.shape[2:]), 'flow must be [B, 2, H, W]'
            flow = c(flow)
        return _TemporalFn.apply(c(old_content), c(old_stylized), c(current_content),
                                 current_stylized.contiguous().float(), float(temporal_weight), flow)

    def closure_loss_video(self, batch, old_images, loss_network, temporal_weight=0.8,
                           style_weight=100_000, content_weight=1, tv_scale=1.0, flow=None):
        """Closure body of video_train (network.py:984-1037); updates ``old_images`` in place."""
        old_content_images, old_styled_images = old_images
        transformed_image = self.forward_pair(batch, old_styled_images)      # = self(cat([batch, old_styled], 1))
        loss_network(transformed_image, content_image=batch)
        style_loss = loss_network.get_total_current_style_loss(weight=style_weight)
        content_loss = loss_network.get_total_current_content_loss(weight=content_weight)
        regularization_loss = self.get_total_variation_regularization_loss(transformed_image)
        if tv_scale != 1.0:
            regularization_loss = regularization_loss * tv_scale
        temporal_loss = self.get_temporal_loss(old_content_images, old_styled_images, batch, transformed_image,
                                               temporal_weight=temporal_weight, flow=flow)
        total_loss = style_loss + content_loss + regularization_loss + temporal_loss
        old_images[0] = batch.detach()
        old_images[1] = transformed_image.detach()
        return total_loss, transformed_image

    def video_training_closure(self, batch, old_content, old_stylized, next_stylized, loss_network, bucket,
                               temporal_weight=0.8, style_weight=100_000, content_weight=1, graph=None, flow=None,
                               optimizer=None):
        """The closure ``video_train`` hands to ``optimizer.step`` (network.py:988-1037) on FIXED tensors: the
        frame batch, the previous frame's content / stylised images as they were before this iteration
        (network.py:980-986 binds them outside the closure) and ``next_stylized``, which receives this frame's
        stylised output (the loop makes it the next ``old_stylized`` after the optimizer step).  zero the
        bucket -> 6-channel forward -> loss network -> TV + temporal loss -> backward (-> all-reduce).  With
        CUDA graphs everything up to the all-reduce is one replayed graph: at one clip per GPU the ~260
        eager launches cost more host time than device time.  ``flow`` (extension): a FIXED [B, 2, H, W]
        tensor of optical flow in pixels from the previous frame to this one; the temporal loss then compares against
        the flow-warped previous frames (the loop copies every frame's flow into it)."""
        world = st_dist.world_size()
        weights = loss_network.loss_weights(style_weight, content_weight)
        tv_seed = torch.full((), float(world), dtype=torch.float32, device=batch.device)
        one_seed = torch.ones((), dtype=torch.float32, device=batch.device)

        def body():
            loss_network._content_key = None
            bucket.rebind()
            bucket.zero()
            if os.environ.get('STRANSFER_FAST_TOTAL', '1') == '0':
                state = [old_content, old_stylized]
                total_loss, transformed = self.closure_loss_video(batch, state, loss_network, temporal_weight, style_weight,
                                                                  content_weight, tv_scale=float(world), flow=flow)
                self._backward_into_bucket(total_loss)
            else:
                # closure_loss_video without the getters (see _weighted_total)
                transformed = self.forward_pair(batch, old_stylized)
                loss_network(transformed, content_image=batch)
                tv = self.get_total_variation_regularization_loss(transformed)
                temporal = self.get_temporal_loss(old_content, old_stylized, batch, transformed,
                                                  temporal_weight=temporal_weight, flow=flow)
                total_loss, terms, seeds = self._weighted_total(loss_network, weights,
                                                                [(tv, float(world), tv_seed), (temporal, 1.0, one_seed)])
                self._backward_into_bucket(terms, seeds)
            with torch.no_grad():
                next_stylized.copy_(transformed)
            return total_loss
        if graph is None:
            graph = cuda_graphs_enabled()
        if graph:
            for node, _ in loss_network.style_losses + loss_network.content_losses + loss_network.feature_losses:
                node.loss = None
            loss_network._loss_vector = None
        return self._step_closure(body, bucket, optimizer, graph, loss_network)

    def video_train(self, style_name='nsp', epochs=50, temporal_weight=0.8, style_weight=100_000,
                    feature_weight=1, content_weight=1, video_loader=None, max_iterations=None) -> None:
        """network.py:905-1069.  ``video_loader`` (extra, optional): an iterable of clips, each an
        iterable of frame batches [B,3,H,W] (replaces VideoDataset + iterate_on_video_batches) -- or of
        ``(frames [B,3,H,W], flow [B,2,H,W])`` pairs: the optical flow (in pixels) from the previous frame, which
        switches the temporal loss to its flow-warped form (extension: the reference's loss has no flow)."""
        from . import dataset
        rank0 = st_dist.rank() == 0
        tb_writer = get_tensorboard_writer(f'runs/video-style-transfer_{style_name}' if rank0 else None)
        style_loss_network = StyleNetwork(self.style_image, torch.rand([1, 3, 256, 256]).to(constants.DEVICE))
        style_loss_network.compute_feature_loss = False
        optimizer = self.get_optimizer(optimizer=optim.Adam)
        LOGGER.info('Training video network with "%s" optimizer', type(optimizer))
        # data parallel over CLIPS (frames of a clip are sequentially dependent);
        # the temporal loss uses the norms of the rank's own clips (per-shard norms)
        world = st_dist.world_size()
        bucket = getattr(optimizer, 'bucket', None) or st_dist.GradBucket(self.parameters())
        iteration = 0
        if video_loader is None:
            video_loader = dataset.VideoDataset(batch_size=self.batch_size)
        bufs, closure, closure_key = None, None, None
        for epoch in range(epochs):
            if epoch == 0 and self.has_external_weights:
                LOGGER.info('Freezing weights imported from fast transfer network for the first epoch')
                for name, param in self.named_parameters():
                    if not name.startswith('0.'):
                        param.requires_grad = False
            if epoch == 1 and self.has_external_weights:
                LOGGER.info('Unfreezing all weights')
                for param in self.parameters():
                    param.requires_grad = True
            epoch_checkpoint_name = os.path.join(constants.PROJECT_ROOT_PATH,
                                                 f'data/models/video_st_{style_name}_epoch{epoch}.pth')
            if os.path.isfile(epoch_checkpoint_name):
                self.load_state_dict(adaptive_torch_load(epoch_checkpoint_name))
                continue
            LOGGER.info('Starting epoch %d', epoch)
            for video_batch in video_loader:
                first_frame = True
                for batch in dataset.iterate_on_video_batches(video_batch):
                    flow = None
                    if isinstance(batch, (tuple, list)):
                        batch, flow = batch
                        flow = flow.to(constants.DEVICE)
                    batch = batch.to(constants.DEVICE)
                    if world > 1:
                        batch = st_dist.shard(batch)
                        flow = None if flow is None else st_dist.shard(flow)
                    batch = batch.contiguous().float()
                    # fixed buffers the closure (a replayed CUDA graph by default) reads: this frame, the previous
                    # frame's content / stylised images as they were BEFORE this iteration (network.py:980-986 binds
                    # them outside the closure), and the slot that receives this frame's stylised output
                    key = (tuple(batch.shape), tuple(p.requires_grad for p in self.parameters()), flow is not None)
                    if key != closure_key:
                        bufs = [torch.empty_like(batch) for _ in range(4)]      # batch, old content, old stylised, next stylised
                        flow_buf = None if flow is None else torch.zeros_like(flow, dtype=torch.float32).contiguous()
                        closure = self.video_training_closure(bufs[0], bufs[1], bufs[2], bufs[3], style_loss_network, bucket,
                                                              temporal_weight, style_weight, content_weight, flow=flow_buf,
                                                              optimizer=optimizer)
                        closure_key, first_frame = key, True
                    bufs[0].copy_(batch)
                    if flow is not None:
                        flow_buf.copy_(flow)
                    if first_frame:                      # new clip: the previous frame is the frame itself (network.py:977-978)
                        bufs[1].copy_(batch)
                        bufs[2].copy_(batch)
                        first_frame = False

                    total_loss = optimizer.step(closure)
                    if iteration % 20 == 0:          # network.py:1042-1048 logs the loss of an extra closure: the same value
                        tb_writer.add_scalar('data/fst_train_loss', total_loss, iteration)
                        LOGGER.info('Epoch: %d\tBatch Loss: %.4f', epoch, total_loss)
                    iteration += 1
                    bufs[1].copy_(bufs[0])               # old_images = [batch, transformed_image] (network.py:1016-1017)
                    bufs[2].copy_(bufs[3])
                    if max_iterations is not None and iteration >= max_iterations:
                        return
            if rank0:
                os.makedirs(os.path.dirname(epoch_checkpoint_name), exist_ok=True)
                torch.save(self.state_dict(), epoch_checkpoint_name)
            st_dist.barrier()

    def process_video(self, video_path: str, style_name='nsp', working_dir='workdir/', out_dir='results/',
                      fps=24.0, frames=None):
        """network.py:1071-1158.  The recurrent per-frame forward is the hot path; decoding the
        input video and muxing the mp4 need imageio (not in this image): ``frames`` (extra,
        optional) supplies an iterable of [1,3,H,W] tensors and the stylised frames are written
        as PNGs into ``working_dir``."""
        from . import dataset, img_utils
        working_dir = os.path.join(constants.PROJECT_ROOT_PATH, working_dir)
        self.load_state_dict(_load_latest_model_weigths(model_name='video_st', style_name=style_name))
        if frames is None:
            frames = dataset.iterate_on_video_batches([dataset.open_video(os.path.join(constants.PROJECT_ROOT_PATH, video_path))])
        shutil.rmtree(working_dir, ignore_errors=True)
        os.makedirs(working_dir, exist_ok=True)
        old_image = None
        LOGGER.info('Starting to process video into stylized frames')
        outputs = []
        # the frame loop never waits for a picture to reach the disk (the reference's per-frame imshow is a
        # device sync + host copy + PNG encode): conversion to uint8 on the device, asynchronous download into
        # pinned buffers, PNG encoding on worker threads (img_utils.FrameWriter); the 6-channel input is read from
        # its two sources by the first layer's loader (no torch.cat)
        writer = img_utils.FrameWriter()
        try:
            with torch.no_grad():
                for i, video_frame in enumerate(frames):
                    video_frame = video_frame.to(constants.DEVICE, non_blocking=True)
                    if old_image is None:
                        old_image = video_frame
                    transformed_image = self.forward_pair(video_frame, old_image)
                    old_image = transformed_image.detach()
                    outputs.append(transformed_image)
                    writer.submit(transformed_image[:1], f'{working_dir}{i}.png')
                    if i % 50 == 0:
                        LOGGER.info('.. processing, currently frame %d', i)
        finally:
            writer.close()
        LOGGER.info('All frames have been stylized.')
        dataset.write_video(working_dir, os.path.join(constants.PROJECT_ROOT_PATH, out_dir), f'video_st_{style_name}.mp4', fps)
        return outputs
