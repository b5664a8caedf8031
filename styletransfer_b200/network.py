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
