"""Single-pass VGG-19 loss network on the B200 kernels.

Replaces the seven (+2) prefix passes that StyleNetwork.forward runs per call
(reference network.py:366-401 -> run_through_pieces :316-340) by ONE forward that
emits the five pre-ReLU taps, and autograd's backward through the replicated
prefixes by ONE fused backward-data chain to the image.  VGG weights are frozen
(no wgrad).  Losses come back as a length-7 tensor
[style x5, content, feature] whose gradient weights (grad_output) are read on the
device, so no host synchronisation happens anywhere in forward or backward.
"""
import os

import torch

from . import ops

# (cin, cout, pool_before): conv1_1, conv1_2, conv2_1, conv2_2, conv3_1 = Conv2d_1..5
PREFIX = [(3, 64, False), (64, 64, False), (64, 128, True), (128, 128, False), (128, 256, True)]
CONTENT_TAP = 3
N_LOSSES = 7
# backward-data of conv1_1: tcgen05 16-channel tile (default) or the exact-fp32 CUDA-core kernel
IMAGE_GRAD_TC = True


# STRANSFER_OVERLAP_LOSSES=1 runs the loss branches of the taps on a side stream.  Measured on B200 (Gatys 512x512:
# 1077 vs 1060-1079 closures/s, fast_st 1426 vs 1431 img/s): within run-to-run noise, because the Gram kernels are
# persistent and occupy every SM anyway -- so the default keeps everything on one stream.
OVERLAP_LOSS_BRANCHES = os.environ.get('STRANSFER_OVERLAP_LOSSES', '0') != '0'
_SIDE_STREAMS = {}


def _side_stream(device):
    key = (device.type, device.index)
    if key not in _SIDE_STREAMS:
        _SIDE_STREAMS[key] = torch.cuda.Stream(device)
    return _SIDE_STREAMS[key]


class VGGPrefix:
    """Packed weights of the five executed convs."""

    def __init__(self, weights):
        assert len(weights) == 5
        self.device = weights[0][0].device
        self.fwd, self.bwd, self.bias = [], [], []
        for i, (w, b) in enumerate(weights):
            ops._require_cuda(w, 'VGGPrefix')
            if i == 0:
                # conv1_1 forward (K = 27) runs in exact fp32 on the CUDA cores straight from the NCHW
                # fp32 image (conv_image.cu): the fp32 weight is the operand
                self.w0 = w.detach().contiguous().float()
                self.fwd.append(None)
                # backward-data to the image: tcgen05 kernel with a 16-channel tile (3 used) that writes
                # the NCHW fp32 gradient from its epilogue
                self.bwd.append(ops.PackedWeight(w, 2, rows_pad=16))
            else:
                self.fwd.append(ops.PackedWeight(w, 0, cols_pad=w.shape[1]))
                # backward-data runs as a forward conv with flipped taps on the tcgen05 kernel
                self.bwd.append(ops.PackedWeight(w, 2, rows_pad=w.shape[1]))
            self.bias.append(b.detach().contiguous().float())

    def taps(self, x, upto=5, keep_all=True, on_tap=None, pool_idx=None):
        """x: NCHW fp32.  Returns the list of pre-ReLU taps (Split) of Conv2d_1..upto.  keep_all=False:
        only the last tap is wanted (content targets); layers whose ReLU copy feeds the next conv then
        skip the store of their pre-ReLU tap (None in the returned list).  ``on_tap(i, tap)`` is called as
        soon as tap i has been launched (the loss branch of the tap can then run beside the next layer).
        ``pool_idx`` (dict): filled with {tap index: routing byte map of the relu+maxpool that follows it}, which
        the backward uses instead of re-reading the full-resolution tap."""
        h = None
        taps = []
        for i in range(upto):
            pool_next = i + 1 < 5 and PREFIX[i + 1][2]
            last = i + 1 == upto
            # a tap feeds either relu+maxpool (separate kernel) or a conv that wants relu(tap)
            want_relu = not last and not pool_next
            want_out = keep_all or not want_relu
            if i == 0:
                tap, relu = ops.conv_image_fwd(x.contiguous().float(), self.w0, self.bias[0], want_out=want_out, want_relu=want_relu)
            else:
                tap, relu, _ = ops.conv2d(h, self.fwd[i], self.bias[i], want_out=want_out, want_relu=want_relu)
            taps.append(tap)
            if on_tap is not None:
                on_tap(i, tap)
            if last:
                break
            if pool_next and pool_idx is not None:
                h, pool_idx[i] = ops.relu_maxpool_fwd(tap, want_idx=True)
            else:
                h = ops.relu_maxpool_fwd(tap) if pool_next else relu
        return taps


class _VGGLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, net, style_targets, content_target, with_feature):
        losses = torch.zeros(N_LOSSES, dtype=torch.float32, device=x.device)
        diffs = [None] * 5
        feat = [None]
        # The loss branch of tap i (Gram -> split-K finish -> style MSE, and the content / feature MSE at
        # tap 4) does not feed the next layer: it runs on a side stream beside conv i+1, so its small
        # finish / reduction kernels fill SMs the convolution leaves idle (inside a captured CUDA graph the
        # two streams become parallel branches).  The streams join before forward returns.
        cur = torch.cuda.current_stream(x.device)
        side = _side_stream(x.device) if OVERLAP_LOSS_BRANCHES else None

        def branch(i, tap):
            G = ops.gram(tap)
            diffs[i] = ops.style_mse_fwd(G, style_targets[i], losses[i:i + 1])      # (diff, max |diff|)
            if i == CONTENT_TAP:
                ops.mse_fwd(tap, content_target, losses[5:6])
                if with_feature:
                    feat[0] = torch.empty(1, dtype=torch.float32, device=x.device)
                    ops.mse_fwd(tap, content_target, feat[0], relu=True)
                    ops.mse_fwd(tap, content_target, losses[6:7], relu=True, feature_form=True)

        def on_tap(i, tap):
            if side is None:
                return branch(i, tap)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                branch(i, tap)
        pool_idx = {} if ctx.needs_input_grad[0] else None       # routing maps of the two max-pools, for the backward
        taps = net.taps(x, on_tap=on_tap, pool_idx=pool_idx)
        if side is not None:
            cur.wait_stream(side)
        feat_mse = feat[0]
        ctx.net, ctx.taps, ctx.diffs, ctx.pool_idx = net, taps, diffs, pool_idx or {}
        ctx.content_target, ctx.feat_mse = content_target, feat_mse
        ctx.in_shape = x.shape
        return losses

    @staticmethod
    def backward(ctx, gl):
        net, taps, diffs = ctx.net, ctx.taps, ctx.diffs
        gl = gl.contiguous().float()
        # dF_l = F_l * D_l  (Gram backward, per-image 1x1 conv on the tensor cores)
        def style_grad(i, add=None):
            tap = taps[i]
            # the per-image weights D take the plane format of the tap they multiply (one format per tensor-core product)
            _keep, wp = ops.style_seed(diffs[i][0], diffs[i][1], tap.H * tap.W, gl[i:i + 1], tap.fmt)
            out, _, _ = ops.conv2d(tap, None, w_ptrs=wp, w_img_rows=tap.C, add=add)
            return out

        t4 = taps[CONTENT_TAP]
        numel4 = t4.buf[0].numel()
        cgrad = ops.mse_bwd(t4, ctx.content_target, gl[5:6], 2.0 / numel4)
        if ctx.feat_mse is not None:
            # d/dx [mse^2/numel] = 2*mse/numel * 2/numel * (relu x - relu t) * (x > 0)
            ops.mse_bwd(t4, ctx.content_target, gl[6:7], 4.0 / (float(numel4) * float(numel4)), out=cgrad,
                        relu=True, mul=ctx.feat_mse)
        g5 = style_grad(4)
        gp2, _, _ = ops.conv2d(g5, net.bwd[4])                                   # -> grad of pool2 output
        g4 = ops.relu_maxpool_bwd(t4, gp2, add=style_grad(3, add=cgrad), idx=ctx.pool_idx.get(3))   # -> grad of tap4
        g3, _, _ = ops.conv2d(g4, net.bwd[3], mask=taps[2], add=style_grad(2))   # relu'(tap3) fused
        gp1, _, _ = ops.conv2d(g3, net.bwd[2])
        g2 = ops.relu_maxpool_bwd(taps[1], gp1, add=style_grad(1), idx=ctx.pool_idx.get(1))
        g1, _, _ = ops.conv2d(g2, net.bwd[1], mask=taps[0], add=style_grad(0))
        if IMAGE_GRAD_TC:
            _, _, gx = ops.conv2d(g1, net.bwd[0], want_out=False, out_nchw_channels=3)
        else:
            gx = ops.conv_image_dgrad(g1, net.w0)
        ctx.taps = ctx.diffs = None
        return gx, None, None, None, None


def vgg_losses(x, net, style_targets, content_target, with_feature=True):
    """losses[0:5] style nodes, [5] content node, [6] feature node (0 when skipped)."""
    ops._require_cuda(x, 'vgg_losses')
    return _VGGLossFn.apply(x.contiguous().float(), net, style_targets, content_target, with_feature)
