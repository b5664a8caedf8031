"""Single-pass VGG-19 loss network on the B200 kernels.

Replaces the prefix passes that StyleNetwork.forward runs once PER LOSS NODE (reference network.py:366-401 ->
run_through_pieces :316-340: seven passes of the input plus two of the content image with the default layer lists)
by ONE forward that emits every tapped layer, and autograd's backward through the replicated prefixes by ONE fused
backward-data chain to the image.  VGG weights are frozen (no wgrad).  Losses come back as one vector
[style nodes..., content nodes..., feature nodes...] whose gradient weights (grad_output) are read on the device, so
no host synchronisation happens anywhere in forward or backward.

The tap set is whatever ``StyleNetwork.content_layers / style_layers / feature_loss_layers`` name (the reference's
piece builder, network.py:257-314, is generic over those class attributes): a loss node may sit on any
``Conv2d_k`` (pre-ReLU output of the k-th conv), ``ReLU_k`` or ``MaxPool2d_k`` layer of torchvision's
``vgg19().features`` (k = 1..16 counts the convolutions).
"""
import os

import torch

from . import ops

# torchvision VGG-19 configuration "E": 16 convs (3x3, zero padding 1), each followed by a ReLU, 5 max-pools
VGG_CFG_E = [64, 64, 'M', 128, 128, 'M', 256, 256, 256, 256, 'M', 512, 512, 512, 512, 'M', 512, 512, 512, 512, 'M']


def _blocks():
    """[(cin, cout, pool_after)] for conv k = 1..16: block k = Conv2d_k -> ReLU_k (-> MaxPool2d_k)."""
    out, cin = [], 3
    for i, v in enumerate(VGG_CFG_E):
        if v == 'M':
            continue
        out.append((cin, v, i + 1 < len(VGG_CFG_E) and VGG_CFG_E[i + 1] == 'M'))
        cin = v
    return out


BLOCKS = _blocks()
# (cin, cout, pool_before) of the five convs the DEFAULT tap set executes (kept for callers that size weights by it)
PREFIX = [(3, 64, False), (64, 64, False), (64, 128, True), (128, 128, False), (128, 256, True)]
KINDS = {'Conv2d': 'C', 'ReLU': 'R', 'MaxPool2d': 'P'}
# backward-data of conv1_1: tcgen05 16-channel tile (default) or the exact-fp32 CUDA-core kernel
IMAGE_GRAD_TC = True
# MaxPool2d + ReLU backward of a block folded into the epilogue of the block's first Gram-backward launch
# (ops.conv2d_routed); STRANSFER_FUSE_POOL_ROUTE=0 keeps the separate relu_maxpool_bwd launch
FUSE_POOL_ROUTE = os.environ.get('STRANSFER_FUSE_POOL_ROUTE', '1') != '0'
# ReLU + MaxPool2d forward folded into the epilogue of the conv in front of it (ops.conv2d_pool); =0 keeps the pool launch
FUSE_POOL_FWD = os.environ.get('STRANSFER_FUSE_POOL_FWD', '1') != '0'


def parse_layer(name):
    """'Conv2d_4' -> ('C', 4), 'ReLU_4' -> ('R', 4), 'MaxPool2d_4' -> ('P', 4) (names of network.py:267-270)."""
    kind, _, idx = name.rpartition('_')
    if kind not in KINDS or not idx.isdigit() or not 1 <= int(idx) <= len(BLOCKS):
        raise ValueError(f'"{name}" is not a layer of vgg19().features (Conv2d_k / ReLU_k / MaxPool2d_k, k = 1..16)')
    k = int(idx)
    if kind == 'MaxPool2d' and not BLOCKS[k - 1][2]:
        raise ValueError(f'"{name}": no max-pool follows conv {k} in VGG-19')
    return KINDS[kind], k


def layer_names(n_convs=len(BLOCKS)):
    """The layer names of vgg19().features in order (all 37, or those of the first ``n_convs`` blocks)."""
    names = []
    for k, (_, _, pool) in enumerate(BLOCKS[:n_convs], start=1):
        names += [f'Conv2d_{k}', f'ReLU_{k}'] + ([f'MaxPool2d_{k}'] if pool else [])
    return names


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
    """Packed weights of the first ``len(weights)`` VGG-19 convs and the fused forward over them."""

    def __init__(self, weights):
        assert 1 <= len(weights) <= len(BLOCKS)
        self.device = weights[0][0].device
        self.fwd, self.bwd, self.bias = [], [], []
        for i, (w, b) in enumerate(weights):
            ops._require_cuda(w, 'VGGPrefix')
            assert tuple(w.shape) == (BLOCKS[i][1], BLOCKS[i][0], 3, 3), f'conv {i + 1}: unexpected weight shape {tuple(w.shape)}'
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

    @property
    def n_convs(self):
        return len(self.bias)

    def run(self, x, points, keep_taps=True, on_point=None, pool_idx=None):
        """One forward over the blocks up to the last requested point.  x: NCHW fp32.  ``points``: set of
        ('C' | 'R' | 'P', k), plus ('R!', k) to ask for a MATERIALISED copy of relu(conv k).
        Returns (vals, taps): vals[('C', k)] = pre-ReLU output of conv k, vals[('R', k)] = its ReLU copy where one was
        materialised, vals[('P', k)] = the pooled tensor; taps[k] = pre-ReLU output of conv k for every executed block
        (only where a requested point or a pool needs it when ``keep_taps`` is False).
        ``on_point(point, vals, taps)`` is called as soon as a requested point has been launched.
        ``pool_idx`` (dict): filled with {k: routing byte map of MaxPool2d_k} for the backward."""
        K = max(k for _, k in points)
        if K > self.n_convs:
            raise ops._lib.StransferError(f'the loss layers need VGG conv {K} but only {self.n_convs} conv weights were given')
        vals, taps = {}, {}
        h = None
        for k in range(1, K + 1):
            _, _, pool = BLOCKS[k - 1]
            last = k == K
            need_pool = pool and (not last or ('P', k) in points)
            need_next = not last or need_pool
            # relu(tap) is materialised by the conv epilogue when the next conv reads it directly, or when a
            # loss node wants the ReLU output as a tensor-core operand (style node on ReLU_k)
            want_relu = (need_next and not pool) or ('R!', k) in points
            want_tap = keep_taps or ('C', k) in points or ('R', k) in points or need_pool or not want_relu
            pooled = None
            if k == 1:
                tap, relu = ops.conv_image_fwd(x.contiguous().float(), self.w0, self.bias[0], want_out=want_tap, want_relu=want_relu)
            else:
                if need_pool and not want_relu and FUSE_POOL_FWD:
                    # ReLU + MaxPool2d_k ride on the epilogue of conv k (the pool does not re-read the tap)
                    pooled = ops.conv2d_pool(h, self.fwd[k - 1], self.bias[k - 1], want_out=want_tap)
                if pooled is not None:
                    tap, relu = pooled[0], None
                else:
                    tap, relu, _ = ops.conv2d(h, self.fwd[k - 1], self.bias[k - 1], want_out=want_tap, want_relu=want_relu)
            taps[k] = tap
            if tap is not None:
                vals[('C', k)] = tap
            if relu is not None:
                vals[('R', k)] = relu
            for pt in (('C', k), ('R', k)):
                if pt in points and on_point is not None:
                    on_point(pt, vals, taps)
            if need_pool:
                if pooled is not None:
                    h = pooled[1]
                    if pool_idx is not None:
                        pool_idx[k] = pooled[2]
                elif pool_idx is not None:
                    h, pool_idx[k] = ops.relu_maxpool_fwd(tap, want_idx=True)
                else:
                    h = ops.relu_maxpool_fwd(tap)
                vals[('P', k)] = h
                if ('P', k) in points and on_point is not None:
                    on_point(('P', k), vals, taps)
            elif need_next:
                h = relu
        return vals, taps

    def taps(self, x, upto=5, keep_all=True):
        """Pre-ReLU outputs of Conv2d_1..upto (list of Split; None where ``keep_all`` is False and the tap is not
        needed).  Convenience for the default five-tap layout."""
        pts = {('C', k) for k in range(1, upto + 1)} if keep_all else {('C', upto)}
        _, taps = self.run(x, pts, keep_taps=keep_all)
        return [taps.get(k) if (keep_all or k == upto) else None for k in range(1, upto + 1)]

    def value_at(self, x, point):
        """NCHW fp32 value of the layer ``point`` for the image batch x (run_through_pieces)."""
        kind, k = point
        vals, taps = self.run(x, {point}, keep_taps=False)
        if kind == 'P':
            return ops.to_nchw(vals[point])
        out = ops.to_nchw(taps[k])
        return out.clamp_(min=0) if kind == 'R' else out


class LossSpec:
    """The loss nodes of a StyleNetwork in loss-vector order: style nodes, content nodes, feature nodes.
    ``style``: [[point, target Gram [1, C, C]]]; ``content`` / ``feature``: [[point, target]] where the target of a
    'C' or 'R' point is the PRE-ReLU tap of the target image at that conv (the kernels apply the ReLU of an 'R' point
    themselves) and the target of a 'P' point is the pooled tensor."""

    def __init__(self, style, content, feature):
        self.style, self.content, self.feature = [list(t) for t in style], [list(t) for t in content], [list(t) for t in feature]

    @property
    def n_losses(self):
        return len(self.style) + len(self.content) + len(self.feature)

    def points(self, with_feature=True, style=True, mse=True):
        pts = set()
        if style:
            for pt, _ in self.style:
                pts.add(pt)
                if pt[0] == 'R':        # a Gram of a ReLU output needs the ReLU copy as a tensor
                    pts.add(('R!', pt[1]))
        if mse:
            for pt, _ in self.content + (self.feature if with_feature else []):
                pts.add(pt)
        return pts


def mse_operand(pt, vals, taps):
    """(tensor, relu flag) the MSE kernels read for a node at ``pt``."""
    if pt[0] == 'P':
        return vals[pt], False
    return taps[pt[1]], pt[0] == 'R'


class _VGGLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, net, spec, with_feature):
        n_style, n_content = len(spec.style), len(spec.content)
        feature = spec.feature if with_feature else []
        losses = torch.zeros(spec.n_losses, dtype=torch.float32, device=x.device)
        diffs = [None] * n_style
        feat_mse = [None] * len(feature)
        by_point = {}
        for i, (pt, _) in enumerate(spec.style):
            by_point.setdefault(pt, []).append(('style', i))
        for i, (pt, _) in enumerate(spec.content):
            by_point.setdefault(pt, []).append(('content', i))
        for i, (pt, _) in enumerate(feature):
            by_point.setdefault(pt, []).append(('feature', i))
        # The loss branch of a tap (Gram -> split-K finish -> style MSE, content / feature MSE) does not feed the
        # next layer: with STRANSFER_OVERLAP_LOSSES it runs on a side stream beside the next conv (inside a captured
        # CUDA graph the two streams become parallel branches).  The streams join before forward returns.
        cur = torch.cuda.current_stream(x.device)
        side = _side_stream(x.device) if OVERLAP_LOSS_BRANCHES else None

        def branch(pt, vals, taps):
            for kind, i in by_point.get(pt, ()):
                if kind == 'style':
                    # Gram + StyleLoss.forward: the finish launch of the Gram also forms diff, the loss and max |diff|
                    diffs[i] = ops.gram_style_fwd(vals[pt], spec.style[i][1], losses[i:i + 1])
                elif kind == 'content':
                    t, relu = mse_operand(pt, vals, taps)
                    ops.mse_fwd(t, spec.content[i][1], losses[n_style + i:n_style + i + 1], relu=relu)
                else:
                    t, relu = mse_operand(pt, vals, taps)
                    feat_mse[i] = torch.empty(1, dtype=torch.float32, device=x.device)
                    ops.mse_fwd(t, feature[i][1], feat_mse[i], relu=relu)
                    j = n_style + n_content + i
                    ops.mse_fwd(t, feature[i][1], losses[j:j + 1], relu=relu, feature_form=True)

        def on_point(pt, vals, taps):
            if side is None:
                return branch(pt, vals, taps)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                branch(pt, vals, taps)
        pool_idx = {} if ctx.needs_input_grad[0] else None       # routing maps of the max-pools, for the backward
        vals, taps = net.run(x, spec.points(with_feature), on_point=on_point, pool_idx=pool_idx)
        if side is not None:
            cur.wait_stream(side)
        ctx.net, ctx.spec, ctx.feature = net, spec, feature
        ctx.vals, ctx.taps, ctx.diffs, ctx.feat_mse, ctx.pool_idx = vals, taps, diffs, feat_mse, pool_idx or {}
        return losses

    @staticmethod
    def backward(ctx, gl):
        net, spec, vals, taps, diffs, feature = ctx.net, ctx.spec, ctx.vals, ctx.taps, ctx.diffs, ctx.feature
        n_style, n_content = len(spec.style), len(spec.content)
        gl = gl.contiguous().float()
        K = max(taps)

        # the per-image weights D of every Gram backward (dF = F * D) in one launch: they depend on the loss gradient
        # only; D takes the plane format of the operand F it multiplies (one format per tensor-core product)
        seeds = ops.style_seed_multi([(diffs[i][0], diffs[i][1], vals[p].H * vals[p].W, gl[i:i + 1], vals[p].fmt)
                                      for i, (p, _) in enumerate(spec.style)]) if n_style else []

        def node_grads(pt, acc=None, skip=None):
            """acc + the loss gradients of the nodes at ``pt`` with respect to the value at ``pt`` -- for 'R' points
            with respect to the pre-ReLU tap (relu' applied).  Returns ``acc`` unchanged when no node sits there.
            ``skip``: index of a style node whose gradient ``acc`` already holds."""
            for i, (p, target) in enumerate(spec.content):
                if p == pt:
                    t, relu = mse_operand(pt, vals, taps)
                    acc = ops.mse_bwd(t, target, gl[n_style + i:n_style + i + 1], 2.0 / t.buf[0].numel(), out=acc, relu=relu)
            for i, (p, target) in enumerate(feature):
                if p == pt:
                    # d/dx [mse^2 / numel] = 2 mse / numel * 2 / numel * (f(x) - f(t)) * f'(x)
                    t, relu = mse_operand(pt, vals, taps)
                    n = float(t.buf[0].numel())
                    j = n_style + n_content + i
                    acc = ops.mse_bwd(t, target, gl[j:j + 1], 4.0 / (n * n), out=acc, relu=relu, mul=ctx.feat_mse[i])
            for i, (p, _) in enumerate(spec.style):
                if p == pt and i != skip:
                    # dF = F * D  (Gram backward, per-image 1x1 conv on the tensor cores); the per-image weights D take
                    # the plane format of the operand F they multiply (one format per tensor-core product)
                    f = vals[pt]
                    _keep, wp = seeds[i]
                    acc, _, _ = ops.conv2d(f, None, w_ptrs=wp, w_img_rows=f.C, add=acc, mask=taps[pt[1]] if pt[0] == 'R' else None)
            return acc

        def tap_node_grads(k):
            return node_grads(('R', k), node_grads(('C', k)))

        def routed_tap_grads(k, gy, idx):
            """Gradient of the pre-ReLU tap of a block that ends in a max-pool: the pool's output gradient ``gy`` routed
            through the pool and the ReLU, + the block's node gradients -- with the routing done by the epilogue of the
            first Gram-backward launch of the block (one pass over the tap-sized gradient instead of three).  None when
            the block has no style node or that launch cannot take the tensor-core kernel."""
            if idx is None or not FUSE_POOL_ROUTE:
                return None
            first = next((i for i, (p, _) in enumerate(spec.style) if p in (('C', k), ('R', k))), None)
            if first is None:
                return None
            pt = spec.style[first][0]
            acc = ops.conv2d_routed(vals[pt], seeds[first][1], vals[pt].C, gy, idx, mask=taps[k] if pt[0] == 'R' else None)
            if acc is None:
                return None
            return node_grads(('R', k), node_grads(('C', k), acc, skip=first), skip=first)

        g_out = None          # gradient of the OUTPUT of block k (what conv k + 1 reads), node gradients of MaxPool2d_k included
        g_tap = None          # gradient of the pre-ReLU output of conv k when the launch below has already produced it
        for k in range(K, 0, -1):
            _, _, pool = BLOCKS[k - 1]
            if g_tap is None:
                gy = None
                if pool and (g_out is not None or ('P', k) in vals):
                    gy = g_out if g_out is not None else node_grads(('P', k))
                if gy is not None:
                    g_tap = routed_tap_grads(k, gy, ctx.pool_idx.get(k))
                    if g_tap is None:
                        # route through the max-pool, relu', + the node gradients of this block, in one kernel
                        g_tap = ops.relu_maxpool_bwd(taps[k], gy, add=tap_node_grads(k), idx=ctx.pool_idx.get(k))
                else:
                    g_tap = tap_node_grads(k)         # top block: the node gradients are the whole gradient
                if g_tap is None:
                    raise ops._lib.StransferError('internal: no gradient reaches the top of the loss network')
            if k == 1:
                if IMAGE_GRAD_TC:
                    _, _, gx = ops.conv2d(g_tap, net.bwd[0], want_out=False, out_nchw_channels=3)
                else:
                    gx = ops.conv_image_dgrad(g_tap, net.w0)
                break
            # backward-data of conv k = gradient of block k - 1's output
            if BLOCKS[k - 2][2]:
                # a max-pool closes block k - 1: its node gradients ride on this launch's epilogue, the pool is routed next
                g_out, _, _ = ops.conv2d(g_tap, net.bwd[k - 1], add=node_grads(('P', k - 1)))
                g_tap = None
            else:
                # no pool: this launch's epilogue applies relu'(tap k - 1) and adds block k - 1's node gradients
                g_tap, _, _ = ops.conv2d(g_tap, net.bwd[k - 1], mask=taps[k - 1], add=tap_node_grads(k - 1))
                g_out = None
        ctx.vals = ctx.taps = ctx.diffs = None
        return gx, None, None, None


def vgg_losses(x, net, spec, with_feature=True):
    """Loss vector [style nodes..., content nodes..., feature nodes...] (feature entries 0 when skipped)."""
    ops._require_cuda(x, 'vgg_losses')
    return _VGGLossFn.apply(x.contiguous().float(), net, spec, with_feature)
