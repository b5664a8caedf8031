"""Fused forward/backward of the Johnson transform net on the B200 kernels.

One ``torch.autograd.Function`` evaluates the whole ``nn.Sequential`` of the
reference's ``ImageTransformNet`` / ``VideoTransformNet`` (network.py:521-611,
:850-855; ``ResidualBlock`` :461-506) in the split-bf16 NHWC format, so that the
23 modules exchange no NCHW fp32 tensors: conv (tcgen05 for the 3x3 stride-1
layers, the gather kernel for 9x9 / stride-2) -> InstanceNorm(+ReLU)(+x2 nearest
upsample on write) -> ... .  The backward runs backward-data and backward-weight
of every conv and the InstanceNorm backward (dx, dgamma, dbeta) and returns the
gradient of all 62 parameter tensors plus the input gradient.

Zero padding is the reference arithmetic (``padding_mode='reflection'`` fell
through to zeros under the pinned torch 1.1.0).
"""
import torch

from . import ops

# (state-dict prefix, kind, cin, cout, k, stride) -- nn.Sequential positions of network.py:521-611
LAYERS = [
    ('0', 'conv', 3, 32, 9, 1), ('1', 'in', 32), ('2', 'relu'),
    ('3', 'conv', 32, 64, 3, 2), ('4', 'in', 64), ('5', 'relu'),
    ('6', 'conv', 64, 128, 3, 2), ('7', 'in', 128), ('8', 'relu'),
    ('9', 'res', 128), ('10', 'res', 128), ('11', 'res', 128), ('12', 'res', 128), ('13', 'res', 128),
    ('14', 'up'), ('15', 'conv', 128, 64, 3, 1), ('16', 'in', 64), ('17', 'relu'),
    ('18', 'up'), ('19', 'conv', 64, 32, 3, 1), ('20', 'in', 32), ('21', 'relu'),
    ('22', 'conv', 32, 3, 9, 1),
]


def param_names():
    """The 62 state-dict keys in module order (weights then bias per module)."""
    names = []
    for spec in LAYERS:
        name, kind = spec[0], spec[1]
        if kind in ('conv', 'in'):
            names += [name + '.weight', name + '.bias']
        elif kind == 'res':
            for sub in ('conv1', 'insn1', 'conv2', 'insn2'):
                names += [f'{name}.{sub}.weight', f'{name}.{sub}.bias']
    return names


PARAM_NAMES = param_names()


def plan():
    """Flatten LAYERS into fused stages:
       ('conv', wname, k, stride, residual_from) and ('in', name, relu, up)."""
    stages = []
    specs = list(LAYERS)
    i = 0

    def norm_flags(j):
        """relu / up flags of the modules that follow position j (an InstanceNorm)."""
        relu = up = False
        k = j + 1
        while k < len(specs) and specs[k][1] in ('relu', 'up'):
            relu |= specs[k][1] == 'relu'
            up |= specs[k][1] == 'up'
            k += 1
        return relu, up, k

    while i < len(specs):
        spec = specs[i]
        name, kind = spec[0], spec[1]
        if kind == 'conv':
            stages.append(('conv', name, spec[4], spec[5], False))
            i += 1
        elif kind == 'in':
            relu, up, nxt = norm_flags(i)
            stages.append(('in', name, relu, up))
            i = nxt
        elif kind == 'res':
            stages.append(('mark',))                                  # residual = x
            stages.append(('conv', name + '.conv1', 3, 1, False))
            stages.append(('in', name + '.insn1', True, False))
            stages.append(('conv', name + '.conv2', 3, 1, True))      # out += residual (epilogue add)
            relu, up, nxt = norm_flags(i)                             # an Upsample may follow the last block
            stages.append(('in', name + '.insn2', False, up))
            i = nxt
        else:
            raise AssertionError(f'unexpected module order at {name}')
    return stages


PLAN = plan()


def _is_conv9_in(w):
    return w.shape[0] == 32 and w.shape[1] in (3, 6) and tuple(w.shape[2:]) == (9, 9)


def _is_conv9_out(w):
    return tuple(w.shape) == (3, 32, 9, 9)


class _PackCache:
    """Packed (split-bf16, K-major) copies of the conv weights, refreshed when the fp32
    master weight changes (optimizer step bumps ``_version``)."""

    def __init__(self):
        self.entries = {}
        self.always_repack = False      # CUDA-graph capture: the packing kernels must be part of every replay
        # {state-dict key: zeroed fp32 buffer}: when set (training closures: the views of the flat gradient
        # bucket that are the parameters' .grad), the backward kernels write the parameter gradients
        # straight into these buffers and autograd gets None for them -- no per-tensor accumulate kernels
        self.grad_sink = None
        # frozen: the entries are kept up to date by the optimizer (optim.FusedAdam.set_packs: the Adam kernel rewrites
        # the packed copies), a lookup never repacks; ``synced`` remembers the parameter versions the packs belong to
        self.frozen = False
        self.synced = None

    def get(self, name, w, mode, **kw):
        key = (name, mode, ops.PRECISION, kw.get('fmt'))
        hit = self.entries.get(key)
        if self.frozen and hit is not None:
            return hit[1]
        tag = (w.data_ptr(), w._version, tuple(w.shape))
        if self.always_repack or hit is None or hit[0] != tag:
            hit = (tag, ops.PackedWeight(w, mode, **kw))
            self.entries[key] = hit
        return hit[1]

    def mark_synced(self, params):
        self.synced = [(p, p._version) for p in params]

    def stale(self):
        """Frozen packs whose fp32 master weight changed behind the optimizer's back (load_state_dict, manual edits)."""
        return self.frozen and self.synced is not None and any(p._version != v for p, v in self.synced)

    def repack(self, named_params):
        """Rewrite every packed copy in place from the current fp32 weights (the buffers keep their addresses: they are
        baked into a captured graph and into the optimizer's table)."""
        for (name, _mode, _prec, _fmt), (_tag, pw) in self.entries.items():
            pw.repack(named_params[name + '.weight'])
        if self.synced is not None:
            self.mark_synced([p for p, _ in self.synced])


def _keep(saved, keep, item):
    saved.append(item if keep else None)


def run_forward(x, cache, params, keep, x2=None):
    """The fused forward over PLAN.  ``keep``: retain what the backward needs (training);
    without it every intermediate is released as soon as the next stage has consumed it
    (inference: process_image / process_video, network.py:798-832, :1117-1134).
    ``x2``: second 3-channel source of the video net's 6-channel input (read in place of torch.cat([x, x2], 1))."""
    p = dict(zip(PARAM_NAMES, params))
    N, cin, H, W = x.shape
    if x2 is not None:
        cin = 6
    h = None            # the first conv reads the NCHW fp32 input itself (9x9 direct kernel)
    saved = []          # per stage: what backward needs
    residual = None
    out_nchw = None
    n_conv = sum(1 for s in PLAN if s[0] == 'conv')
    seen = 0
    for st in PLAN:
        if st[0] == 'mark':
            residual = h
            _keep(saved, keep, None)
        elif st[0] == 'conv':
            _, name, k, stride, add_res = st
            seen += 1
            w, b = p[name + '.weight'], p[name + '.bias']
            last = seen == n_conv
            if seen == 1 and _is_conv9_in(w):
                # layer 0: 3|6 -> 32 channels, 9x9: exact-fp32 direct kernel on the NCHW input
                h = ops.conv9_expand(x, w, b, x2=x2)
                _keep(saved, keep, None)
                continue
            if h is None:
                h = ops.from_nchw(x, x2, cpad=ops.pad8(cin))
            if last and _is_conv9_out(w) and h.C == 32:
                # layer 22: 32 -> 3 channels, 9x9: direct kernel writing the NCHW fp32 output
                out_nchw = ops.conv9_reduce(h, w, b)
                _keep(saved, keep, h)
                continue
            pw = cache.get(name, w, 0, cols_pad=h.C)
            if last:
                _, _, out_nchw = ops.conv2d(h, pw, b, stride=stride, want_out=False, out_nchw_channels=w.shape[0])
                _keep(saved, keep, h)
            else:
                y, _, _ = ops.conv2d(h, pw, b, stride=stride, add=residual if add_res else None)
                _keep(saved, keep, h)
                h = y
        else:
            _, name, relu, up = st
            y, stats = ops.instnorm_fwd(h, p[name + '.weight'], p[name + '.bias'], relu=relu, up=up)
            _keep(saved, keep, (h, stats))
            h = y
    return out_nchw, saved


class TransformNetFn(torch.autograd.Function):
    """y = net(x) (or net(cat([x, x2], 1)) read from its two sources).  args: x NCHW fp32 [N, Cin, H, W],
    x2 (None or a second [N, 3, H, W] source), cache, then the 62 parameters in PARAM_NAMES order."""

    @staticmethod
    def forward(ctx, x, x2, cache, *params):
        out_nchw, saved = run_forward(x, cache, params, keep=True, x2=x2)
        shape = x.shape if x2 is None else (x.shape[0], 6) + tuple(x.shape[2:])
        ctx.saved, ctx.cache, ctx.params, ctx.in_shape, ctx.x, ctx.x2 = saved, cache, params, shape, x, x2
        return out_nchw

    @staticmethod
    def backward(ctx, gy):
        p = dict(zip(PARAM_NAMES, ctx.params))
        need = dict(zip(PARAM_NAMES, ctx.needs_input_grad[3:]))
        want_gx = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        grads = {}
        cache = ctx.cache
        sink = cache.grad_sink or {}

        def out_of(name):
            """The sink buffer of a parameter gradient (the kernel writes it; autograd then gets None)."""
            return sink.get(name) if need[name] else None
        N, cin, H, W = ctx.in_shape
        gy = gy.contiguous().float()
        # f16x3 mode: gradients that meet forward activations in a tensor-core product are fp16 planes under a
        # power-of-two scale slot (InstanceNorm-backward outputs carry their own; the gradient image gets one here)
        dyf = ops.dy_fmt()
        gy_slot = None
        g = None             # built lazily: the 9x9 output layer consumes the NCHW fp32 gradient directly
        res_grad = None      # gradient flowing into the residual branch of the current block
        pending_db = None    # sum of dy per channel from the InstanceNorm backward right after a conv
        first_conv = next(i for i, s in enumerate(PLAN) if s[0] == 'conv')
        last_conv = max(i for i, s in enumerate(PLAN) if s[0] == 'conv')
        for idx in range(len(PLAN) - 1, -1, -1):
            st = PLAN[idx]
            if st[0] == 'conv' and idx == last_conv and _is_conv9_out(p[st[1] + '.weight']) and ctx.saved[idx].C == 32:
                name = st[1]
                w = p[name + '.weight']
                if need[name + '.weight'] or need[name + '.bias']:
                    if dyf:
                        gy_slot = ops.absmax_slot(gy)
                    grads[name + '.weight'], grads[name + '.bias'] = ops.conv9_wgrad(
                        ctx.saved[idx], gy, 1, dw_out=out_of(name + '.weight'), db_out=out_of(name + '.bias'), small_slot=gy_slot)
                g = ops.conv9_expand(gy, w, flip=True, planes=ops.bwd_planes(), fmt=0)      # feeds an InstanceNorm backward
                continue
            if g is None:
                g = ops.from_nchw(gy, cpad=8, planes=ops.bwd_planes())
            if st[0] == 'conv' and idx == first_conv and ctx.saved[idx] is None:
                name = st[1]
                w = p[name + '.weight']
                if need[name + '.weight'] or need[name + '.bias']:
                    grads[name + '.weight'], grads[name + '.bias'] = ops.conv9_wgrad(
                        g, ctx.x, 0, dw_out=out_of(name + '.weight'), db_out=out_of(name + '.bias'), small2=ctx.x2)
                if want_gx:
                    # rare (tests / image optimisation through the net): gather-form backward-data
                    pw = cache.get(name, w, 1, rows_pad=ops.pad8(cin), cols_pad=g.C, fmt=g.fmt)
                    _, _, gx = ops.conv2d(g, pw, stride=1, transposed=1, out_hw=(H, W), want_out=False, out_nchw_channels=cin)
                    grads['__input__'] = gx
                break
            if st[0] == 'mark':
                # g is the gradient wrt conv1's input; the residual branch was added by conv1's dgrad epilogue
                res_grad = None
            elif st[0] == 'conv':
                _, name, k, stride, add_res = st
                xin = ctx.saved[idx]
                w = p[name + '.weight']
                cout, cin_w = w.shape[0], w.shape[1]
                if need[name + '.weight'] or need[name + '.bias']:
                    # the bias gradient (sum of dy) came out of the InstanceNorm backward that produced dy
                    dw, db = ops.conv2d_wgrad(g, xin, k, stride, k // 2, 0, cout, cin_w, want_bias=pending_db is None,
                                              dw_out=out_of(name + '.weight'),
                                              db_out=out_of(name + '.bias') if pending_db is None else None)
                    grads[name + '.weight'], grads[name + '.bias'] = dw, (db if pending_db is None else pending_db[:cout])
                pending_db = None
                if add_res:
                    res_grad = g                    # d(out)/d(residual) = identity
                if idx == first_conv:
                    if want_gx:
                        pw = cache.get(name, w, 1, rows_pad=xin.C, cols_pad=g.C, fmt=g.fmt)
                        _, _, gx = ops.conv2d(g, pw, stride=stride, transposed=1, out_hw=(xin.H, xin.W),
                                              want_out=False, out_nchw_channels=cin)
                        grads['__input__'] = gx
                    break
                tc = k == 3 and stride == 1
                # the conv after a residual 'mark' (conv1) adds the block's residual gradient
                add = res_grad if (PLAN[idx - 1][0] == 'mark') else None
                # backward-data weights in the format of the gradient they multiply
                if tc:
                    pw = cache.get(name, w, 2, rows_pad=xin.C, cols_pad=g.C, fmt=g.fmt)
                    g, _, _ = ops.conv2d(g, pw, add=add)
                else:
                    pw = cache.get(name, w, 1, rows_pad=xin.C, cols_pad=g.C, fmt=g.fmt)
                    g, _, _ = ops.conv2d(g, pw, stride=stride, transposed=1, out_hw=(xin.H, xin.W), add=add)
            else:
                _, name, relu, up = st
                xin, stats = ctx.saved[idx]
                # the bias gradient of the convolution in front (the stage before an 'in' stage is its conv)
                conv_bias = PLAN[idx - 1][1] + '.bias'
                dsum_out = out_of(conv_bias) if (PLAN[idx - 1][0] == 'conv' and p[conv_bias].numel() == xin.C) else None
                g, dgam, dbet, pending_db = ops.instnorm_bwd(xin, g, stats, p[name + '.weight'], p[name + '.bias'], relu=relu, up=up,
                                                             want_dsum=True, dgamma_out=out_of(name + '.weight'),
                                                             dbeta_out=out_of(name + '.bias'), dsum_out=dsum_out)
                grads[name + '.weight'], grads[name + '.bias'] = dgam, dbet
        ctx.saved = None
        gx = grads.get('__input__')
        if ctx.x2 is not None and gx is not None:
            out = [gx[:, :3].contiguous(), gx[:, 3:].contiguous(), None]
        else:
            out = [gx, None, None]
        for n in PARAM_NAMES:
            gn = grads.get(n) if need[n] else None
            # a gradient the kernels wrote into its sink (same storage) is already where autograd would add it
            if gn is not None and n in sink and gn.data_ptr() == sink[n].data_ptr():
                gn = None
            out.append(gn)
        return tuple(out)
