"""Thin Python wrappers over the C ABI: device buffers in the split-bf16 NHWC
format and one function per entry point.  PyTorch is used for memory and streams
only; every numeric op below is a kernel of libstransfer_b200.so.
"""
import ctypes
import os

import torch

from . import _lib

F16 = 16        # plane-code bit (include/stransfer_b200.h ST_PLANES_F16): the planes hold fp16 instead of bf16

# precision mode -> (planes, format) of forward activations / weights, planes of gradients, format of the gradients
# that feed tensor-core products together with forward activations ("dy": the InstanceNorm-backward outputs)
#  'f16x3' : forward 2 fp16 planes (22 mantissa bits) / 3 tensor-core passes per product; gradients 2 bf16 planes /
#            3 passes, except the InstanceNorm-backward outputs, which are 2 fp16 planes under a per-tensor power-of-two
#            scale (the tensor cores cannot multiply fp16 activations with bf16 gradients).  4 bytes per element.
#  'bf16x6': forward 3 bf16 planes / 6 passes, backward 2 planes / 3 passes; 6 bytes per forward element.
#            Forward values of both are fp32-accurate, so ReLU masks and max-pool argmaxes agree with the
#            reference (each flip costs a whole gradient entry).
#  'bf16x3': 2 bf16 planes / 3 passes everywhere (loss 1e-5, gradient ~2e-3 at >= 128^2 images)
#  'bf16'  : 1 plane / 1 pass (fast, ~3e-2 error on the Gatys gradient)
_MODES = {'f16x3': (2, F16, 2, F16), 'bf16x6': (3, 0, 2, 0), 'bf16x3': (2, 0, 2, 0), 'bf16': (1, 0, 1, 0)}
PRECISION = os.environ.get('STRANSFER_PRECISION', 'f16x3')
assert PRECISION in _MODES, PRECISION
W_F16_INV_SCALE = 1.0 / 256.0      # fp16-packed conv weights carry the fixed factor 256 (csrc/common.cuh)


def set_precision(mode):
    global PRECISION
    assert mode in _MODES
    PRECISION = mode


def fwd_planes():
    return _MODES[PRECISION][0]


def fwd_fmt():
    return _MODES[PRECISION][1]


def bwd_planes():
    return _MODES[PRECISION][2]


def dy_fmt():
    """Format of the gradients that meet forward activations in a tensor-core product."""
    return _MODES[PRECISION][3]


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _require_cuda(t, what):
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise _lib.StransferError(f'{what}: expected a CUDA tensor (there is no CPU fallback)')


def pad8(c):
    return (c + 7) // 8 * 8


class Split:
    """NHWC activation stored as 16-bit hi (+ lo (+ third)) planes in one allocation.  ``fmt``: 0 = bf16 planes,
    F16 = fp16 planes (the storage dtype stays torch.bfloat16: a 16-bit slot).  ``slot``: optional 2-float device
    tensor {s, 1/s}: the planes hold s * value (a gradient under a power-of-two scale)."""
    __slots__ = ('buf', 'N', 'H', 'W', 'C', 'hi', 'lo', 'planes', 'fmt', 'slot')

    def __init__(self, N, H, W, C, device, planes=None, zero=False, fmt=None, scaled=False):
        if planes is None:
            planes, fmt = fwd_planes(), (fwd_fmt() if fmt is None else fmt)
        fmt = fmt or 0
        assert not (fmt and planes > 2), 'fp16 tensors have at most two planes'
        alloc = torch.zeros if zero else torch.empty
        self.buf = alloc((planes, N, H, W, C), dtype=torch.bfloat16, device=device)
        self.N, self.H, self.W, self.C, self.planes, self.fmt = N, H, W, C, planes, fmt
        self.hi = self.buf[0].data_ptr()
        self.lo = self.buf[1].data_ptr() if planes > 1 else None   # plane 3 sits at lo + (lo - hi)
        self.slot = torch.empty(2, dtype=torch.float32, device=device) if scaled else None

    @property
    def shape(self):
        return (self.N, self.H, self.W, self.C)

    @property
    def code(self):
        """Plane code of the C ABI."""
        return self.planes | self.fmt

    def read_code(self, max_planes):
        """Plane code for reading at most ``max_planes`` planes."""
        return min(self.planes, max_planes) | self.fmt

    @property
    def slot_ptr(self):
        return None if self.slot is None else self.slot.data_ptr()

    def float(self):
        """fp32 NHWC view of the value (sum of the planes, unscaled) -- for tests/debugging."""
        planes = self.buf.view(torch.float16) if self.fmt else self.buf
        v = planes[0].float()
        for p in range(1, self.planes):
            v = v + planes[p].float()
        return v if self.slot is None else v * self.slot[1]

    def set_from_float(self, v):
        """Fill the planes from an fp32 NHWC tensor with torch ops (targets / tests, off the hot path)."""
        planes = self.buf.view(torch.float16) if self.fmt else self.buf
        rest = v.float()
        for p in range(self.planes):
            plane = rest.clamp(-65504.0, 65504.0).to(torch.float16) if self.fmt else rest.to(torch.bfloat16)
            planes[p].copy_(plane)
            rest = rest - plane.float()
        return self

    def like(self):
        return Split(self.N, self.H, self.W, self.C, self.buf.device, planes=self.planes, fmt=self.fmt)

    def grad_like(self):
        return Split(self.N, self.H, self.W, self.C, self.buf.device, planes=bwd_planes(), fmt=0)

    def dy_like(self):
        """A gradient of this shape that will meet forward activations in a tensor-core product: fp16 planes
        under a scale slot in the f16x3 mode, plain bf16 planes otherwise."""
        return Split(self.N, self.H, self.W, self.C, self.buf.device, planes=bwd_planes(), fmt=dy_fmt(), scaled=bool(dy_fmt()))


def _p(t):
    return None if t is None else t.data_ptr()


def from_nchw(x, x2=None, cpad=None, planes=None, fmt=None):
    """NCHW fp32 -> Split (channel-concat of x and x2, padded with zeros to cpad).  Default: the forward format;
    with ``planes`` given: bf16 planes unless ``fmt`` says otherwise."""
    _require_cuda(x, 'from_nchw')
    x = x.contiguous().float()
    N, C1, H, W = x.shape
    C2 = 0
    if x2 is not None:
        x2 = x2.contiguous().float()
        C2 = x2.shape[1]
    cpad = cpad or pad8(C1 + C2)
    out = Split(N, H, W, cpad, x.device, planes=planes, fmt=fmt)
    _lib.check(_lib.lib().st_nchw_to_nhwc(x.data_ptr(), _p(x2), out.hi, out.lo, out.code, N, C1, C2, H, W, cpad, _stream()),
               'st_nchw_to_nhwc')
    return out


def to_nchw(s, C=None):
    C = C or s.C
    out = torch.empty((s.N, C, s.H, s.W), dtype=torch.float32, device=s.buf.device)
    _lib.check(_lib.lib().st_nhwc_to_nchw(s.hi, s.lo, s.code, s.slot_ptr, out.data_ptr(), s.N, C, s.H, s.W, s.C, _stream()), 'st_nhwc_to_nchw')
    return out


class PackedWeight:
    """nn.Conv2d weight packed for one of the conv modes (see st_pack_conv_weight).  Forward weights (mode 0)
    carry the forward precision and format; backward-data weights (modes 1, 2) the gradient precision, in the
    format of the gradient they multiply (``fmt``: bf16 unless the caller says fp16)."""
    __slots__ = ('buf', 'hi', 'lo', 'rows', 'cols', 'ks', 'mode', 'cout', 'cin', 'planes', 'fmt')

    def __init__(self, w, mode, rows_pad=None, cols_pad=None, planes=None, fmt=None):
        _require_cuda(w, 'PackedWeight')
        w = w.detach().contiguous().float()
        cout, cin, ks, _ = w.shape
        rows, cols = (cout, cin) if mode == 0 else (cin, cout)
        self.rows, self.cols = rows_pad or pad8(rows), cols_pad or pad8(cols)
        self.ks, self.mode, self.cout, self.cin = ks, mode, cout, cin
        self.planes = planes or (fwd_planes() if mode == 0 else bwd_planes())
        self.fmt = (fwd_fmt() if mode == 0 else 0) if fmt is None else fmt
        self.buf = torch.empty((self.planes, self.rows, ks * ks, self.cols), dtype=torch.bfloat16, device=w.device)
        self.hi = self.buf[0].data_ptr()
        self.lo = self.buf[1].data_ptr() if self.planes > 1 else None
        self.repack(w)

    def repack(self, w):
        """(Re)write the packed planes in place from the fp32 weight."""
        w = w.detach().contiguous().float()
        assert tuple(w.shape) == (self.cout, self.cin, self.ks, self.ks)
        _lib.check(_lib.lib().st_pack_conv_weight(w.data_ptr(), self.hi, self.lo, self.code, self.cout, self.cin, self.ks, self.rows,
                                                  self.cols, self.mode, _stream()), 'st_pack_conv_weight')

    @property
    def code(self):
        return self.planes | self.fmt

    @property
    def acc_mul(self):
        """What the consumer multiplies its accumulators by (fp16 weights are packed as 256 * w)."""
        return W_F16_INV_SCALE if self.fmt else 1.0


def conv2d(x, w, bias=None, stride=1, pad=None, up=0, transposed=0, out_hw=None,
           mask=None, add=None, want_out=True, want_relu=False, out_nchw_channels=0, w_img_rows=0, w_ptrs=None):
    """Generic entry to st_conv2d.  `w` is a PackedWeight (or None with w_ptrs =
    (hi, lo, rows_total, ks, cols, plane code, slot tensor | None) for the per-image Gram-backward weights).
    Returns (out Split|None, relu Split|None, out_nchw Tensor|None)."""
    if w is not None:
        w_hi, w_lo, rows_total, ks, cols = w.hi, w.lo, w.rows, w.ks, w.cols
        cout_pad, w_code, forward, acc_mul, w_slot = w.rows, w.code, w.mode == 0, w.acc_mul, None
    else:
        w_hi, w_lo, rows_total, ks, cols, w_code, w_slot = w_ptrs
        cout_pad, forward, acc_mul = w_img_rows, False, 1.0
    # forward convs run at the forward precision and write the forward format; backward convs read at most the
    # gradient precision from their (possibly 3-plane) operands and write bf16 gradient planes
    if forward:
        out_planes, out_fmt, x_code = fwd_planes(), fwd_fmt(), x.code
    else:
        out_planes, out_fmt, x_code = bwd_planes(), 0, x.read_code(bwd_planes())
    assert cols == x.C, f'weight expects {cols} input channels, activation has {x.C}'
    assert not (x.slot is not None and w_slot is not None), 'at most one scaled operand'
    acc_slot = x.slot if x.slot is not None else w_slot
    pad = ks // 2 if pad is None else pad
    Hl, Wl = x.H << up, x.W << up
    if out_hw is not None:
        Ho, Wo = out_hw
    elif transposed:
        Ho, Wo = x.H * stride, x.W * stride
    else:
        Ho, Wo = (Hl + 2 * pad - ks) // stride + 1, (Wl + 2 * pad - ks) // stride + 1
    dev = x.buf.device
    out = Split(x.N, Ho, Wo, cout_pad, dev, planes=out_planes, fmt=out_fmt) if want_out else None
    relu = Split(x.N, Ho, Wo, cout_pad, dev, planes=out_planes, fmt=out_fmt) if want_relu else None
    nchw = torch.empty((x.N, out_nchw_channels, Ho, Wo), dtype=torch.float32, device=dev) if out_nchw_channels else None
    for t in (mask, add):
        assert t is None or t.shape == (x.N, Ho, Wo, cout_pad), 'mask/add shape mismatch'
    cbias = 0 if bias is None else bias.numel()
    _lib.check(_lib.lib().st_conv2d(
        x.hi, x.lo if (x_code & 15) > 1 else None, x_code, x.N, x.H, x.W, x.C, up, w_hi, w_lo, w_code, rows_total, w_img_rows, ks, stride, pad, transposed,
        Ho, Wo, cout_pad, cbias, _p(bias), mask.hi if mask else None,
        add.hi if add else None, add.lo if add else None, add.code if add else 1,
        out.hi if out else None, out.lo if out else None, relu.hi if relu else None, relu.lo if relu else None,
        out_planes | out_fmt, _p(nchw), out_nchw_channels, float(acc_mul), _p(acc_slot), add.slot_ptr if add else None,
        _stream()), 'st_conv2d')
    return out, relu, nchw


def conv2d_pool(x, w, bias, want_out=True):
    """Forward 3x3 conv whose epilogue also applies the ReLU + MaxPool2d(2, 2) that follows it in the VGG pieces
    (reference network.py:257-314): returns (tap Split | None, pooled Split, routing byte map) -- what ``conv2d`` +
    ``relu_maxpool_fwd(want_idx=True)`` return -- or None when the launch cannot take the tensor-core kernel."""
    assert w.mode == 0 and w.ks == 3 and w.cols == x.C
    L = _lib.lib()
    dev = x.buf.device
    out_planes, out_fmt = fwd_planes(), fwd_fmt()
    if (x.code & 16) != (w.code & 16) or x.H < 2 or x.W < 2:
        return None
    out = Split(x.N, x.H, x.W, w.rows, dev, planes=out_planes, fmt=out_fmt) if want_out else None
    pooled = Split(x.N, x.H // 2, x.W // 2, w.rows, dev, planes=out_planes, fmt=out_fmt)
    idx = torch.empty((x.N, x.H // 2, x.W // 2, w.rows), dtype=torch.uint8, device=dev)
    cbias = 0 if bias is None else bias.numel()
    rc = L.st_conv2d_pool(x.hi, x.lo if (x.code & 15) > 1 else None, x.code, x.N, x.H, x.W, x.C, w.hi, w.lo, w.code, w.rows, w.ks, w.rows, cbias,
                          _p(bias), out.hi if out else None, out.lo if out else None, out_planes | out_fmt, pooled.hi, pooled.lo, idx.data_ptr(),
                          float(w.acc_mul), _p(x.slot), _stream())
    if rc == -3:      # ST_ERR_UNSUPPORTED
        return None
    _lib.check(rc, 'st_conv2d_pool')
    return out, pooled, idx


def conv2d_routed(x, w_ptrs, w_img_rows, gy, idx, mask=None):
    """out = x * D (per-image 1x1 weights, the Gram backward) [* (mask > 0)] + the gradient ``gy`` of the OUTPUT of the
    2x2 max-pool that follows, routed through the byte map ``idx`` of relu_maxpool_fwd(want_idx=True): the MaxPool2d and
    ReLU backward of torch autograd (reference network.py:257-314 builds the VGG pieces they belong to) folded into the
    epilogue of the tensor-core launch.  Returns None when this launch cannot take the tensor-core path (the caller then
    runs relu_maxpool_bwd + conv2d)."""
    L = _lib.lib()
    _bf16_grad(gy)
    w_hi, w_lo, rows_total, ks, cols, w_code, w_slot = w_ptrs
    assert cols == x.C and ks == 1 and tuple(idx.shape) == (x.N, x.H // 2, x.W // 2, w_img_rows) and gy.shape == tuple(idx.shape)
    assert not (x.slot is not None and w_slot is not None), 'at most one scaled operand'
    x_code = x.read_code(bwd_planes())
    if (x_code & 16) != (w_code & 16):
        return None
    acc_slot = x.slot if x.slot is not None else w_slot
    out = Split(x.N, x.H, x.W, w_img_rows, x.buf.device, planes=bwd_planes(), fmt=0)
    rc = L.st_conv2d_routed(x.hi, x.lo if (x_code & 15) > 1 else None, x_code, x.N, x.H, x.W, x.C, w_hi, w_lo, w_code, rows_total, w_img_rows, ks,
                            w_img_rows, mask.hi if mask else None, gy.hi, gy.lo, gy.code, idx.data_ptr(), out.hi, out.lo, bwd_planes(),
                            1.0, _p(acc_slot), gy.slot_ptr, _stream())
    if rc == -3:      # ST_ERR_UNSUPPORTED: forced generic path, channel counts outside the tensor-core kernel
        return None
    _lib.check(rc, 'st_conv2d_routed')
    return out


def conv_image_fwd(x, w, bias, want_out=True, want_relu=False):
    """VGG conv1_1 on the NCHW fp32 image (exact fp32, CUDA cores) -> (tap Split, relu Split)."""
    _require_cuda(x, 'conv_image_fwd')
    N, C, H, W = x.shape
    assert C == 3 and tuple(w.shape[1:]) == (3, 3, 3), 'conv_image_fwd: 3-channel image, 3x3 kernel'
    cout = w.shape[0]
    out = Split(N, H, W, cout, x.device) if want_out else None
    relu = Split(N, H, W, cout, x.device) if want_relu else None
    code = (out or relu).code
    _lib.check(_lib.lib().st_conv3x3_image_fwd(x.data_ptr(), w.data_ptr(), _p(bias), out.hi if out else None, out.lo if out else None,
                                               relu.hi if relu else None, relu.lo if relu else None, code, N, H, W, cout, _stream()),
               'st_conv3x3_image_fwd')
    return out, relu


def conv_image_dgrad(g, w):
    """Gradient of the image (NCHW fp32) from the gradient of the conv1_1 tap."""
    gx = torch.empty((g.N, 3, g.H, g.W), dtype=torch.float32, device=g.buf.device)
    _lib.check(_lib.lib().st_conv3x3_image_dgrad(g.hi, g.lo, g.code, w.data_ptr(), gx.data_ptr(), g.N, g.H, g.W, g.C, _stream()),
               'st_conv3x3_image_dgrad')
    return gx


def _small_ptrs(x, x2=None):
    """(s0, s1, channels per sample, total channels) of the "small" side of a 9x9 layer: one NCHW fp32 tensor with
    3 or 6 channels, or TWO 3-channel tensors read as channels 0..2 and 3..5 (the video net's
    cat([frame, previous stylised frame], 1), network.py:984-986, without materialising the concatenation)."""
    N, C, H, W = x.shape
    if x2 is not None:
        assert C == 3 and tuple(x2.shape) == tuple(x.shape), 'two-source input: two [N, 3, H, W] tensors'
        return x.data_ptr(), x2.data_ptr(), 3, 6
    assert C in (3, 6), 'the 9x9 kernels take 3- or 6-channel NCHW tensors'
    return x.data_ptr(), (x.data_ptr() + 3 * H * W * 4 if C == 6 else None), C, C


def conv9_expand(x, w, bias=None, flip=False, planes=None, fmt=None, x2=None):
    """9x9 conv from a 3/6-channel NCHW fp32 tensor to 32 channels (Split).  flip=False: layer-0
    forward (w [32][C][9][9]); flip=True: layer-22 backward-data of dy (w [3][32][9][9]).  The operands
    are built in shared memory in the format of the output planes."""
    _require_cuda(x, 'conv9_expand')
    x = x.contiguous().float()
    x2 = None if x2 is None else x2.contiguous().float()
    N, _, H, W = x.shape
    s0, s1, sc, C = _small_ptrs(x, x2)
    assert tuple(w.shape) == ((C, 32, 9, 9) if flip else (32, C, 9, 9))
    out = Split(N, H, W, 32, x.device, planes=planes, fmt=fmt)
    _lib.check(_lib.lib().st_conv9_expand(s0, s1, sc, C, w.data_ptr(), _p(bias), int(flip), out.hi, out.lo, out.code,
                                          N, H, W, _stream()), 'st_conv9_expand')
    return out


def conv9_reduce(x, w, bias=None):
    """9x9 conv from a 32-channel Split to a 3-channel NCHW fp32 tensor (layer-22 forward)."""
    assert x.C == 32 and tuple(w.shape) == (3, 32, 9, 9)
    L = _lib.lib()
    out = torch.empty((x.N, 3, x.H, x.W), dtype=torch.float32, device=x.buf.device)
    ws = workspace(L.st_conv9_reduce_workspace_bytes(x.N, x.H, x.W), x.buf.device, 'conv9')
    _lib.check(L.st_conv9_reduce(x.hi, x.lo, x.code, w.data_ptr(), _p(bias), out.data_ptr(), ws.data_ptr(), ws.numel(),
                                 x.N, x.H, x.W, _stream()), 'st_conv9_reduce')
    return out


def _grad_out(out, shape, dev):
    """A caller-provided gradient buffer (a view of the flat gradient bucket) or a fresh tensor."""
    if out is None:
        return torch.empty(shape, dtype=torch.float32, device=dev)
    assert out.is_contiguous() and out.dtype == torch.float32 and tuple(out.shape) == tuple(shape), (out.shape, shape)
    return out


def conv9_wgrad(big, small, mode, want_bias=True, dw_out=None, db_out=None, small_slot=None, small2=None):
    """Backward-weight of the 9x9 layers.  mode 0: layer 0 (big = dy Split, small = input NCHW) ->
    dw [32][C][9][9]; mode 1: layer 22 (big = input Split, small = dy NCHW) -> dw [3][32][9][9].
    ``small_slot`` (absmax_slot of the small tensor): the scale under which a gradient image enters fp16 operands."""
    L = _lib.lib()
    small = small.contiguous().float()
    small2 = None if small2 is None else small2.contiguous().float()
    N, _, H, W = small.shape
    s0, s1, sc, C = _small_ptrs(small, small2)
    assert big.C == 32 and big.shape[:3] == (N, H, W)
    dev = big.buf.device
    ws = workspace(L.st_conv9_wgrad_workspace_bytes(C), dev)
    dw = _grad_out(dw_out, (32, C, 9, 9) if mode == 0 else (C, 32, 9, 9), dev)
    db = _grad_out(db_out, (32 if mode == 0 else C,), dev) if want_bias else None
    _lib.check(L.st_conv9_wgrad(big.hi, big.lo, big.read_code(2), big.slot_ptr, _p(small_slot), s0, s1, sc, C, mode, dw.data_ptr(), _p(db), ws.data_ptr(), ws.numel(),
                                N, H, W, _stream()), 'st_conv9_wgrad')
    return dw, db


_ws_cache = {}


def workspace(nbytes, device, tag='ws'):
    """Grow-only scratch buffer per (device, tag, stream)."""
    key = (device, tag, _stream())
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 1 << 16), dtype=torch.uint8, device=device)
        _ws_cache[key] = buf
    return buf


def conv2d_wgrad(dy, x, ks, stride, pad, up, cout, cin, want_bias=True, dw_out=None, db_out=None):
    L = _lib.lib()
    dev = x.buf.device
    nbytes = L.st_conv2d_wgrad_workspace_bytes(dy.N, dy.H, dy.W, dy.C, x.C, ks)
    ws = workspace(nbytes, dev)
    dw = _grad_out(dw_out, (cout, cin, ks, ks), dev)
    db = _grad_out(db_out, (cout,), dev) if want_bias else None
    _lib.check(L.st_conv2d_wgrad(dy.hi, dy.lo, dy.code, dy.slot_ptr, dy.C, x.hi, x.lo, x.read_code(2), x.N, x.H, x.W, x.C, up, dy.H, dy.W, ks, stride, pad,
                                 cout, cin, dw.data_ptr(), _p(db), ws.data_ptr(), ws.numel(), _stream()), 'st_conv2d_wgrad')
    return dw, db


def relu_maxpool_fwd(x, want_idx=False):
    """want_idx: also return the byte map of routing decisions (which of the 4 window positions receives the
    gradient, 4 = none) that relu_maxpool_bwd(idx=...) consumes instead of re-reading x."""
    y = Split(x.N, x.H // 2, x.W // 2, x.C, x.buf.device, planes=x.planes, fmt=x.fmt)
    if want_idx:
        idx = torch.empty((x.N, x.H // 2, x.W // 2, x.C), dtype=torch.uint8, device=x.buf.device)
        _lib.check(_lib.lib().st_relu_maxpool2x2_fwd_idx(x.hi, x.lo, y.hi, y.lo, idx.data_ptr(), x.code, x.N, x.H, x.W, x.C,
                                                         _stream()), 'st_relu_maxpool2x2_fwd_idx')
        return y, idx
    _lib.check(_lib.lib().st_relu_maxpool2x2_fwd(x.hi, x.lo, y.hi, y.lo, x.code, x.N, x.H, x.W, x.C, _stream()), 'st_relu_maxpool2x2_fwd')
    return y


def _bf16_grad(*ts):
    for t in ts:
        assert t is None or (t.fmt == 0 and t.slot is None), 'this kernel reads its gradient operands as bf16 planes'


def relu_maxpool_bwd(x, gy, add=None, idx=None):
    _bf16_grad(gy, add)
    gx = x.grad_like()
    if idx is not None:
        _lib.check(_lib.lib().st_relu_maxpool2x2_bwd_idx(idx.data_ptr(), gy.hi, gy.lo, add.hi if add else None, add.lo if add else None,
                                                         gx.hi, gx.lo, x.N, x.H, x.W, x.C, _stream()), 'st_relu_maxpool2x2_bwd_idx')
        return gx
    _lib.check(_lib.lib().st_relu_maxpool2x2_bwd(x.hi, x.lo, gy.hi, gy.lo, add.hi if add else None, add.lo if add else None,
                                                 gx.hi, gx.lo, x.code, x.N, x.H, x.W, x.C, _stream()), 'st_relu_maxpool2x2_bwd')
    return gx


def gram(f):
    """[B][C][C] fp32 Gram matrices, already divided by C*H*W (network.py:92-108)."""
    L = _lib.lib()
    dev = f.buf.device
    HW = f.H * f.W
    nbytes = L.st_gram_workspace_bytes(f.N, HW, f.C)
    ws = workspace(nbytes, dev)
    G = torch.empty((f.N, f.C, f.C), dtype=torch.float32, device=dev)
    _lib.check(L.st_gram_fwd(f.hi, f.lo, f.read_code(2), f.N, HW, f.C, G.data_ptr(), ws.data_ptr(), ws.numel(), _stream()), 'st_gram_fwd')
    return G


_tickets = {}


def gram_style_fwd(f, T, loss_out):
    """StyleLoss.gram_matrix + StyleLoss.forward of one node in two launches (network.py:92-123): the loss
    mean((G - T)^2) goes to the 1-element view ``loss_out``; returns (diff = G - T, max|diff| as a 1-element tensor)."""
    L = _lib.lib()
    dev = f.buf.device
    HW = f.H * f.W
    ws = workspace(L.st_gram_style_workspace_bytes(f.N, HW, f.C), dev)
    key = (dev, _stream())
    ticket = _tickets.get(key)
    if ticket is None:
        ticket = _tickets[key] = torch.zeros(1, dtype=torch.int32, device=dev)      # the finish kernel hands it back at zero
    diff = torch.empty((f.N, f.C, f.C), dtype=torch.float32, device=dev)
    amax = torch.empty(1, dtype=torch.float32, device=dev)
    _lib.check(L.st_gram_style_fwd(f.hi, f.lo, f.read_code(2), f.N, HW, f.C, T.data_ptr(), diff.data_ptr(), loss_out.data_ptr(),
                                   amax.data_ptr(), ticket.data_ptr(), ws.data_ptr(), ws.numel(), _stream()), 'st_gram_style_fwd')
    return diff, amax


def style_mse_fwd(G, T, loss_out):
    """loss_out: 1-element fp32 view that receives mean((G-T)^2); returns (diff, max|diff| as a 1-element tensor)."""
    B, C, _ = G.shape
    diff = torch.empty_like(G)
    amax = torch.empty(1, dtype=torch.float32, device=G.device)
    ws = workspace(4096 * 4, G.device, 'small')
    _lib.check(_lib.lib().st_style_mse_fwd(G.data_ptr(), T.data_ptr(), B, C, diff.data_ptr(), loss_out.data_ptr(), amax.data_ptr(),
                                           ws.data_ptr(), _stream()), 'st_style_mse_fwd')
    return diff, amax


def style_seed(diff, amax, HW, gw, fmt):
    """Packed per-image weights D = gw * 4/(B C^2 C HW) * diff for the dF = F*D conv, in the plane format ``fmt``
    of F (fp16: under a power-of-two scale slot).  Returns (buffers to keep alive, w_ptrs of ops.conv2d)."""
    B, C, _ = diff.shape
    planes = bwd_planes()
    buf = torch.empty((planes, B * C, C), dtype=torch.bfloat16, device=diff.device)
    hi, lo_p = buf[0].data_ptr(), (buf[1].data_ptr() if planes > 1 else None)
    slot = torch.empty(2, dtype=torch.float32, device=diff.device) if fmt else None
    _lib.check(_lib.lib().st_style_mse_bwd_seed(diff.data_ptr(), B, C, HW, gw.data_ptr(), hi, lo_p, planes | fmt, amax.data_ptr(),
                                                _p(slot), _stream()), 'st_style_mse_bwd_seed')
    return (buf, slot), (hi, lo_p, B * C, 1, C, planes | fmt, slot)


def style_seed_multi(nodes):
    """``style_seed`` for every style node of a loss network in ONE launch.  nodes: [(diff, amax, HW, gw, fmt)];
    returns [(buffers to keep alive, w_ptrs of ops.conv2d)] in the same order."""
    n = len(nodes)
    planes = bwd_planes()
    out, rows = [], []
    for diff, amax, HW, gw, fmt in nodes:
        B, C, _ = diff.shape
        buf = torch.empty((planes, B * C, C), dtype=torch.bfloat16, device=diff.device)
        hi, lo_p = buf[0].data_ptr(), (buf[1].data_ptr() if planes > 1 else None)
        slot = torch.empty(2, dtype=torch.float32, device=diff.device) if fmt else None
        out.append(((buf, slot), (hi, lo_p, B * C, 1, C, planes | fmt, slot)))
        rows.append((diff.data_ptr(), B, C, HW, gw.data_ptr(), hi, lo_p or 0, planes | fmt, amax.data_ptr(), _p(slot) or 0))
    VP, IA = ctypes.c_void_p * n, ctypes.c_int * n
    col = lambda j, T: T(*[r[j] for r in rows])
    _lib.check(_lib.lib().st_style_mse_bwd_seed_multi(n, col(0, VP), col(1, IA), col(2, IA), col(3, IA), col(4, VP), col(5, VP), col(6, VP),
                                                      col(7, IA), col(8, VP), col(9, VP), _stream()), 'st_style_mse_bwd_seed_multi')
    return out


def loss_total(losses, weights, extras=()):
    """sum(losses * weights) + sum(extra * scale) as a 0-dim device tensor, one launch.  extras: up to two
    (0-dim tensor, float scale) pairs."""
    out = torch.empty((), dtype=torch.float32, device=losses.device)
    ex = list(extras) + [(None, 0.0)] * (2 - len(extras))
    _lib.check(_lib.lib().st_loss_total(losses.data_ptr(), weights.data_ptr(), losses.numel(), _p(ex[0][0]), float(ex[0][1]),
                                        _p(ex[1][0]), float(ex[1][1]), out.data_ptr(), _stream()), 'st_loss_total')
    return out


def absmax_slot(t, target_log2=10.0):
    """{s, 1/s} with the power of two s that maps max|t| to 2^target_log2 (device tensor, no host sync)."""
    _require_cuda(t, 'absmax_slot')
    t = t.contiguous().float()
    slot = torch.empty(2, dtype=torch.float32, device=t.device)
    ws = workspace(4096 * 4, t.device, 'small')
    _lib.check(_lib.lib().st_absmax_slot(t.data_ptr(), t.numel(), float(target_log2), slot.data_ptr(), ws.data_ptr(), _stream()),
               'st_absmax_slot')
    return slot


def mse_fwd(x, t, loss_out, relu=False, feature_form=False):
    ws = workspace(4096 * 4, x.buf.device, 'small')
    n = x.buf[0].numel()
    _lib.check(_lib.lib().st_mse_fwd(x.hi, x.lo, x.read_code(2), t.hi, t.lo, t.read_code(2), n, int(relu), int(feature_form),
                                     loss_out.data_ptr(), ws.data_ptr(), _stream()), 'st_mse_fwd')


def mse_bwd(x, t, gw, scale, out=None, relu=False, mul=None):
    acc = out is not None
    out = out or x.grad_like()
    _bf16_grad(out)
    n = x.buf[0].numel()
    _lib.check(_lib.lib().st_mse_bwd(x.hi, x.lo, x.read_code(2), t.hi, t.lo, t.read_code(2), n, int(relu), float(scale), gw.data_ptr(),
                                     _p(mul), int(acc), out.hi, out.lo, _stream()), 'st_mse_bwd')
    return out


def instnorm_fwd(x, gamma, beta, relu=False, up=False, eps=1e-5):
    L = _lib.lib()
    dev = x.buf.device
    ws = workspace(L.st_instnorm_workspace_bytes(x.N, x.C), dev)
    y = Split(x.N, x.H << int(up), x.W << int(up), x.C, dev, planes=x.planes, fmt=x.fmt)
    stats = torch.empty((2, x.N, x.C), dtype=torch.float32, device=dev)
    _lib.check(L.st_instnorm_fwd(x.hi, x.lo, gamma.data_ptr(), beta.data_ptr(), eps, y.hi, y.lo, x.code,
                                 stats[0].data_ptr(), stats[1].data_ptr(), x.N, x.H, x.W, x.C, int(relu), int(up),
                                 ws.data_ptr(), _stream()), 'st_instnorm_fwd')
    return y, stats


def instnorm_bwd(x, gy, stats, gamma, beta, relu=False, up=False, want_dsum=False, dgamma_out=None, dbeta_out=None,
                 dsum_out=None):
    """Returns (dx, dgamma, dbeta[, dsum]); dsum = per-channel sum of dx = the bias gradient of the
    convolution in front of the norm.  The saved input is read at the gradient precision.  dx feeds the
    backward-data / backward-weight products of that convolution: ``Split.dy_like`` (scaled fp16 planes in the
    f16x3 mode)."""
    L = _lib.lib()
    _bf16_grad(gy)
    dev = x.buf.device
    ws = workspace(L.st_instnorm_workspace_bytes(x.N, x.C), dev)
    dx = x.dy_like()
    dgamma = _grad_out(dgamma_out, gamma.shape, dev)
    dbeta = _grad_out(dbeta_out, beta.shape, dev)
    dsum = _grad_out(dsum_out, (x.C,), dev) if want_dsum else None
    _lib.check(L.st_instnorm_bwd(x.hi, x.lo, gy.hi, gy.lo, stats[0].data_ptr(), stats[1].data_ptr(),
                                 gamma.data_ptr(), beta.data_ptr(), dx.hi, dx.lo, dx.code, dx.slot_ptr, dgamma.data_ptr(), dbeta.data_ptr(), _p(dsum),
                                 x.read_code(bwd_planes()), x.N, x.H, x.W, x.C, int(relu), int(up), ws.data_ptr(), _stream()),
               'st_instnorm_bwd')
    return (dx, dgamma, dbeta, dsum) if want_dsum else (dx, dgamma, dbeta)


def tv_fwd(x, factor):
    _require_cuda(x, 'tv_fwd')
    N, C, H, W = x.shape
    loss = torch.empty((), dtype=torch.float32, device=x.device)
    ws = workspace(4096 * 4, x.device, 'small')
    _lib.check(_lib.lib().st_tv_fwd(x.data_ptr(), N, C, H, W, float(factor), loss.data_ptr(), ws.data_ptr(), _stream()), 'st_tv_fwd')
    return loss


def tv_bwd(x, factor, gw):
    N, C, H, W = x.shape
    g = torch.empty_like(x)
    _lib.check(_lib.lib().st_tv_bwd(x.data_ptr(), N, C, H, W, float(factor), gw.data_ptr(), g.data_ptr(), _stream()), 'st_tv_bwd')
    return g


def temporal_fwd(old_content, old_stylized, cur_content, cur_stylized, tw):
    _require_cuda(cur_stylized, 'temporal_fwd')
    out3 = torch.empty((3,), dtype=torch.float32, device=cur_stylized.device)
    ws = workspace(4096 * 4, cur_stylized.device, 'small')
    _lib.check(_lib.lib().st_temporal_fwd(old_content.data_ptr(), old_stylized.data_ptr(), cur_content.data_ptr(),
                                          cur_stylized.data_ptr(), cur_stylized.numel(), float(tw), out3.data_ptr(),
                                          ws.data_ptr(), _stream()), 'st_temporal_fwd')
    return out3


def temporal_warp_fwd(old_content, old_stylized, cur_content, cur_stylized, flow, tw):
    """Extension: temporal loss against the flow-warped previous frame."""
    _require_cuda(cur_stylized, 'temporal_warp_fwd')
    N, C, H, W = cur_stylized.shape
    out3 = torch.empty((3,), dtype=torch.float32, device=cur_stylized.device)
    ws = workspace(4096 * 4, cur_stylized.device, 'small')
    _lib.check(_lib.lib().st_temporal_warp_fwd(old_content.data_ptr(), old_stylized.data_ptr(), cur_content.data_ptr(),
                                               cur_stylized.data_ptr(), flow.data_ptr(), N, C, H, W, float(tw), out3.data_ptr(),
                                               ws.data_ptr(), _stream()), 'st_temporal_warp_fwd')
    return out3


def temporal_warp_bwd(old_stylized, cur_stylized, flow, tw, stats3, gw):
    N, C, H, W = cur_stylized.shape
    g = torch.empty_like(cur_stylized)
    _lib.check(_lib.lib().st_temporal_warp_bwd(old_stylized.data_ptr(), cur_stylized.data_ptr(), flow.data_ptr(), N, C, H, W,
                                               float(tw), stats3.data_ptr(), gw.data_ptr(), g.data_ptr(), _stream()),
               'st_temporal_warp_bwd')
    return g


def temporal_bwd(old_stylized, cur_stylized, tw, stats3, gw):
    g = torch.empty_like(cur_stylized)
    _lib.check(_lib.lib().st_temporal_bwd(old_stylized.data_ptr(), cur_stylized.data_ptr(), cur_stylized.numel(), float(tw),
                                          stats3.data_ptr(), gw.data_ptr(), g.data_ptr(), _stream()), 'st_temporal_bwd')
    return g


def _stats3():
    import ctypes
    from . import constants
    mean = (ctypes.c_float * 3)(*constants.IMAGENET_MEAN)
    std = (ctypes.c_float * 3)(*constants.IMAGENET_STD)
    return mean, std


def image_u8_to_nchw(u8):
    """[N, H, W, 3] uint8 (device) -> ImageNet-normalised [N, 3, H, W] fp32 (reference img_utils.py:33-42)."""
    _require_cuda(u8, 'image_u8_to_nchw')
    assert u8.dtype == torch.uint8 and u8.dim() == 4 and u8.shape[3] == 3
    u8 = u8.contiguous()
    N, H, W, _ = u8.shape
    out = torch.empty((N, 3, H, W), dtype=torch.float32, device=u8.device)
    mean, std = _stats3()
    _lib.check(_lib.lib().st_image_u8_to_nchw(u8.data_ptr(), out.data_ptr(), N, H, W, mean, std, _stream()), 'st_image_u8_to_nchw')
    return out


def image_nchw_to_u8(x, out=None):
    """[N, 3, H, W] normalised fp32 -> [N, H, W, 3] uint8 picture (de-normalise, clamp: reference img_utils.py:105-109)."""
    _require_cuda(x, 'image_nchw_to_u8')
    x = x.contiguous().float()
    N, C, H, W = x.shape
    assert C == 3
    if out is None:
        out = torch.empty((N, H, W, 3), dtype=torch.uint8, device=x.device)
    mean, std = _stats3()
    _lib.check(_lib.lib().st_image_nchw_to_u8(x.data_ptr(), out.data_ptr(), N, H, W, mean, std, _stream()), 'st_image_nchw_to_u8')
    return out
