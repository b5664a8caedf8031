"""Per-kernel parity: each C-ABI entry point against a plain PyTorch fp32 op on the
same GPU (TF32 off), in both parity-grade precision modes (f16x3: two fp16 planes / 3 tensor-core passes;
bf16x6: three bf16 planes / 6 passes).  Tolerances: gradients carry ~16 mantissa bits, so
rel-L2 <= 1e-4 on single ops (north star: features/Grams <= 1e-2)."""
import pytest
import torch
import torch.nn.functional as F

from tests.util import fp32_reference_mode, nhwc, rel_l2, seeded

pytestmark = pytest.mark.gpu
TOL = 1e-4


@pytest.fixture(scope='module', params=['f16x3', 'bf16x6'])
def ops(request):
    fp32_reference_mode()
    from styletransfer_b200 import _lib, ops as o
    _lib.lib()
    prev = o.PRECISION
    o.set_precision(request.param)
    yield o
    o.set_precision(prev)


def _grad(ops, gy, cpad=None):
    """A gradient tensor as the product makes them where they meet forward activations in a tensor-core product
    (fp16 planes in the f16x3 mode -- unscaled here: the test values fit the fp16 range)."""
    return ops.from_nchw(gy, cpad=cpad, planes=ops.bwd_planes(), fmt=ops.dy_fmt())


def _bf16_grad(ops, gy):
    """A gradient tensor of the bf16 chain (conv backward-data outputs, pooling / loss gradients)."""
    return ops.from_nchw(gy, planes=ops.bwd_planes(), fmt=0)


def test_layout_roundtrip(ops):
    x = seeded((2, 3, 20, 36), 1, device='cuda')
    s = ops.from_nchw(x)
    assert s.shape == (2, 20, 36, 8)
    assert rel_l2(nhwc(s, 3), x) < 1e-5
    assert float(nhwc(s)[:, 3:].abs().max()) == 0.0
    assert rel_l2(ops.to_nchw(s, 3), x) < 1e-5
    a, b = seeded((1, 3, 8, 8), 2, device='cuda'), seeded((1, 3, 8, 8), 3, device='cuda')
    s2 = ops.from_nchw(a, b)                      # channel concat (video net input)
    assert rel_l2(nhwc(s2, 6), torch.cat([a, b], 1)) < 1e-5


def _conv_case(ops, N, H, W, cin, cout, ks, stride, path, relu=True, up=0, seed=0):
    from styletransfer_b200 import _lib
    _lib.lib().st_set_conv_path(path)
    try:
        x = seeded((N, cin, H, W), seed + 1, device='cuda')
        w = seeded((cout, cin, ks, ks), seed + 2, 1.0 / (cin * ks * ks) ** 0.5, device='cuda')
        b = seeded((cout,), seed + 3, 0.1, device='cuda')
        xs = ops.from_nchw(x)
        pw = ops.PackedWeight(w, 0, cols_pad=xs.C)
        out, rl, _ = ops.conv2d(xs, pw, b, stride=stride, up=up, want_relu=relu)
        xin = F.interpolate(x, scale_factor=2, mode='nearest') if up else x
        ref = F.conv2d(xin, w, b, stride=stride, padding=ks // 2)
        e = rel_l2(nhwc(out, cout), ref)
        er = rel_l2(nhwc(rl, cout), F.relu(ref)) if relu else 0.0
        return max(e, er)
    finally:
        _lib.lib().st_set_conv_path(0)


@pytest.mark.parametrize('cin,cout,ks,stride,up', [
    (3, 64, 3, 1, 0), (3, 32, 9, 1, 0), (6, 32, 9, 1, 0), (32, 64, 3, 2, 0), (64, 128, 3, 2, 0),
    (32, 3, 9, 1, 0), (128, 64, 3, 1, 1), (64, 32, 3, 1, 1), (64, 64, 3, 1, 0)])
def test_conv_generic(ops, cin, cout, ks, stride, up):
    assert _conv_case(ops, 2, 20, 28, cin, cout, ks, stride, path=1, up=up) < TOL


@pytest.mark.parametrize('cin,cout,H,W', [
    (64, 64, 16, 32), (64, 64, 20, 36), (64, 128, 24, 40), (128, 128, 8, 16),
    (128, 256, 12, 20), (128, 64, 16, 16), (64, 32, 16, 48), (128, 128, 33, 17)])
def test_conv_tcgen05_3x3(ops, cin, cout, H, W):
    assert _conv_case(ops, 2, H, W, cin, cout, 3, 1, path=0) < TOL


def test_conv_tcgen05_many_tiles(ops):
    # more work items than SMs: exercises the persistent loop, both TMEM accumulators
    # and every pipeline phase wrap
    assert _conv_case(ops, 3, 96, 160, 64, 64, 3, 1, path=0) < TOL
    assert _conv_case(ops, 1, 64, 128, 128, 256, 3, 1, path=0) < TOL


def test_conv_tcgen05_plain_bf16(ops):
    """the cheaper precision modes: 1 plane / 1 pass and 2 planes / 3 passes"""
    prev = ops.PRECISION
    try:
        ops.set_precision('bf16')
        assert _conv_case(ops, 1, 16, 32, 64, 64, 3, 1, path=0) < 1e-2
        assert _conv_case(ops, 1, 16, 32, 32, 64, 3, 2, path=1) < 1e-2
        ops.set_precision('bf16x3')
        assert _conv_case(ops, 2, 24, 40, 128, 128, 3, 1, path=0) < 1e-4
        assert _conv_case(ops, 1, 16, 32, 3, 32, 9, 1, path=1) < 1e-4
    finally:
        ops.set_precision(prev)


def test_forward_conv_is_fp32_accurate(ops):
    """Parity-grade forward (3 bf16 planes / 6 passes = 24 mantissa bits, or 2 fp16 planes / 3 passes = 22 bits):
    what remains is the tensor core's truncating fp32 accumulation over the K loop (~1e-5 at K = 1152,
    DESIGN.md "Precision").  Compared against an fp64 convolution."""
    x = seeded((1, 128, 24, 40), 91, device='cuda')
    w = seeded((128, 128, 3, 3), 92, 0.03, device='cuda')
    ref = F.conv2d(x.double(), w.double(), None, padding=1)
    out, _, _ = ops.conv2d(ops.from_nchw(x), ops.PackedWeight(w, 0))
    assert rel_l2(nhwc(out), ref) < 2e-5
    x8 = seeded((1, 3, 24, 40), 93, device='cuda')
    w8 = seeded((64, 3, 3, 3), 94, 0.2, device='cuda')
    ref = F.conv2d(x8.double(), w8.double(), None, padding=1)
    out, _, _ = ops.conv2d(ops.from_nchw(x8), ops.PackedWeight(w8, 0, cols_pad=8))
    assert rel_l2(nhwc(out, 64), ref) < 2e-5


@pytest.mark.parametrize('N,H,W', [(2, 20, 36), (1, 7, 9), (1, 64, 64), (3, 150, 77)])
def test_conv_tcgen05_image_gradient(ops, N, H, W):
    """backward-data of conv1_1: the project-then-gather tcgen05 kernel (conv_imgrad_tc.cu, kernel id 4) and the
    16-channel tile of conv_tc it replaced (st_set_conv_path(2) selects it), fp32 NCHW output"""
    from styletransfer_b200 import _lib
    x = seeded((N, 3, H, W), 71, device='cuda').requires_grad_(True)
    w = seeded((64, 3, 3, 3), 72, 0.2, device='cuda')
    gy = seeded((N, 64, H, W), 74, device='cuda')
    F.conv2d(x, w, None, padding=1).backward(gy)
    pw = ops.PackedWeight(w, 2, rows_pad=16)
    _, _, gx = ops.conv2d(_bf16_grad(ops, gy), pw, want_out=False, out_nchw_channels=3)
    assert _lib.lib().st_last_conv_kernel() == 4
    assert rel_l2(gx, x.grad) < TOL
    _lib.lib().st_set_conv_path(2)
    try:
        _, _, gx16 = ops.conv2d(_bf16_grad(ops, gy), pw, want_out=False, out_nchw_channels=3)
        assert _lib.lib().st_last_conv_kernel() == 1
    finally:
        _lib.lib().st_set_conv_path(0)
    assert rel_l2(gx16, x.grad) < TOL and rel_l2(gx, gx16) < 1e-5


@pytest.mark.parametrize('cin,cout,H,W', [(32, 64, 20, 28), (64, 128, 16, 32), (32, 64, 33, 47), (64, 128, 9, 19)])
def test_conv_tcx_stride2_forward(ops, cin, cout, H, W):
    """3x3 stride-2 forward on the run-time-geometry tcgen05 kernel (strided TMA boxes)"""
    from styletransfer_b200 import _lib
    assert _conv_case(ops, 2, H, W, cin, cout, 3, 2, path=0) < TOL
    assert _lib.lib().st_last_conv_kernel() == 2


@pytest.mark.parametrize('cin,cout,H,W', [(32, 64, 16, 32), (64, 128, 24, 40), (64, 128, 34, 18)])
def test_conv_tcx_stride2_dgrad(ops, cin, cout, H, W):
    """backward-data of the stride-2 convs as 4 output-parity classes"""
    from styletransfer_b200 import _lib
    N = 2
    w = seeded((cout, cin, 3, 3), 5, 1.0 / (cin * 9) ** 0.5, device='cuda')
    gy = seeded((N, cout, H // 2, W // 2), 6, device='cuda')
    x = seeded((N, cin, H, W), 7, device='cuda').requires_grad_(True)
    F.conv2d(x, w, None, stride=2, padding=1).backward(gy)
    gys = _grad(ops, gy)
    pw = ops.PackedWeight(w, 1, rows_pad=cin, cols_pad=gys.C, fmt=gys.fmt)
    out, _, _ = ops.conv2d(gys, pw, stride=2, transposed=1, out_hw=(H, W))
    assert _lib.lib().st_last_conv_kernel() == 2
    assert rel_l2(nhwc(out, cin), x.grad) < TOL


@pytest.mark.parametrize('cin,cout,H,W', [(32, 64, 16, 32), (32, 64, 21, 37), (32, 32, 8, 16), (96, 64, 12, 20)])
def test_conv_tcx_32_channel_chunks(ops, cin, cout, H, W):
    """stride-1 3x3 with a K side of 32-channel chunks (64-byte swizzle), forward and masked dgrad"""
    from styletransfer_b200 import _lib
    assert _conv_case(ops, 2, H, W, cin, cout, 3, 1, path=0) < TOL
    assert _lib.lib().st_last_conv_kernel() == 2
    N = 2
    w = seeded((cin, cout, 3, 3), 5, 1.0 / (cout * 9) ** 0.5, device='cuda')     # forward layer cout -> cin
    gy = seeded((N, cin, H, W), 6, device='cuda')
    x = seeded((N, cout, H, W), 7, device='cuda').requires_grad_(True)
    F.conv2d(x, w, None, padding=1).backward(gy)
    gys = _grad(ops, gy)
    pw = ops.PackedWeight(w, 2, rows_pad=cout, cols_pad=gys.C, fmt=gys.fmt)
    tap, add = seeded((N, cout, H, W), 8, device='cuda'), seeded((N, cout, H, W), 9, device='cuda')
    out, _, _ = ops.conv2d(gys, pw, mask=ops.from_nchw(tap), add=_grad(ops, add))
    assert _lib.lib().st_last_conv_kernel() == 2
    assert rel_l2(nhwc(out, cout), x.grad * (tap > 0) + add) < TOL


@pytest.mark.parametrize('cin,cout,ks,stride,path', [
    (64, 64, 3, 1, 0), (128, 256, 3, 1, 0), (64, 128, 3, 1, 0), (3, 64, 3, 1, 1),
    (32, 64, 3, 2, 1), (64, 128, 3, 2, 1), (32, 3, 9, 1, 1), (3, 32, 9, 1, 1)])
def test_conv_dgrad(ops, cin, cout, ks, stride, path):
    """backward-data with the fused epilogue: gx = dgrad(gy) * (tap > 0) + add"""
    N, H, W = 2, 16, 32
    Ho, Wo = (H + 2 * (ks // 2) - ks) // stride + 1, (W + 2 * (ks // 2) - ks) // stride + 1
    w = seeded((cout, cin, ks, ks), 5, 1.0 / (cin * ks * ks) ** 0.5, device='cuda')
    gy = seeded((N, cout, Ho, Wo), 6, device='cuda')
    x = seeded((N, cin, H, W), 7, device='cuda').requires_grad_(True)
    F.conv2d(x, w, None, stride=stride, padding=ks // 2).backward(gy)
    gys = _grad(ops, gy)
    cin_pad = ops.pad8(cin)
    if path == 0:
        pw = ops.PackedWeight(w, 2, rows_pad=cin_pad, cols_pad=gys.C, fmt=gys.fmt)
        tap = seeded((N, cin, H, W), 8, device='cuda')
        add = seeded((N, cin, H, W), 9, device='cuda')
        # the addend of a backward-data epilogue is a bf16 gradient, or a (scaled) fp16 one (residual blocks)
        for a in (_bf16_grad(ops, add), _grad(ops, add)):
            out, _, _ = ops.conv2d(gys, pw, mask=ops.from_nchw(tap), add=a)
            ref = x.grad * (tap > 0) + add
            assert rel_l2(nhwc(out, cin), ref) < TOL
    else:
        pw = ops.PackedWeight(w, 1, rows_pad=cin_pad, cols_pad=gys.C, fmt=gys.fmt)
        out, _, nchw = ops.conv2d(gys, pw, stride=stride, transposed=1, out_hw=(H, W), out_nchw_channels=cin)
        assert rel_l2(nhwc(out, cin), x.grad) < TOL
        assert rel_l2(nchw, x.grad) < TOL


@pytest.mark.parametrize('cin,cout,ks,stride,up', [
    (128, 128, 3, 1, 0), (3, 32, 9, 1, 0), (32, 64, 3, 2, 0), (128, 64, 3, 1, 1), (32, 3, 9, 1, 0),
    (128, 64, 3, 1, 0), (64, 32, 3, 1, 0), (64, 64, 3, 1, 0)])
def test_conv_wgrad(ops, cin, cout, ks, stride, up):
    N, H, W = 2, 12, 20
    _wgrad_case(ops, N, H, W, cin, cout, ks, stride, up)


def test_conv_wgrad_tcgen05_many_tiles(ops):
    """more pixel tiles than splits (several pipeline wraps per CTA) and a ragged image"""
    _wgrad_case(ops, 3, 64, 96, 128, 128, 3, 1, 0)
    _wgrad_case(ops, 2, 37, 53, 64, 32, 3, 1, 0)
    _wgrad_case(ops, 1, 128, 128, 128, 64, 3, 1, 0)


def _wgrad_case(ops, N, H, W, cin, cout, ks, stride, up):
    x = seeded((N, cin, H, W), 11, device='cuda')
    w = seeded((cout, cin, ks, ks), 12, 0.05, device='cuda').requires_grad_(True)
    b = torch.zeros(cout, device='cuda', requires_grad=True)
    xin = F.interpolate(x, scale_factor=2, mode='nearest') if up else x
    y = F.conv2d(xin, w, b, stride=stride, padding=ks // 2)
    gy = seeded(tuple(y.shape), 13, device='cuda')
    y.backward(gy)
    dw, db = ops.conv2d_wgrad(_grad(ops, gy), ops.from_nchw(x), ks, stride, ks // 2, up, cout, cin)
    assert rel_l2(dw, w.grad) < TOL
    assert rel_l2(db, b.grad) < TOL
    # a bf16 gradient against (possibly fp16) activations: the converting mma.sync kernel takes mixed formats
    dw, db = ops.conv2d_wgrad(_bf16_grad(ops, gy), ops.from_nchw(x), ks, stride, ks // 2, up, cout, cin)
    assert rel_l2(dw, w.grad) < TOL and rel_l2(db, b.grad) < TOL


@pytest.mark.parametrize('H,W', [(16, 32), (15, 21)])
def test_relu_maxpool(ops, H, W):
    x = seeded((2, 64, H, W), 21, device='cuda').requires_grad_(True)
    y = F.max_pool2d(F.relu(x), 2, 2)
    gy = seeded(tuple(y.shape), 22, device='cuda')
    y.backward(gy)
    xs = ops.from_nchw(x.detach())
    ys = ops.relu_maxpool_fwd(xs)
    # the pooled value is rounded to split-bf16 twice (input and output)
    assert rel_l2(nhwc(ys), y) < 1e-5
    add = seeded((2, 64, H, W), 23, device='cuda')
    gx = ops.relu_maxpool_bwd(xs, _bf16_grad(ops, gy), add=_bf16_grad(ops, add))
    assert rel_l2(nhwc(gx), x.grad + add) < 1e-5
    # the variant whose forward saves the routing decision (1 byte per pooled element) and whose backward does not
    # read x again: bit-identical outputs, odd borders included; ties (relu'(x) = 0 windows) route nothing
    ys2, idx = ops.relu_maxpool_fwd(xs, want_idx=True)
    assert torch.equal(ys2.buf, ys.buf) and idx.shape == (2, H // 2, W // 2, 64) and int(idx.max()) <= 4
    for a in (_bf16_grad(ops, add), None):
        g_ref = ops.relu_maxpool_bwd(xs, _bf16_grad(ops, gy), add=a)
        g_idx = ops.relu_maxpool_bwd(xs, _bf16_grad(ops, gy), add=a, idx=idx)
        assert torch.equal(g_idx.buf, g_ref.buf)
    neg = ops.from_nchw(-x.detach().abs())
    _, idx_neg = ops.relu_maxpool_fwd(neg, want_idx=True)
    assert int(idx_neg.min()) == 4


@pytest.mark.parametrize('B,C,H,W', [(1, 64, 32, 32), (2, 128, 20, 12), (3, 256, 9, 7), (1, 64, 128, 128), (16, 256, 32, 32), (16, 64, 64, 64)])
def test_gram_and_style_loss(ops, B, C, H, W):
    f = seeded((B, C, H, W), 31, device='cuda').requires_grad_(True)
    T = seeded((1, C, C), 32, 0.01, device='cuda')
    T = 0.5 * (T + T.transpose(1, 2))            # Gram targets are symmetric (the backward relies on it)
    G = torch.bmm(f.view(B, C, -1), f.view(B, C, -1).transpose(1, 2)) / (C * H * W)
    loss = F.mse_loss(G, T.expand_as(G))
    loss.backward()
    fs = ops.from_nchw(f.detach())
    Gk = ops.gram(fs)
    assert rel_l2(Gk, G) < TOL
    lk = torch.zeros(1, device='cuda')
    diff, amax = ops.style_mse_fwd(Gk, T, lk)
    assert abs(float(lk) - float(loss)) / float(loss) < 1e-4
    assert abs(float(amax) - float((G - T).abs().max())) <= 1e-4 * float(amax)
    # loss weights of very different size: the per-tensor power-of-two scale of the fp16 seed follows them
    for weight in (1.0, 1e5, 1e-6):
        gw = torch.tensor([weight], device='cuda')
        _keep, wp = ops.style_seed(diff, amax, H * W, gw, fs.fmt)
        dF, _, _ = ops.conv2d(fs, None, w_ptrs=wp, w_img_rows=C)
        assert rel_l2(nhwc(dF), weight * f.grad) < 2e-4


@pytest.mark.parametrize('B,C,H,W,masked', [(1, 64, 32, 48, False), (2, 128, 21, 19, False), (1, 128, 40, 24, True),
                                            (2, 256, 17, 32, True), (1, 64, 130, 70, False)])
def test_gram_backward_with_routed_pool_gradient(ops, B, C, H, W, masked):
    """st_conv2d_routed = st_relu_maxpool2x2_bwd_idx + st_conv2d(add=...) in one launch: dF = F * D [* relu'(tap)] +
    MaxPool2d/ReLU backward of the pooled gradient (torch autograd on F.max_pool2d(F.relu(x), 2) is the reference),
    odd borders included."""
    tap = seeded((B, C, H, W), 71, device='cuda').requires_grad_(True)
    y = F.max_pool2d(F.relu(tap), 2)
    gy = seeded(tuple(y.shape), 72, device='cuda')
    y.backward(gy)
    routed_ref = tap.grad.clone()
    taps = ops.from_nchw(tap.detach())
    _, idx = ops.relu_maxpool_fwd(taps, want_idx=True)
    # the operand of the Gram backward: the tap itself ('Conv2d_k' style node) or its ReLU copy ('ReLU_k' node, masked)
    f = ops.from_nchw(F.relu(tap.detach())) if masked else taps
    fn = F.relu(tap.detach()) if masked else tap.detach()
    diff = seeded((B, C, C), 73, 1e-2, device='cuda')
    diff = 0.5 * (diff + diff.transpose(1, 2))
    amax = diff.abs().max().reshape(1)
    gw = torch.tensor([2.0], device='cuda')
    _keep, wp = ops.style_seed(diff, amax, H * W, gw, f.fmt)
    gys = _bf16_grad(ops, gy)
    out = ops.conv2d_routed(f, wp, C, gys, idx, mask=taps if masked else None)
    assert out is not None, 'the routed launch must take the tensor-core path at these shapes'
    # two-launch form of the product, same operands
    two = ops.relu_maxpool_bwd(taps, gys, add=ops.conv2d(f, None, w_ptrs=wp, w_img_rows=C, mask=taps if masked else None)[0], idx=idx)
    assert rel_l2(nhwc(out), nhwc(two)) < 1e-5
    # fp32 reference: d/dF of sum(diff * Gram) scaled as StyleLoss does (4 / (C C B) * gw / (C H W)) -- through the two-launch
    # form's own pieces this is checked in test_gram_and_style_loss; here the routed part against autograd
    dF = ops.conv2d(f, None, w_ptrs=wp, w_img_rows=C, mask=taps if masked else None)[0]
    assert rel_l2(nhwc(out) - nhwc(dF), routed_ref) < 1e-4
    # the generic path has no routed form: the wrapper reports it and the caller falls back
    from styletransfer_b200 import _lib
    _lib.lib().st_set_conv_path(1)
    try:
        assert ops.conv2d_routed(f, wp, C, gys, idx) is None
    finally:
        _lib.lib().st_set_conv_path(0)


@pytest.mark.parametrize('N,Cin,Cout,H,W', [(1, 64, 64, 32, 48), (2, 64, 128, 21, 19), (1, 128, 128, 40, 24), (1, 64, 64, 130, 70)])
def test_conv_with_fused_relu_maxpool(ops, N, Cin, Cout, H, W):
    """st_conv2d_pool = st_conv2d + st_relu_maxpool2x2_fwd_idx in one launch (nn.Conv2d -> nn.ReLU -> nn.MaxPool2d of the
    VGG pieces): the tap is bit-identical and the pooled values are identical to the two-launch form (rounding to the planes
    is monotone), the routing bytes agree except where two window values round to the same stored number."""
    x = seeded((N, Cin, H, W), 81, device='cuda')
    w = seeded((Cout, Cin, 3, 3), 82, 0.05, device='cuda')
    b = seeded((Cout,), 83, 0.1, device='cuda')
    xs = ops.from_nchw(F.relu(x))
    pw = ops.PackedWeight(w, 0, cols_pad=Cin)
    got = ops.conv2d_pool(xs, pw, b)
    assert got is not None, 'these shapes take the tensor-core kernel'
    tap, pooled, idx = got
    tap2, _, _ = ops.conv2d(xs, pw, b)
    pooled2, idx2 = ops.relu_maxpool_fwd(tap2, want_idx=True)
    assert torch.equal(tap.buf, tap2.buf)
    # same VALUES (hi + lo); the split of a value into planes is not unique when lo is exactly half an ulp of hi
    assert torch.equal(nhwc(pooled), nhwc(pooled2))
    assert idx.shape == idx2.shape and float((idx != idx2).float().mean()) < 1e-4
    ref = F.max_pool2d(F.relu(F.conv2d(F.relu(x), w, b, padding=1)), 2)
    assert rel_l2(nhwc(pooled), ref) < 1e-4
    # the routing bytes point at a window element whose stored value is the pooled value (4: nothing positive)
    t = nhwc(tap2)[:, :, :2 * (H // 2), :2 * (W // 2)]
    win = torch.stack([t[:, :, 0::2, 0::2], t[:, :, 0::2, 1::2], t[:, :, 1::2, 0::2], t[:, :, 1::2, 1::2]], 0)
    sel = idx.permute(0, 3, 1, 2).long()
    picked = torch.gather(torch.cat([win, torch.zeros_like(win[:1])], 0), 0, sel.unsqueeze(0)).squeeze(0)
    assert torch.equal(picked.clamp(min=0), nhwc(pooled))


def test_mse(ops):
    x = seeded((2, 128, 10, 12), 41, device='cuda').requires_grad_(True)
    t = seeded((2, 128, 10, 12), 42, device='cuda')
    loss = F.mse_loss(x, t)
    loss.backward()
    xs, ts = ops.from_nchw(x.detach()), ops.from_nchw(t)
    lk = torch.zeros(1, device='cuda')
    ops.mse_fwd(xs, ts, lk)
    assert abs(float(lk) - float(loss)) / float(loss) < 1e-5
    g = ops.mse_bwd(xs, ts, torch.tensor([3.0], device='cuda'), 2.0 / x.numel())
    assert rel_l2(nhwc(g), 3.0 * x.grad) < 1e-5
    # FeatureReconstructionLoss form on relu'd maps (network.py:186-201)
    x2 = x.detach().clone().requires_grad_(True)
    l2 = F.mse_loss(F.relu(x2), F.relu(t)).pow(2) / x2.numel()
    l2.backward()
    ops.mse_fwd(xs, ts, lk, relu=True, feature_form=True)
    assert abs(float(lk) - float(l2)) / float(l2) < 1e-4
    mse = torch.zeros(1, device='cuda')
    ops.mse_fwd(xs, ts, mse, relu=True)
    n = float(x.numel())
    g2 = ops.mse_bwd(xs, ts, torch.tensor([1.0], device='cuda'), 4.0 / (n * n), relu=True, mul=mse)
    assert rel_l2(nhwc(g2), x2.grad) < 1e-4


@pytest.mark.parametrize('C,H,W,relu,up', [(32, 16, 24, True, False), (128, 8, 8, False, True),
                                           (64, 12, 20, True, True), (128, 7, 9, False, False)])
def test_instance_norm(ops, C, H, W, relu, up):
    N = 3
    x = seeded((N, C, H, W), 51, device='cuda').requires_grad_(True)
    gamma = (1 + 0.1 * seeded((C,), 52, device='cuda')).requires_grad_(True)
    beta = (0.1 * seeded((C,), 53, device='cuda')).requires_grad_(True)
    y = F.instance_norm(x, None, None, gamma, beta, True, 0.0, 1e-5)
    if relu:
        y = F.relu(y)
    if up:
        y = F.interpolate(y, scale_factor=2, mode='nearest')
    gy = seeded(tuple(y.shape), 54, device='cuda')
    y.backward(gy)
    xs = ops.from_nchw(x.detach())
    ys, stats = ops.instnorm_fwd(xs, gamma.detach(), beta.detach(), relu=relu, up=up)
    assert rel_l2(nhwc(ys), y) < 1e-4
    # gradients of very different size: dx is stored under a per-tensor power-of-two scale in the f16x3 mode
    for weight in (1.0, 3e-7, 2e4):
        dx, dg, db = ops.instnorm_bwd(xs, _bf16_grad(ops, weight * gy), stats, gamma.detach(), beta.detach(), relu=relu, up=up)
        assert (dx.slot is not None) == bool(ops.dy_fmt())
        assert rel_l2(nhwc(dx), weight * x.grad) < 2e-4
        assert rel_l2(dg, weight * gamma.grad) < 2e-4
        assert rel_l2(db, weight * beta.grad) < 2e-4
        assert rel_l2(ops.to_nchw(dx), weight * x.grad) < 2e-4


def test_tv_and_temporal(ops):
    x = seeded((2, 3, 17, 23), 61, device='cuda').requires_grad_(True)
    tv = 1e-6 * ((x[:, :, :, :-1] - x[:, :, :, 1:]).abs().sum() + (x[:, :, :-1, :] - x[:, :, 1:, :]).abs().sum())
    tv.backward()
    lk = ops.tv_fwd(x.detach(), 1e-6)
    assert abs(float(lk) - float(tv)) / float(tv) < 1e-5
    g = ops.tv_bwd(x.detach(), 1e-6, torch.tensor([2.0], device='cuda'))
    assert rel_l2(g, 2.0 * x.grad) < 1e-6
    oc, os_, cc = (seeded((2, 3, 17, 23), s, device='cuda') for s in (62, 63, 64))
    cs = seeded((2, 3, 17, 23), 65, device='cuda').requires_grad_(True)
    tl = ((cs - os_).norm() / ((cc - oc).norm() + 1)) * 0.8
    tl.backward()
    out3 = ops.temporal_fwd(oc, os_, cc, cs.detach(), 0.8)
    assert abs(float(out3[0]) - float(tl)) / float(tl) < 1e-5
    g = ops.temporal_bwd(os_, cs.detach(), 0.8, out3, torch.tensor([1.0], device='cuda'))
    assert rel_l2(g, cs.grad) < 1e-5


def test_temporal_flow_warp_extension(ops):
    """EXTENSION, parity unpinned (the reference's temporal loss has no optical flow): the flow-warped
    temporal loss against torch's grid_sample (bilinear, zeros, align_corners=True, flow in pixels), including
    samples that leave the frame; zero flow reproduces the reference form; the gradient reaches cur_stylized only."""
    N, C, H, W = 2, 3, 19, 27
    oc, os_, cc = (seeded((N, C, H, W), s, device='cuda') for s in (62, 63, 64))
    cs = seeded((N, C, H, W), 65, device='cuda').requires_grad_(True)
    flow = seeded((N, 2, H, W), 66, 2.5, device='cuda')
    ys, xs = torch.meshgrid(torch.arange(H, device='cuda', dtype=torch.float32),
                            torch.arange(W, device='cuda', dtype=torch.float32), indexing='ij')
    gx = 2 * (xs[None] + flow[:, 0]) / (W - 1) - 1
    gy = 2 * (ys[None] + flow[:, 1]) / (H - 1) - 1
    grid = torch.stack([gx, gy], dim=-1)
    warp = lambda t: F.grid_sample(t, grid, mode='bilinear', padding_mode='zeros', align_corners=True)
    tl = ((cs - warp(os_)).norm() / ((cc - warp(oc)).norm() + 1)) * 0.8
    tl.backward()
    out3 = ops.temporal_warp_fwd(oc, os_, cc, cs.detach(), flow, 0.8)
    assert abs(float(out3[0]) - float(tl)) / float(tl) < 1e-5
    g = ops.temporal_warp_bwd(os_, cs.detach(), flow, 0.8, out3, torch.tensor([1.0], device='cuda'))
    assert rel_l2(g, cs.grad) < 1e-5
    zero = torch.zeros_like(flow)
    assert torch.allclose(ops.temporal_warp_fwd(oc, os_, cc, cs.detach(), zero, 0.8), ops.temporal_fwd(oc, os_, cc, cs.detach(), 0.8), rtol=1e-6)
    from styletransfer_b200 import network
    net = network.VideoTransformNet.__new__(network.VideoTransformNet)
    cs2 = cs.detach().clone().requires_grad_(True)
    loss = network.VideoTransformNet.get_temporal_loss(net, oc, os_, cc, cs2, temporal_weight=0.8, flow=flow)
    loss.backward()
    assert abs(float(loss) - float(tl)) / float(tl) < 1e-5 and rel_l2(cs2.grad, cs.grad) < 1e-5


@pytest.mark.parametrize('N,H,W', [(2, 20, 36), (1, 7, 9), (1, 64, 64), (16, 200, 203)])
def test_conv_image_fwd_and_dgrad(ops, N, H, W):
    """VGG conv1_1 (3 -> 64) in exact fp32 on the CUDA cores, NCHW fp32 image in, image gradient out.  The last case has
    enough tiles for the forward kernel with 8 pixels per thread (conv_image.cu picks by tile count), ragged in both axes."""
    x = seeded((N, 3, H, W), 71, device='cuda').requires_grad_(True)
    w = seeded((64, 3, 3, 3), 72, 0.2, device='cuda')
    b = seeded((64,), 73, 0.1, device='cuda')
    ref = F.conv2d(x, w, b, padding=1)
    gy = seeded(tuple(ref.shape), 74, device='cuda')
    ref.backward(gy)
    tap, relu = ops.conv_image_fwd(x.detach(), w, b, want_out=True, want_relu=True)
    assert rel_l2(nhwc(tap), ref) < 2e-6
    assert rel_l2(nhwc(relu), F.relu(ref)) < 2e-6
    only_relu = ops.conv_image_fwd(x.detach(), w, b, want_out=False, want_relu=True)
    assert only_relu[0] is None and rel_l2(nhwc(only_relu[1]), F.relu(ref)) < 2e-6
    gx = ops.conv_image_dgrad(_bf16_grad(ops, gy), w)
    assert gx.shape == x.shape and rel_l2(gx, x.grad) < 2e-5


@pytest.mark.parametrize('path', [0, 1])
@pytest.mark.parametrize('C,N,H,W', [(3, 2, 20, 36), (6, 1, 13, 45), (3, 1, 64, 160)])
def test_conv9_layers(ops, C, N, H, W, path):
    """The 9x9 layers on the tensor cores (path 0: hand-built shifted-window operands, vertical-stage
    reduce) and as exact-fp32 direct convolutions (path 1): layer 0 (C -> 32) forward + backward-weight,
    layer 22 (32 -> 3) forward + backward-data + backward-weight, against torch fp32."""
    from styletransfer_b200 import _lib
    _lib.lib().st_set_conv_path(path)
    try:
        _conv9_layers(ops, C, N, H, W)
    finally:
        _lib.lib().st_set_conv_path(0)


def _conv9_layers(ops, C, N, H, W):
    # layer 0
    x = seeded((N, C, H, W), 81, device='cuda')
    w0 = seeded((32, C, 9, 9), 82, 0.06, device='cuda').requires_grad_(True)
    b0 = seeded((32,), 83, 0.1, device='cuda').requires_grad_(True)
    y = F.conv2d(x, w0, b0, padding=4)
    gy = seeded(tuple(y.shape), 84, device='cuda')
    y.backward(gy)
    out = ops.conv9_expand(x, w0.detach(), b0.detach())
    assert rel_l2(nhwc(out), y) < 2e-6
    dw, db = ops.conv9_wgrad(_grad(ops, gy), x, 0)
    assert rel_l2(dw, w0.grad) < 2e-5 and rel_l2(db, b0.grad) < 2e-5
    if C != 3:
        return
    # layer 22
    xb = seeded((N, 32, H, W), 85, device='cuda').requires_grad_(True)
    w2 = seeded((3, 32, 9, 9), 86, 0.02, device='cuda').requires_grad_(True)
    b2 = seeded((3,), 87, 0.1, device='cuda').requires_grad_(True)
    y2 = F.conv2d(xb, w2, b2, padding=4)
    g2 = seeded(tuple(y2.shape), 88, device='cuda')
    y2.backward(g2)
    xs = ops.from_nchw(xb.detach())
    assert rel_l2(ops.conv9_reduce(xs, w2.detach(), b2.detach()), y2) < 2e-6
    dx = ops.conv9_expand(g2, w2.detach(), flip=True, planes=2, fmt=0)
    assert rel_l2(nhwc(dx), xb.grad) < 2e-5
    # a tiny gradient image: it enters the fp16 operands of the f16x3 mode under the power-of-two scale of its maximum
    for weight in (1.0, 1e-7):
        slot = ops.absmax_slot(weight * g2) if ops.dy_fmt() else None
        dw2, db2 = ops.conv9_wgrad(xs, weight * g2, 1, small_slot=slot)
        assert rel_l2(dw2, weight * w2.grad) < 2e-5 and rel_l2(db2, weight * b2.grad) < 2e-6
