"""st_instnorm_fwd_static (include/stransfer_b200.h): the InstanceNorm2d forward with its statistics partials in a
static device array -- for callers that cannot allocate -- against st_instnorm_fwd with a caller workspace, through
the C ABI, on two calls in a row that share the array (different shapes, so the second call finds the first call's
partials in it).  Same kernels, same launch geometry: the results are compared bit for bit.

Collected last on purpose (zz)."""
import pytest
import torch

from tests.util import record, seeded

pytestmark = pytest.mark.gpu

MAX_NC = 1024          # ST_INSTNORM_STATIC_MAX_NC


def _static_fwd(ops, x, gamma, beta, relu, up, eps=1e-5):
    from styletransfer_b200 import _lib
    dev = x.buf.device
    y = ops.Split(x.N, x.H << int(up), x.W << int(up), x.C, dev, planes=x.planes, fmt=x.fmt)
    stats = torch.empty((2, x.N, x.C), dtype=torch.float32, device=dev)
    rc = _lib.lib().st_instnorm_fwd_static(x.hi, x.lo, gamma.data_ptr(), beta.data_ptr(), eps, y.hi, y.lo, x.code,
                                           stats[0].data_ptr(), stats[1].data_ptr(), x.N, x.H, x.W, x.C, int(relu), int(up),
                                           ops._stream())
    return rc, y, stats


# (N, H, W, C, relu, up): the transform net's InstanceNorm shapes (network.py:531-600) at small frame sizes;
# every N*C is within the limit, the last one exactly on it
CASES = [(2, 48, 40, 32, True, False), (3, 17, 23, 128, False, False), (4, 24, 24, 64, True, True), (8, 12, 20, 128, True, False)]


@pytest.mark.parametrize('precision', ['f16x3', 'bf16x6'])
def test_static_scratch_instnorm_matches_workspace_entry_point_on_consecutive_calls(precision):
    from styletransfer_b200 import _lib, ops
    prev = ops.PRECISION
    ops.set_precision(precision)
    try:
        dev = torch.device('cuda', torch.cuda.current_device())
        for rounds in range(2):                        # the whole sequence twice: no call depends on what the array held
            for i, (N, H, W, C, relu, up) in enumerate(CASES):
                assert N * C <= MAX_NC
                x = ops.Split(N, H, W, C, dev).set_from_float(seeded((N, H, W, C), 40 + i, scale=3.0, device=dev) + 0.5)
                gamma, beta = seeded((C,), 50 + i, device=dev) + 1.0, seeded((C,), 60 + i, device=dev)
                y_ref, stats_ref = ops.instnorm_fwd(x, gamma, beta, relu=relu, up=up)
                torch.cuda.synchronize()
                rc, y, stats = _static_fwd(ops, x, gamma, beta, relu, up)
                _lib.check(rc, 'st_instnorm_fwd_static')
                torch.cuda.synchronize()
                assert torch.equal(stats, stats_ref), f'case {i}: mean / rstd differ'
                assert torch.equal(y.buf.view(torch.int16), y_ref.buf.view(torch.int16)), f'case {i}: planes differ'
                record(f'{precision}/instnorm_static/case{i}_round{rounds}_bit_identical', 1.0)
    finally:
        ops.set_precision(prev)


def test_static_scratch_instnorm_rejects_sizes_beyond_the_limit():
    from styletransfer_b200 import _lib, ops
    dev = torch.device('cuda', torch.cuda.current_device())
    N, C = 9, 128                                       # 1152 > 1024
    x = ops.Split(N, 8, 8, C, dev, zero=True)
    gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    rc, _, _ = _static_fwd(ops, x, gamma, beta, False, False)
    assert rc != 0
    msg = _lib.lib().st_last_error().decode()
    assert 'exceeds the built-in scratch' in msg and str(MAX_NC) in msg, msg
    with pytest.raises(_lib.StransferError):
        _lib.check(rc, 'st_instnorm_fwd_static')
