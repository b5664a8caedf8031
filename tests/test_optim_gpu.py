"""Device-resident optimizer steps (csrc/optim.cu) against torch's own optimizers -- the ones
the reference calls (optim.Adam network.py:643-649/:765, optim.LBFGS network.py:435/:456) --
and against the CPU model oracle/lbfgs_model.py."""
import numpy as np
import pytest
import torch

from oracle.lbfgs_model import LBFGSModel
from tests.util import record, rel_l2

pytestmark = pytest.mark.gpu


def _problem(n, seed, spread=0.0, device='cuda'):
    gen = torch.Generator().manual_seed(seed)
    A = (torch.randn(n, n, generator=gen) / n ** 0.5 * torch.logspace(0, spread, n)).to(device)
    b = (torch.randn(n, generator=gen) * 3).to(device)
    x0 = torch.randn(n, generator=gen).to(device)

    def f(x):
        r = A @ x - b
        return torch.sqrt(1 + r * r).sum() + 0.05 * (x ** 4).sum()
    return f, x0


def test_fused_adam_matches_torch_adam():
    from styletransfer_b200.optim import FusedAdam
    gen = torch.Generator().manual_seed(5)
    shapes = [(32, 3, 9, 9), (32,), (1,), (2049,), (128, 128, 3, 3), (7, 5)]
    ref = [torch.nn.Parameter(torch.randn(*s, generator=gen).cuda()) for s in shapes]
    ours = [torch.nn.Parameter(p.detach().clone()) for p in ref]
    o_ref, o_ours = torch.optim.Adam(ref), FusedAdam(ours)
    for step in range(7):
        for i, (a, b) in enumerate(zip(ref, ours)):
            frozen = (i == 3 and step < 2) or (i == 1 and step == 4)      # a tensor without a gradient is skipped
            g = None if frozen else (torch.randn(*a.shape, generator=gen) * 10.0 ** (step % 3 - 1)).cuda()
            a.grad = g
            b.grad = None if g is None else g.clone()                      # not a view of the bucket: copied in
        o_ref.step()
        o_ours.step()
    torch.cuda.synchronize()
    worst = max(float((a - b).abs().max() / a.abs().max()) for a, b in zip(ref, ours))
    assert record('fused_adam/param_rel_max', worst) < 2e-6
    assert o_ours.steps.tolist() == [7, 6, 7, 5, 7, 7]
    assert all(b._version > 0 for b in ours)                               # packed-weight caches see the update


def test_fused_adam_reads_the_bucket_in_place():
    from styletransfer_b200.optim import FusedAdam
    ps = [torch.nn.Parameter(torch.ones(5000, device='cuda')), torch.nn.Parameter(torch.ones(3, device='cuda'))]
    opt = FusedAdam(ps, lr=0.1)
    assert ps[0].grad.data_ptr() == opt.bucket.flat.data_ptr()
    assert ps[0].data.data_ptr() == opt.flat_param.data_ptr()
    opt.bucket.flat.fill_(2.0)
    opt.step()
    assert torch.allclose(ps[0], torch.full_like(ps[0], 0.9), atol=1e-6) and torch.allclose(ps[1], torch.full_like(ps[1], 0.9), atol=1e-6)


@pytest.mark.parametrize('n,hist,steps', [(1003, 7, 2), (4096, 100, 3), (5, 3, 1)])
def test_device_lbfgs_matches_torch_and_model(n, hist, steps):
    """Same closure, same start: torch.optim.LBFGS (GPU fp32), the CPU model and the device kernels.
    Without a line search the iteration amplifies rounding noise, so the comparison with torch is
    made after the first step (20 iterations); the model (same arithmetic) is followed throughout."""
    from styletransfer_b200.optim import DeviceLBFGS
    f, x0 = _problem(n, 11)
    xt = x0.clone().requires_grad_(True)
    ot = torch.optim.LBFGS([xt], history_size=hist)
    xd = x0.clone().requires_grad_(True)
    od = DeviceLBFGS([xd], history_size=hist)
    mod = LBFGSModel(x0.cpu().numpy().copy(), history_size=hist)

    def make(x, opt):
        def closure():
            opt.zero_grad()
            loss = f(x)
            loss.backward()
            return loss
        return closure

    def model_closure():
        x = torch.from_numpy(mod.x).cuda().requires_grad_(True)
        loss = f(x)
        loss.backward()
        return float(loss.detach()), x.grad.cpu().numpy().copy()
    errs_t, errs_m = [], []
    for s in range(steps):
        ot.step(make(xt, ot))
        od.step(make(xd, od))
        mod.step(model_closure)
        errs_t.append(rel_l2(xd, xt))
        errs_m.append(rel_l2(xd, torch.from_numpy(mod.x)))
    info = od.info()
    record(f'lbfgs/n{n}_h{hist}/vs_torch_step1', errs_t[0])
    record(f'lbfgs/n{n}_h{hist}/vs_model_last', errs_m[-1])
    assert errs_t[0] < 2e-4, errs_t
    assert errs_m[0] < 2e-4, errs_m
    assert int(info['n_iter']) == mod.n_iter and int(info['k']) == mod.k and int(info['head']) == mod.head, (info, mod.n_iter, mod.k)
    assert float(f(xd.detach())) <= float(f(x0)) and torch.isfinite(xd).all()


def test_device_lbfgs_stops_like_torch():
    """A quadratic bowl is minimised in a few iterations: torch leaves the loop on its tolerances;
    the device flag must freeze x at the same point and later steps must stay there."""
    from styletransfer_b200.optim import DeviceLBFGS
    target = torch.linspace(-1, 1, 50, device='cuda')

    def run(make_opt):
        x = torch.zeros(50, device='cuda', requires_grad=True)
        opt = make_opt([x])

        def closure():
            opt.zero_grad()
            loss = ((x - target) ** 2).sum()
            loss.backward()
            return loss
        for _ in range(3):
            opt.step(closure)
        return x.detach().clone(), opt
    xt, _ = run(torch.optim.LBFGS)
    xd, od = run(DeviceLBFGS)
    info = od.info()
    assert info['halt'] == 1.0 and int(info['n_iter']) < 60
    assert float((xd - target).abs().max()) < 1e-5 and float((xd - xt).abs().max()) < 1e-5


def test_train_gatys_uses_the_device_lbfgs():
    from styletransfer_b200 import network
    from styletransfer_b200.optim import DeviceLBFGS
    x = torch.zeros(1, 3, 8, 8, device='cuda')
    net = network.StyleNetwork.__new__(network.StyleNetwork)
    assert isinstance(network.StyleNetwork.get_content_optimizer(net, x, optt=torch.optim.LBFGS), DeviceLBFGS)
    assert isinstance(network.StyleNetwork.get_content_optimizer(net, x, optt=torch.optim.Adam), torch.optim.Adam)
