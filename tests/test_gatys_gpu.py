"""Parity of the Gatys hot path (StyleNetwork forward+backward to the image) against
the committed golden vectors (outputs of the UNMODIFIED reference, see
oracle/make_golden.py) and against the CPU oracle on fresh seeded inputs.

North-star tolerances: per-layer features and Gram matrices <= 1e-2
relative L2; total loss and input gradient <= 1e-3 relative; PSNR >= 40 dB after
100 L-BFGS steps."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import oracle
from tests.util import nhwc, record, rel_l2, relu_decision_flips, seeded

pytestmark = pytest.mark.gpu
TOL_FEATURE, TOL_LOSS_GRAD = 1e-2, 1e-3
# input-gradient bound of the ten-conv loss networks when ReLU decisions of near-zero pre-activations differ from the
# oracle's (tests.util.relu_decision_flips): each flipped element of an 8x12 ... 16x24 feature map moves a whole receptive
# field of the gradient.  With no flipped decision the bound is TOL_LOSS_GRAD.
TOL_GRAD_WITH_FLIPS, FLIP_MARGIN = 5e-3, 1e-4


def _deep_grad_tol(tag, x0_cpu, vw_cpu, n_convs):
    flips, worst = relu_decision_flips(x0_cpu, vw_cpu, n_convs)
    record(tag + '/relu_decision_flips', flips)
    record(tag + '/flipped_value_over_layer_max', worst)
    assert worst < FLIP_MARGIN, 'a ReLU decision differs on an element that is NOT near zero: a forward error, not rounding'
    return TOL_LOSS_GRAD if flips == 0 else TOL_GRAD_WITH_FLIPS


@pytest.fixture(scope='module', params=['f16x3', 'bf16x6'])
def net_mod(request):
    """Both parity-grade precision modes (DESIGN.md "Precision")."""
    from styletransfer_b200 import network, ops
    prev = ops.PRECISION
    ops.set_precision(request.param)
    yield network
    ops.set_precision(prev)


def _vgg_cuda():
    return [(w.cuda(), b.cuda()) for w, b in oracle.make_vgg_weights(1)]


def _closure(network, content, style, x0, sw, cw):
    net = network.StyleNetwork(style, content, vgg_weights=_vgg_cuda())
    x = x0.clone().requires_grad_(True)
    net(x, content)
    total = net.get_total_current_style_loss(weight=sw) + net.get_total_current_content_loss(weight=cw)
    total.backward()
    per_node = torch.stack([n[0].loss for n in net.style_losses + net.content_losses + net.feature_losses])
    return net, total.detach(), x.grad.detach(), per_node.detach()


@pytest.mark.parametrize('name', ['gatys_64x64', 'gatys_40x56'])
def test_closure_matches_reference_golden(net_mod, golden_dir, name):
    g = np.load(os.path.join(golden_dir, name + '.npz'))
    h, w = int(g['h']), int(g['w'])
    content = seeded((1, 3, h, w), 10, device='cuda')
    style = seeded((1, 3, h, w), 11, device='cuda')
    x0 = content + 0.3 * seeded((1, 3, h, w), 12, device='cuda')
    net, total, grad, per_node = _closure(net_mod, content, style, x0, float(g['sw']), float(g['cw']))
    assert abs(float(total) - float(g['total'])) / abs(float(g['total'])) < TOL_LOSS_GRAD
    assert rel_l2(per_node, torch.from_numpy(g['losses'])) < TOL_LOSS_GRAD
    assert record(name + '/grad_rel_l2', rel_l2(grad, torch.from_numpy(g['grad']))) < TOL_LOSS_GRAD
    record(name + '/total_rel', abs(float(total) - float(g['total'])) / abs(float(g['total'])))
    assert rel_l2(net.style_losses[0][0].target, torch.from_numpy(g['style_target_gram0'])) < TOL_FEATURE


def test_features_and_grams_match_oracle(net_mod):
    from styletransfer_b200 import ops, vgg_loss
    vw = oracle.make_vgg_weights(1)
    x = seeded((2, 3, 48, 80), 70)
    taps_ref = oracle.vgg_taps(x, vw)
    prefix = vgg_loss.VGGPrefix(_vgg_cuda())
    taps = prefix.taps(x.cuda())
    for t, r in zip(taps, taps_ref):
        assert record('tap_rel_l2', rel_l2(nhwc(t), r)) < TOL_FEATURE
        assert record('gram_rel_l2', rel_l2(ops.gram(t), oracle.gram_matrix(r))) < TOL_FEATURE


def test_run_through_pieces_and_modules(net_mod):
    vw = oracle.make_vgg_weights(1)
    style, content = seeded((1, 3, 32, 32), 71), seeded((1, 3, 32, 32), 72)
    net = net_mod.StyleNetwork(style.cuda(), content.cuda(), vgg_weights=_vgg_cuda())
    taps = oracle.vgg_taps(content, vw)
    for piece, ref in [(0, taps[0]), (1, taps[1]), (2, taps[2]), (3, taps[3]), (4, torch.relu(taps[3])), (5, taps[4])]:
        assert rel_l2(net.run_through_pieces(content.cuda(), piece), ref) < TOL_FEATURE
    # all pieces (until = -1): with five conv weights the last layer is ReLU_5
    assert rel_l2(net.run_through_pieces(content.cuda()), torch.relu(taps[4])) < TOL_FEATURE
    # stand-alone loss modules with autograd (network.py:79-201)
    f = seeded((2, 64, 12, 20), 73)
    tgt = seeded((1, 64, 12, 20), 74)
    fr = f.clone().requires_grad_(True)
    lref = oracle.style_loss(fr, oracle.gram_matrix(tgt))
    lref.backward()
    sl = net_mod.StyleLoss(tgt.cuda())
    fc = f.cuda().requires_grad_(True)
    assert sl(fc) is fc
    sl.loss.backward()
    assert abs(float(sl.loss) - float(lref)) / float(lref) < TOL_LOSS_GRAD
    assert rel_l2(fc.grad, fr.grad) < TOL_LOSS_GRAD
    for cls, fn in [(net_mod.ContentLoss, oracle.content_loss), (net_mod.FeatureReconstructionLoss, oracle.feature_loss)]:
        fr = f.clone().requires_grad_(True)
        tgt2 = seeded((2, 64, 12, 20), 75)
        lref = fn(fr, tgt2)
        lref.backward()
        m = cls(tgt2.cuda())
        fc = f.cuda().requires_grad_(True)
        m(fc)
        m.loss.backward()
        assert abs(float(m.loss) - float(lref)) / float(lref) < TOL_LOSS_GRAD
        assert rel_l2(fc.grad, fr.grad) < TOL_LOSS_GRAD


def test_feature_loss_gradient_path(net_mod):
    """static_test uses the feature node (network.py:788-791); check its value + gradient."""
    vw = oracle.make_vgg_weights(1)
    content, style = seeded((1, 3, 32, 48), 76), seeded((1, 3, 32, 48), 77)
    x0 = seeded((1, 3, 32, 48), 78)
    xr = x0.clone().requires_grad_(True)
    losses, _ = oracle.loss_network(xr, content, oracle.style_targets(style, vw), vw)
    (1e5 * losses[:5].sum() + 1e9 * losses[6]).backward()
    net = net_mod.StyleNetwork(style.cuda(), content.cuda(), vgg_weights=_vgg_cuda())
    x = x0.cuda().requires_grad_(True)
    net(x, content.cuda())
    tot = net.get_total_current_style_loss(1e5) + net.get_total_current_feature_loss(1e9)
    tot.backward()
    assert rel_l2(net.feature_losses[0][0].loss, losses[6]) < TOL_LOSS_GRAD
    assert record('feature_path/grad_rel_l2', rel_l2(x.grad, xr.grad)) < TOL_LOSS_GRAD


def test_lbfgs_100_steps_psnr(net_mod):
    """100 L-BFGS steps (train_gatys, network.py:411-458) vs the same loop on the CPU oracle."""
    vw = oracle.make_vgg_weights(1)
    s = 32
    content, style = seeded((1, 3, s, s), 80), seeded((1, 3, s, s), 81)
    targets = oracle.style_targets(style, vw)
    xr = content.clone().requires_grad_(True)
    opt = torch.optim.LBFGS([xr])
    for _ in range(100):
        def closure():
            opt.zero_grad()
            losses, _ = oracle.loss_network(xr, content, targets, vw)
            total = 1e5 * losses[:5].sum() + losses[5]
            total.backward()
            return total
        opt.step(closure)
    net = net_mod.StyleNetwork(style.cuda(), content.cuda(), vgg_weights=_vgg_cuda())
    out = net.train_gatys(style.cuda(), content.cuda(), steps=100)
    mse = float(((out.detach().cpu() - xr.detach()) ** 2).mean())
    peak = float(xr.detach().max() - xr.detach().min())
    psnr = 10 * math.log10(peak * peak / max(mse, 1e-30))
    record('lbfgs100/psnr_db', psnr)
    assert psnr >= 40.0, psnr


def test_full_size_properties(net_mod):
    """512x512 (BASELINE configs[1]): size-independent properties -- the loss of the
    content image against itself has zero content term, the gradient is finite, and
    two evaluations are bit-identical (deterministic reductions)."""
    content, style = seeded((1, 3, 512, 512), 90, device='cuda'), seeded((1, 3, 512, 512), 91, device='cuda')
    net, total, grad, per_node = _closure(net_mod, content, style, content, 1e5, 1.0)
    assert float(per_node[5]) == 0.0
    assert torch.isfinite(grad).all() and float(grad.abs().max()) > 0
    _, total2, grad2, _ = _closure(net_mod, content, style, content, 1e5, 1.0)
    assert float(total) == float(total2) and torch.equal(grad, grad2)
    # linearity of backward in the loss weights: grad(2*sw, 2*cw) == 2*grad(sw, cw)
    x0 = content + 0.1 * style
    _, _, g1, _ = _closure(net_mod, content, style, x0, 1e5, 1.0)
    _, _, g2, _ = _closure(net_mod, content, style, x0, 2e5, 2.0)
    assert rel_l2(g2, 2 * g1) < 1e-4


def test_cuda_graph_closure_matches_eager(net_mod):
    """train_gatys replays the closure as a CUDA graph: same loss and gradient as the eager
    closure, and a replay sees in-place updates of the image (what L-BFGS does between calls)."""
    h, w = 48, 80
    content, style = seeded((1, 3, h, w), 81, device='cuda'), seeded((1, 3, h, w), 82, device='cuda')
    net = net_mod.StyleNetwork(style, content, vgg_weights=_vgg_cuda())
    net.compute_feature_loss = False
    x = (content + 0.2 * seeded((1, 3, h, w), 83, device='cuda')).requires_grad_(True)
    eager = net.gatys_closure(x, content, 1e5, 1, graph=False)
    l0 = eager().detach().clone()
    g0 = x.grad.detach().clone()
    graphed = net.gatys_closure(x, content, 1e5, 1, graph=True)
    assert graphed.launches > 20
    l1 = graphed().detach().clone()
    assert torch.equal(l0, l1) and torch.equal(g0, x.grad)
    with torch.no_grad():
        x.add_(0.05 * seeded((1, 3, h, w), 84, device='cuda'))
    l2, g2 = graphed().detach().clone(), x.grad.detach().clone()
    l3 = eager().detach().clone()
    assert not torch.equal(l2, l1)
    assert torch.equal(l2, l3) and torch.equal(g2, x.grad)
    # everything the captured closure reads must stay alive: small allocations between replays
    # must not be able to land on memory the graph has baked in
    for k in range(4):
        junk = [torch.full((7,), float(k + i), device='cuda') for i in range(64)]
        with torch.no_grad():
            x.add_(0.05 * seeded((1, 3, h, w), 85 + k, device='cuda'))
        lg, gg = graphed().detach().clone(), x.grad.detach().clone()
        le = eager().detach().clone()
        assert torch.equal(lg, le) and torch.equal(gg, x.grad), k
        del junk
    out = net.train_gatys(style, content, steps=2)          # L-BFGS through the graphed closure
    assert out.shape == content.shape and torch.isfinite(out).all()


def _generic_net(net_mod):
    from oracle.make_golden import GENERIC_LAYERS

    class Net(net_mod.StyleNetwork):
        style_layers, content_layers, feature_loss_layers = GENERIC_LAYERS['style'], GENERIC_LAYERS['content'], GENERIC_LAYERS['feature']
    return Net


def test_non_default_loss_layers_match_reference_golden(net_mod, golden_dir):
    """The loss layers are class attributes a user may edit (network.py:214-232) and the piece builder is generic over
    them (network.py:257-314): style nodes on conv / ReLU / max-pool layers up to conv 9 (256 -> 256, 256 -> 512
    channels), content nodes on ReLU_7 and Conv2d_10 (512 -> 512), a feature node on ReLU_10 -- against the
    fixture written by the UNMODIFIED reference with the same edited lists (oracle/make_golden.py generic_case)."""
    from oracle.make_golden import GENERIC_CONVS
    g = np.load(os.path.join(golden_dir, 'gatys_generic_64x96.npz'))
    h, w = int(g['h']), int(g['w'])
    vw = [(a.cuda(), b.cuda()) for a, b in oracle.make_vgg_weights(1, GENERIC_CONVS)]
    content, style = seeded((1, 3, h, w), 10, device='cuda'), seeded((1, 3, h, w), 11, device='cuda')
    x0 = content + 0.3 * seeded((1, 3, h, w), 12, device='cuda')
    net = _generic_net(net_mod)(style, content, vgg_weights=vw)
    # same pieces as the reference cuts (the last piece holds what follows the last loss layer of the available blocks)
    assert [p for _, p in net.style_losses] == g['piece_of_style'].tolist()
    assert [p for _, p in net.content_losses] == g['piece_of_content'].tolist()
    assert [p for _, p in net.feature_losses] == g['piece_of_feature'].tolist()
    x = x0.clone().requires_grad_(True)
    net(x, content)
    total = net.get_total_current_style_loss(weight=float(g['sw'])) + net.get_total_current_content_loss(weight=float(g['cw']))
    total.backward()
    per_node = torch.stack([n[0].loss for n in net.style_losses + net.content_losses + net.feature_losses]).detach()
    assert record('generic_golden/total_rel', abs(float(total) - float(g['total'])) / abs(float(g['total']))) < TOL_LOSS_GRAD
    assert record('generic_golden/losses_rel_l2', rel_l2(per_node, torch.from_numpy(g['losses']))) < TOL_LOSS_GRAD
    tol = _deep_grad_tol('generic_golden', x0.cpu(), oracle.make_vgg_weights(1, GENERIC_CONVS), GENERIC_CONVS)
    assert record('generic_golden/grad_rel_l2', rel_l2(x.grad, torch.from_numpy(g['grad']))) < tol
    # the feature node takes part in the gradient when asked for (static_test's combination)
    xf = x0.clone().requires_grad_(True)
    net(xf, content)
    (net.get_total_current_style_loss(1e5) + net.get_total_current_feature_loss(1e12)).backward()
    o = oracle.make_vgg_weights(1, GENERIC_CONVS)
    from oracle.make_golden import GENERIC_LAYERS as GL
    xr = x0.cpu().clone().requires_grad_(True)
    outs = oracle.vgg_layer_outputs(xr, o, set(GL['style']) | set(GL['feature']))
    with torch.no_grad():
        s_out = oracle.vgg_layer_outputs(style.cpu(), o, GL['style'])
        c_out = oracle.vgg_layer_outputs(content.cpu(), o, GL['feature'])
    ref = 1e5 * sum(oracle.style_loss(outs[n], oracle.gram_matrix(s_out[n])) for n in GL['style']) + \
        1e12 * sum(oracle.feature_loss(outs[n], c_out[n]) for n in GL['feature'])
    ref.backward()
    assert record('generic_feature/grad_rel_l2', rel_l2(xf.grad, xr.grad)) < tol


def test_non_default_loss_layers_match_oracle_and_all_pieces(net_mod):
    """The same edited layer lists at a larger, ragged size against the CPU oracle; run_through_pieces through
    every piece (until = -1, all 16 convs and 5 pools) and up to single pieces; the graph-replayed closure."""
    from oracle.make_golden import GENERIC_LAYERS as GL
    h, w = 136, 200
    vw16 = oracle.make_vgg_weights(1, 16)
    content, style = seeded((1, 3, h, w), 20), seeded((1, 3, h, w), 21)
    x0 = content + 0.3 * seeded((1, 3, h, w), 22)
    o_total, o_grad, o_losses = oracle.generic_closure(x0, content, style, vw16, GL['style'], GL['content'], GL['feature'])
    net = _generic_net(net_mod)(style.cuda(), content.cuda(), vgg_weights=[(a.cuda(), b.cuda()) for a, b in vw16])
    x = x0.cuda().requires_grad_(True)
    net(x, content.cuda())
    total = net.get_total_current_style_loss(weight=1e5) + net.get_total_current_content_loss(weight=1)
    total.backward()
    per_node = torch.stack([n[0].loss for n in net.style_losses + net.content_losses + net.feature_losses]).detach()
    assert record('generic_136x200/total_rel', abs(float(total) - float(o_total)) / abs(float(o_total))) < TOL_LOSS_GRAD
    assert record('generic_136x200/losses_rel_l2', rel_l2(per_node, o_losses)) < TOL_LOSS_GRAD
    tol = _deep_grad_tol('generic_136x200', x0, vw16, 10)
    assert record('generic_136x200/grad_rel_l2', rel_l2(x.grad, o_grad)) < tol
    # pieces: all of them (MaxPool2d_16), the piece of a max-pool node, of a ReLU node and of a deep conv node
    names = ['MaxPool2d_16', 'MaxPool2d_4', 'ReLU_7', 'Conv2d_10']
    outs = oracle.vgg_layer_outputs(content, vw16, names)
    assert len(net.net_pieces) == 10 and sum(len(list(p.children())) for p in net.net_pieces) == 37
    assert record('generic/all_pieces_rel_l2', rel_l2(net.run_through_pieces(content.cuda()), outs['MaxPool2d_16'])) < TOL_FEATURE
    for (node, piece), name in ((net.style_losses[3], 'MaxPool2d_4'), (net.content_losses[0], 'ReLU_7'), (net.content_losses[1], 'Conv2d_10')):
        assert rel_l2(net.run_through_pieces(content.cuda(), piece), outs[name]) < TOL_FEATURE
    assert rel_l2(net.content_losses[0][0].target, outs['ReLU_7']) < TOL_FEATURE
    # the closure train_gatys replays
    xg = x0.cuda().requires_grad_(True)
    net.compute_feature_loss = False
    closure = net.gatys_closure(xg, content.cuda(), 1e5, 1)
    loss = closure()
    assert abs(float(loss) - float(o_total)) / abs(float(o_total)) < TOL_LOSS_GRAD and rel_l2(xg.grad, o_grad) < tol
