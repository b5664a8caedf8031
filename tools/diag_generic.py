#!/usr/bin/env python
"""Diagnostic: the non-default loss-layer case at 64x96 -- per-layer forward error against the oracle, ReLU decisions that
differ, and where the gradient error sits (a flipped decision is a localised patch; a kernel bug follows tile edges)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def seeded(shape, seed):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed))


def main():
    from oracle import oracle
    from oracle.make_golden import GENERIC_LAYERS as GL, GENERIC_CONVS
    from styletransfer_b200 import network, ops, vgg_loss
    ops.set_precision(sys.argv[1] if len(sys.argv) > 1 else 'f16x3')
    h, w = (64, 96) if len(sys.argv) < 3 else (int(sys.argv[2]), int(sys.argv[3]))
    vw = oracle.make_vgg_weights(1, GENERIC_CONVS)
    content, style = seeded((1, 3, h, w), 10), seeded((1, 3, h, w), 11)
    x0 = content + 0.3 * seeded((1, 3, h, w), 12)
    names = [f'Conv2d_{k}' for k in range(1, GENERIC_CONVS + 1)]
    ref = oracle.vgg_layer_outputs(x0, vw, names)
    prefix = vgg_loss.VGGPrefix([(a.cuda(), b.cuda()) for a, b in vw])
    _, taps = prefix.run(x0.cuda(), {('C', k) for k in range(1, GENERIC_CONVS + 1)})
    for k in range(1, GENERIC_CONVS + 1):
        got = ops.to_nchw(taps[k]).cpu()
        r = ref[f'Conv2d_{k}']
        err = (got - r).abs()
        flips = int(((got > 0) != (r > 0)).sum())
        iy = err.amax(dim=(0, 1))
        yy, xx = divmod(int(iy.argmax()), iy.shape[1])
        print(f'conv {k:2d} {tuple(r.shape)}: rel L2 {float((got - r).norm() / r.norm()):.3e}  max abs {float(err.max()):.3e} at (y {yy}, x {xx})  '
              f'sign flips {flips}  min |ref| at flips {float(r.abs()[(got > 0) != (r > 0)].max()) if flips else 0:.2e}')

    class Net(network.StyleNetwork):
        style_layers, content_layers, feature_loss_layers = GL['style'], GL['content'], GL['feature']
    o_total, o_grad, _ = oracle.generic_closure(x0, content, style, vw, GL['style'], GL['content'], GL['feature'])
    net = Net(style.cuda(), content.cuda(), vgg_weights=[(a.cuda(), b.cuda()) for a, b in vw])
    x = x0.cuda().requires_grad_(True)
    net(x, content.cuda())
    (net.get_total_current_style_loss(weight=1e5) + net.get_total_current_content_loss(weight=1)).backward()
    d = (x.grad.cpu() - o_grad)
    print(f'grad rel L2 {float(d.norm() / o_grad.norm()):.3e}')
    e = d.pow(2).sum(dim=(0, 1))
    tot = float(e.sum())
    # share of the squared error inside the 16x16 block that holds most of it
    blocks = e.unfold(0, 16, 16).unfold(1, 16, 16).sum(dim=(2, 3))
    by, bx = divmod(int(blocks.argmax()), blocks.shape[1])
    print(f'squared error: worst 16x16 block at (y {by * 16}, x {bx * 16}) holds {float(blocks.max()) / tot:.2%}; '
          f'rows {[round(float(v) / tot, 3) for v in e.sum(dim=1).unfold(0, 8, 8).sum(dim=1)]}')
    print('cols', [round(float(v) / tot, 3) for v in e.sum(dim=0).unfold(0, 8, 8).sum(dim=1)])


if __name__ == '__main__':
    main()
