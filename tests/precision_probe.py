"""Measurement script (not a test): Gatys closure error vs the CPU oracle at several sizes and
precision modes.  Run on the GPU box: python -m tests.precision_probe [sizes...]"""
import json
import sys
import time

import torch

from oracle import oracle
from tests.util import rel_l2, seeded


def main():
    from styletransfer_b200 import network, ops
    sizes = [int(a) for a in sys.argv[1:]] or [128, 256, 512]
    vw = oracle.make_vgg_weights(1)
    vwc = [(w.cuda(), b.cuda()) for w, b in vw]
    for s in sizes:
        content, style = seeded((1, 3, s, s), 100), seeded((1, 3, s, s), 101)
        x0 = content + 0.3 * seeded((1, 3, s, s), 102)
        t0 = time.time()
        ref_total, ref_grad, ref_losses = oracle.gatys_closure(x0, content, oracle.style_targets(style, vw), vw)
        t_cpu = time.time() - t0
        # fp64 oracle: how well-defined is the fp32 reference itself?
        vw64 = [(w.double(), b.double()) for w, b in vw]
        t64, g64, _ = oracle.gatys_closure(x0.double(), content.double(), oracle.style_targets(style.double(), vw64), vw64)
        print(json.dumps({'size': s, 'mode': 'cpu_fp32_vs_fp64', 'grad_rel_l2': rel_l2(ref_grad, g64),
                          'loss_rel': abs(float(ref_total) - float(t64)) / abs(float(t64)), 'cpu_s': t_cpu}), flush=True)
        for mode in ops._MODES:
            ops.set_precision(mode)
            net = network.StyleNetwork(style.cuda(), content.cuda(), vgg_weights=vwc)
            x = x0.cuda().requires_grad_(True)
            net(x, content.cuda())
            total = net.get_total_current_style_loss(weight=1e5) + net.get_total_current_content_loss(weight=1)
            total.backward()
            print(json.dumps({'size': s, 'mode': mode,
                              'grad_rel_l2_vs_fp32': rel_l2(x.grad, ref_grad), 'grad_rel_l2_vs_fp64': rel_l2(x.grad, g64),
                              'loss_rel': abs(float(total) - float(ref_total)) / abs(float(ref_total))}), flush=True)


if __name__ == '__main__':
    main()
