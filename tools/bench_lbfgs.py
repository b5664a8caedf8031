#!/usr/bin/env python
"""The train_gatys optimizer loop (bench.run_lbfgs_loop) at several loss scales: prints one JSON line per
scale with the history fill, the losses and closures/s of optim.DeviceLBFGS and torch.optim.LBFGS."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    from styletransfer_b200 import _lib, network
    dev = torch.device('cuda', 0)
    L = _lib.lib()
    content = bench.seeded((1, 3, bench.SIZE, bench.SIZE), 100).to(dev)
    blocky = torch.nn.functional.interpolate(bench.seeded((1, 3, bench.SIZE // 16, bench.SIZE // 16), 111) * 2.0,
                                             scale_factor=16, mode='nearest').to(dev)
    net = network.StyleNetwork(blocky, content, vgg_weights=network.load_vgg19_prefix_weights(seed=1))
    net.compute_feature_loss = False
    variants = [int(v) for v in os.environ.get('LBFGS_VARIANTS', '0').split(',')]
    for scale in [float(a) for a in sys.argv[1:]] or [1e3, 1e4, 1e5]:
        for variant in variants:
            L.st_set_lbfgs_variant(variant)
            bench.LBFGS_LOSS_SCALE = scale
            out = bench.run_lbfgs_loop(net, content, dev, L)
            out['scale'], out['dots_variant'] = scale, variant
            print(json.dumps(out), flush=True)
    L.st_set_lbfgs_variant(0)


if __name__ == '__main__':
    main()
