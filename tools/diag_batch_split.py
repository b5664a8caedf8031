#!/usr/bin/env python
"""Single GPU: the fast_st gradient of a global batch against the mean of the gradients of its two halves (what two
data-parallel ranks compute: TV term scaled by 2, then averaged).  The two differ only by summation order unless some
kernel's arithmetic depends on the batch size; prints the relative L2 distance per parameter tensor (worst first)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def seeded(shape, seed):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed))


def main():
    dev = torch.device('cuda', 0)
    from oracle import oracle
    from styletransfer_b200 import dist as st_dist, network
    B, S = 8, 96
    style = seeded((1, 3, S, S), 1).to(dev)
    vw = network.load_vgg19_prefix_weights(seed=1)
    params = oracle.make_transform_params(2, 3)
    batch = seeded((B, 3, S, S), 2).to(dev)

    def grads_of(b, tv_scale):
        net = network.ImageTransformNet(style, b.shape[0]).to(dev)
        net.load_state_dict({k: v.clone() for k, v in params.items()})
        loss_net = network.StyleNetwork(style, torch.rand([1, 3, S, S]).to(dev), vgg_weights=vw)
        bucket = st_dist.GradBucket(net.parameters())
        total, _ = net.closure_loss(b, loss_net, tv_scale=tv_scale)
        total.backward()
        torch.cuda.synchronize()
        names = [n for n, _ in net.named_parameters()]
        sizes = [p.numel() for p in net.parameters()]
        return bucket.flat.clone(), names, sizes

    g_ref, names, sizes = grads_of(batch, 1.0)
    g0, _, _ = grads_of(batch[:B // 2], 2.0)
    g1, _, _ = grads_of(batch[B // 2:], 2.0)
    g_dp = 0.5 * (g0 + g1)
    print(f'all parameters: rel L2 {float((g_dp - g_ref).norm() / g_ref.norm()):.3e}')
    rows, o = [], 0
    for n, k in zip(names, sizes):
        a, b = g_dp[o:o + k], g_ref[o:o + k]
        rows.append((float((a - b).norm()), float((a - b).norm() / (b.norm() + 1e-30)), float(b.norm()), n))
        o += k
    # by absolute contribution: conv biases in front of an InstanceNorm have analytically zero gradients (pure rounding noise)
    for ab, e, nb, n in sorted(rows, reverse=True)[:10]:
        print(f'  |diff| {ab:.3e}  rel {e:.3e}  |g| {nb:.3e}  {n}')


if __name__ == '__main__':
    main()
