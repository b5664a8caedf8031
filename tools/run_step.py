#!/usr/bin/env python
"""One step of a benchmark workload between cudaProfilerStart/Stop, for ncu:

    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
        --log-file gpurun_out/launches.csv python tools/run_step.py --workload fast_st

Workloads: gatys (512x512 closure), fast_st (batch 16, 256x256 closure + Adam),
infer (ImageTransformNet forward, batch --batch at --size)."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def seeded(shape, seed):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--workload', default='gatys', choices=['gatys', 'gatys_graph', 'fast_st', 'fast_st_graph', 'infer', 'lbfgs'])
    ap.add_argument('--size', type=int, default=0)
    ap.add_argument('--batch', type=int, default=0)
    ap.add_argument('--width', type=int, default=0, help='infer: frame width (height = --size)')
    ap.add_argument('--warmup', type=int, default=2)
    ap.add_argument('--steps', type=int, default=1)
    ap.add_argument('--precision', default=None)
    args = ap.parse_args()
    from oracle import oracle
    from styletransfer_b200 import network, ops
    ops.set_precision(args.precision or ops.PRECISION)
    dev = torch.device('cuda', 0)
    vw = network.load_vgg19_prefix_weights(seed=1)
    if args.workload == 'lbfgs':
        # one optimizer.step of the train_gatys loop (20 closures + 20 device L-BFGS iterations), history full
        S = args.size or 512
        content = seeded((1, 3, S, S), 100).to(dev)
        style = torch.nn.functional.interpolate(seeded((1, 3, S // 16, S // 16), 111) * 2.0, scale_factor=16, mode='nearest').to(dev)
        net = network.StyleNetwork(style, content, vgg_weights=vw)
        net.compute_feature_loss = False
        x = content.clone().requires_grad_(True)
        evaluate = net.gatys_closure(x, content, style_weight=1e10, content_weight=1e5, graph=False)    # see bench.LBFGS_LOSS_SCALE
        opt = net.get_content_optimizer(x, optt=torch.optim.LBFGS)
        args.warmup = max(args.warmup, 6)

        def step():
            opt.step(evaluate)
    elif args.workload == 'gatys_graph':
        # the closure train_gatys replays (loss-vector dot product instead of the getters), as a CUDA graph
        S = args.size or 512
        content, style = seeded((1, 3, S, S), 100).to(dev), seeded((1, 3, S, S), 101).to(dev)
        net = network.StyleNetwork(style, content, vgg_weights=vw)
        net.compute_feature_loss = False
        x = content.clone().requires_grad_(True)
        step = net.gatys_closure(x, content, 1e5, 1.0)
    elif args.workload == 'gatys':
        S = args.size or 512
        content, style = seeded((1, 3, S, S), 100).to(dev), seeded((1, 3, S, S), 101).to(dev)
        net = network.StyleNetwork(style, content, vgg_weights=vw)
        net.compute_feature_loss = False
        x = content.clone().requires_grad_(True)

        def step():
            x.grad = None
            net(x, content)
            (net.get_total_current_style_loss(weight=1e5) + net.get_total_current_content_loss(weight=1)).backward()
    else:
        S, B = args.size or 256, args.batch or 16
        style = seeded((1, 3, 256, 256), 201).to(dev)
        tnet = network.ImageTransformNet(style, B).to(dev)
        tnet.load_state_dict({k: v.clone() for k, v in oracle.make_transform_params(2, 3).items()})
        batch = seeded((B, 3, S, args.width or S), 202).to(dev)
        if args.workload == 'infer':
            def step():
                with torch.no_grad():
                    tnet(batch)
        else:
            loss_net = network.StyleNetwork(style, torch.rand([1, 3, 256, 256]).to(dev), vgg_weights=vw)
            loss_net.compute_feature_loss = False
            opt = tnet.get_optimizer()

            from styletransfer_b200 import dist as st_dist
            bucket = getattr(opt, 'bucket', None) or st_dist.GradBucket(tnet.parameters())
            graph = args.workload == 'fast_st_graph'
            closure = tnet.training_closure(batch, loss_net, bucket, graph=graph, optimizer=opt if graph else None)

            def step():
                opt.step(closure)
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.cudart().cudaProfilerStart()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    print(f'{args.workload}: {e0.elapsed_time(e1) / args.steps:.3f} ms per step')


if __name__ == '__main__':
    main()
