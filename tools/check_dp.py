#!/usr/bin/env python
"""Run under torchrun on >= 2 GPUs: the data-parallel fast_st gradient (local shard + TV scaling +
flat-bucket all-reduce) must equal the single-GPU gradient of the global batch.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/check_dp.py
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def seeded(shape, seed):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed))


def main():
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    dist.init_process_group('nccl', device_id=dev)
    from oracle import oracle
    from styletransfer_b200 import dist as st_dist, network
    B, S = 4 * world, 96
    style = seeded((1, 3, S, S), 1).to(dev)
    vw = network.load_vgg19_prefix_weights(seed=1)
    params = oracle.make_transform_params(2, 3)
    batch = seeded((B, 3, S, S), 2).to(dev)

    def grads_of(b, tv_scale, reduce):
        net = network.ImageTransformNet(style, b.shape[0]).to(dev)
        net.load_state_dict({k: v.clone() for k, v in params.items()})
        loss_net = network.StyleNetwork(style, torch.rand([1, 3, S, S]).to(dev), vgg_weights=vw)
        bucket = st_dist.GradBucket(net.parameters())
        total, _ = net.closure_loss(b, loss_net, tv_scale=tv_scale)
        total.backward()
        if reduce:
            bucket.all_reduce()
        torch.cuda.synchronize()
        return bucket.flat.clone(), float(total)

    g_dp, l_dp = grads_of(st_dist.shard(batch), float(world), True)
    g_ref, l_ref = grads_of(batch, 1.0, False)
    err = float((g_dp - g_ref).norm() / g_ref.norm())
    print(f'rank {rank}: dp{world} gradient vs single-GPU global batch: rel L2 {err:.3e} (local loss {l_dp:.6f}, global {l_ref:.6f})', flush=True)
    assert err < 1e-4, err
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
