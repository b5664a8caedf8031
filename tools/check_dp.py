#!/usr/bin/env python
"""Run under torchrun on >= 2 GPUs: the data-parallel fast_st gradient (local shard + TV scaling +
flat-bucket all-reduce) must equal the single-GPU gradient of the global batch, and three Adam steps of the
data-parallel loop -- eager, and with the whole step (closure + NCCL all-reduce + Adam) replayed as ONE CUDA graph --
must land where three single-process steps on the global batches land.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/check_dp.py
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
# InstanceNorm statistics with a pixel split that does not depend on the batch size: a shard of 4 images and the global
# batch of 8 then run bit-identical forwards, and this script compares the data-parallel PLUMBING (sharding, TV scaling,
# all-reduce, 1/world in Adam).  With the default, batch-size-tuned split the per-image statistics differ in the last
# bit, a handful of ReLU decisions of near-zero activations flip, and conv weight gradients differ by ~1e-3 relative
# (tools/diag_batch_split.py) -- the same kind of difference cuDNN's batch-dependent algorithm choice produces.
os.environ.setdefault('STRANSFER_IN_SPLITS_INVARIANT', '1')


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

    # ---- three optimizer steps: single process on the global batches / DP eager / DP with one graph per step
    from styletransfer_b200 import optim as st_optim
    STEPS = 3
    batches = [seeded((B, 3, S, S), 10 + i).to(dev) for i in range(STEPS)]
    probe = seeded((2, 3, S, S), 99).to(dev)

    def train(b_of, dp, graph):
        net = network.ImageTransformNet(style, b_of(batches[0]).shape[0]).to(dev)
        net.load_state_dict({k: v.clone() for k, v in params.items()})
        loss_net = network.StyleNetwork(style, torch.rand([1, 3, S, S]).to(dev), vgg_weights=vw)
        loss_net.compute_feature_loss = False
        opt = st_optim.FusedAdam(list(net.parameters()))
        buf = b_of(batches[0]).clone().contiguous()
        losses = []
        if dp:
            closure = net.training_closure(buf, loss_net, opt.bucket, graph=graph, optimizer=opt)
            assert bool(getattr(closure, 'includes_optimizer', None)) == bool(graph)
            for b in batches:
                buf.copy_(b_of(b))
                losses.append(float(opt.step(closure)))
        else:
            # what one process does with the global batch: no sharding, no TV scaling, no all-reduce
            for b in batches:
                opt.bucket.rebind()
                opt.bucket.zero()
                total, _ = net.closure_loss(b, loss_net)
                net._backward_into_bucket(total)
                losses.append(float(total))
                opt.step()
        torch.cuda.synchronize()
        with torch.no_grad():
            out = net(probe).clone()
        return opt.flat_param.clone(), out, losses

    w0 = torch.cat([params[k].reshape(-1) for k, _ in network.ImageTransformNet(style, 1).named_parameters()]).to(dev)
    w_ref, o_ref, l_ref = train(lambda b: b, False, False)
    moved = float((w_ref - w0).norm() / w0.norm())
    rl = lambda a, b: float((a - b).norm() / b.norm())
    results = {}
    for name, graph in (('eager', False), ('one graph per step, NCCL captured', True)):
        w, o, l = train(st_dist.shard, True, graph)
        results[name] = (w, o)
        perr, oerr = rl(w, w_ref), rl(o, o_ref)
        # a rank's loss is the loss of its shard with the TV term scaled by world: the MEAN over ranks is the global one
        lt = torch.tensor(l, device=dev, dtype=torch.float64)
        dist.all_reduce(lt)
        lerr = float(((lt / world).cpu() - torch.tensor(l_ref, dtype=torch.float64)).abs().max() / abs(l_ref[0]))
        print(f'rank {rank}: dp{world} {STEPS} Adam steps ({name}): weights vs single process rel L2 {perr:.3e} '
              f'(weights moved by {moved:.3e}), output on a probe batch rel L2 {oerr:.3e}, mean-over-ranks loss rel {lerr:.3e}', flush=True)
        # Adam turns the rounding noise of analytically-zero gradients (conv biases in front of InstanceNorm) into
        # +-lr steps of random sign, so the WEIGHTS are compared against how far training moved them and the
        # function the net computes (which those biases do not enter) is compared tightly
        assert perr < 0.1 * moved, (perr, moved)
        assert oerr < 1e-2, oerr
        assert lerr < 2e-3, lerr
    (we, oe), (wg, og) = results['eager'], results['one graph per step, NCCL captured']
    gerr, goerr = rl(wg, we), rl(og, oe)
    print(f'rank {rank}: graph vs eager DP step: weights rel L2 {gerr:.3e}, probe output rel L2 {goerr:.3e}', flush=True)
    assert gerr < 0.1 * moved and goerr < 1e-2, (gerr, goerr)
    # every rank holds the same weights after the loop
    wl = [torch.empty_like(wg) for _ in range(world)]
    dist.all_gather(wl, wg)
    assert all(torch.equal(wl[0], t) for t in wl), 'ranks diverged'
    print(f'rank {rank}: weights bit-identical across ranks', flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
