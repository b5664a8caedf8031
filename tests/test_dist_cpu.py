"""World-size-2 gloo tests (CPU) of the data-parallel host logic in styletransfer_b200/dist.py:
batch sharding, the flat gradient bucket, the single all-reduce and the TV-term scaling that
makes an N-rank step reproduce the 1-process step on the global batch."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _toy_params(seed=0):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(3, 3, 3, 3, generator=g) * 0.2, torch.randn(3, generator=g) * 0.1]


def _toy_loss(params, batch, tv_scale=1.0):
    """Same structure as the static_train closure: mean-type terms + a sum-type TV term."""
    w, b = params
    out = torch.nn.functional.conv2d(batch, w, b, padding=1)
    mean_term = torch.nn.functional.mse_loss(out, batch)                      # mean over the batch
    gram = oracle.gram_matrix(out)
    style_term = torch.nn.functional.mse_loss(gram, torch.zeros(1, 3, 3).expand_as(gram))
    return mean_term + style_term + tv_scale * oracle.tv_loss(out, 1e-2)      # TV: sum over the batch


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from styletransfer_b200 import dist as st_dist
    r, w = st_dist.init(backend='gloo')
    assert (r, w) == (rank, world) and st_dist.world_size() == world and st_dist.tv_scale() == float(world)
    params = [p.clone().requires_grad_(True) for p in _toy_params()]
    bucket = st_dist.GradBucket(params)
    assert bucket.numel == sum(p.numel() for p in params)
    batch = torch.randn(4, 3, 8, 8, generator=torch.Generator().manual_seed(5))
    local = st_dist.shard(batch)
    assert local.shape[0] == 2 and torch.equal(local, batch[rank * 2:rank * 2 + 2])
    opt = torch.optim.Adam(params)
    for _ in range(2):
        opt.zero_grad()                       # drops the views: rebind() must restore them
        bucket.rebind()
        bucket.zero()
        _toy_loss(params, local, tv_scale=st_dist.tv_scale()).backward()
        assert params[0].grad.data_ptr() == bucket.flat.data_ptr()
        bucket.all_reduce()
        grads = [p.grad.clone() for p in params]
        opt.step()
    # frozen parameters keep their slot but carry no gradient
    params[1].requires_grad_(False)
    bucket.rebind()
    assert params[1].grad is None and params[0].grad is not None
    q.put((rank, [g.numpy() for g in grads], [p.detach().numpy() for p in params]))
    dist.barrier()
    dist.destroy_process_group()


def test_data_parallel_step_matches_single_process():
    world, port = 2, _free_port()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted([q.get(timeout=240) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # single process on the global batch
    params = [p.clone().requires_grad_(True) for p in _toy_params()]
    batch = torch.randn(4, 3, 8, 8, generator=torch.Generator().manual_seed(5))
    opt = torch.optim.Adam(params)
    for _ in range(2):
        opt.zero_grad()
        _toy_loss(params, batch).backward()
        ref_grads = [p.grad.clone() for p in params]
        opt.step()
    for rank, grads, new_params in results:
        for g, rg in zip(grads, ref_grads):
            assert torch.allclose(torch.from_numpy(g), rg, rtol=1e-4, atol=1e-7), rank
        for p, rp in zip(new_params, params):
            assert torch.allclose(torch.from_numpy(p), rp.detach(), rtol=1e-4, atol=1e-6), rank
    # both ranks hold identical weights after the step
    for a, b in zip(results[0][2], results[1][2]):
        assert (a == b).all()


def test_shard_rejects_ragged_batches():
    from styletransfer_b200 import dist as st_dist
    import pytest
    with pytest.raises(ValueError):
        st_dist.shard(torch.zeros(5, 3, 4, 4), rank=0, world=2)
    assert st_dist.shard(torch.arange(8).reshape(8, 1), rank=3, world=4).flatten().tolist() == [6, 7]
