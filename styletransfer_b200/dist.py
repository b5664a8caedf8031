"""Data-parallel training of the transform nets: one process per GPU, one flat gradient
bucket, one all-reduce per step (NCCL over NVLink / NVSwitch; gloo in the CPU tests).

The reference has no distributed code; this module is the new C1 row of
SURVEY section 2b.  The global batch of ``static_train`` / ``video_train`` (network.py:687,
:969) is sharded over the ranks; weights, the frozen VGG prefix and the style targets are
replicated; InstanceNorm statistics are per sample, so no activation crosses GPUs.

Exactness against a 1-GPU run of the global batch: the style and content terms are MEANS
over the local batch (F.mse_loss, network.py:118, :163), so their gradients are averaged;
the total-variation term is a SUM over the batch (network.py:634-641), so the local TV
term is scaled by the world size before the average.  ``tv_scale()`` returns that factor.
The 1/world of the average is folded into the same pass that un-flattens nothing: the
parameters' ``.grad`` tensors are views into the bucket, so the optimiser reads the
reduced values in place.
"""
import os

import torch
import torch.distributed as dist


def env_rank_world():
    return (int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)),
            int(os.environ.get('LOCAL_RANK', 0)))


def init(backend=None, device=None):
    """Initialise torch.distributed from the torchrun environment (no-op for one process)."""
    rank, world, local = env_rank_world()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', '29500')
        backend = backend or ('nccl' if torch.cuda.is_available() else 'gloo')
        kw = {}
        if backend == 'nccl':
            # one GPU per rank: bind it BEFORE the process group exists (NCCL otherwise sees every rank on cuda:0)
            if device is None:
                device = torch.device('cuda', local % max(torch.cuda.device_count(), 1))
            torch.cuda.set_device(device)
            kw = {'device_id': device}
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, world


_FORCED_SINGLE = False


class single_process:
    """Context manager: inside it this process behaves as a world of one although torch.distributed is initialised
    (no sharding, no TV scaling, no all-reduce, no barrier).  bench.py uses it on rank 0 to time the same global
    batch on ONE GPU beside the data-parallel run (strong-scaling efficiency); the other ranks wait outside."""

    def __enter__(self):
        global _FORCED_SINGLE
        self.prev, _FORCED_SINGLE = _FORCED_SINGLE, True

    def __exit__(self, *exc):
        global _FORCED_SINGLE
        _FORCED_SINGLE = self.prev


def forced_single():
    return _FORCED_SINGLE


def _active():
    return dist.is_initialized() and not _FORCED_SINGLE


def rank():
    return dist.get_rank() if _active() else 0


def barrier():
    if _active():
        dist.barrier()


def world_size():
    return dist.get_world_size() if _active() else 1


def tv_scale():
    """Factor for the local total-variation term so that averaging the gradients over the
    ranks reproduces the single-process sum over the global batch."""
    return float(world_size())


def shard(batch, rank=None, world=None):
    """The rank's slice of a global batch (dim 0); the batch must divide evenly."""
    world = world_size() if world is None else world
    rank = (dist.get_rank() if _active() else 0) if rank is None else rank
    n = batch.shape[0]
    if n % world != 0:
        raise ValueError(f'global batch {n} does not divide over {world} ranks')
    per = n // world
    return batch[rank * per:(rank + 1) * per]


class GradBucket:
    """All parameter gradients of a module as views into ONE flat fp32 buffer.

    ``all_reduce()`` launches a single collective on a side stream as soon as the caller's
    stream has produced the gradients, scales by 1/world in the same stream and makes the
    caller's stream wait for it, so the optimiser step that follows sees averaged gradients
    and nothing synchronises with the host."""

    def __init__(self, params):
        self.params = list(params)          # frozen ones keep their slot (video_train epoch 0)
        if not self.params:
            raise ValueError('GradBucket: no parameters')
        dev = self.params[0].device
        self.numel = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(self.numel, dtype=torch.float32, device=dev)
        self.rebind()
        self._side = torch.cuda.Stream(device=dev) if dev.type == 'cuda' else None

    def zero(self):
        self.flat.zero_()

    def rebind(self):
        """Re-attach the views (an optimizer.zero_grad(set_to_none=True) drops them)."""
        off = 0
        for p in self.params:
            n = p.numel()
            if not p.requires_grad:
                p.grad = None
            elif p.grad is None or p.grad.data_ptr() != self.flat.data_ptr() + 4 * off:
                p.grad = self.flat[off:off + n].view_as(p)
            off += n

    def all_reduce(self, average=True):
        """``average=False`` leaves the SUM over the ranks in the bucket: the fused Adam kernel folds the 1/world into
        its gradient read (``optim.FusedAdam.grad_scale``), which saves the separate scaling pass."""
        world = world_size()
        if world == 1:
            return
        if self._side is None:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
            if average:
                self.flat.mul_(1.0 / world)
            return
        cur = torch.cuda.current_stream(self.flat.device)
        self._side.wait_stream(cur)
        with torch.cuda.stream(self._side):
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
            if average:
                self.flat.mul_(1.0 / world)
        cur.wait_stream(self._side)

    def nbytes(self):
        return self.numel * 4
