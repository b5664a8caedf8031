"""Device-resident optimizer steps of the two training loops.

``FusedAdam``   -- ``torch.optim.Adam`` (reference network.py:643-649, step at :765) over a
                   transform net whose parameters, gradients and moments are flat buffers:
                   one ``st_adam_step`` launch instead of torch's per-tensor / foreach kernels,
                   reading the data-parallel gradient bucket in place.
``DeviceLBFGS`` -- ``torch.optim.LBFGS`` at the defaults the reference uses for the Gatys loop
                   (network.py:435 / :456, no line search) as three launches per iteration and no
                   host synchronisation (``csrc/optim.cu``); the closure is typically a replayed
                   CUDA graph, so an ``optimizer.step`` of 20 iterations is 20 x 4 launches.

Both are ``torch.optim.Optimizer`` subclasses with the reference's call shape
(``optimizer.step(closure)``); PyTorch supplies memory and streams only.
"""
import ctypes

import torch

from . import _lib
from .dist import GradBucket


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _bump_versions(tensors):
    """The kernels write through raw pointers: tell autograd / the packed-weight cache that the
    tensors changed."""
    inc = torch.autograd.graph.increment_version
    for t in tensors:
        inc(t)


class FusedAdam(torch.optim.Optimizer):
    """torch.optim.Adam semantics (per-parameter step count, parameters without a gradient are
    skipped) on flat storage.  ``bucket``: the ``dist.GradBucket`` whose flat buffer holds the
    gradients (created here if not given); the parameters' storage is moved into one flat buffer
    (``p.data`` become views; ``state_dict`` / ``load_state_dict`` of the module are unaffected)."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, bucket=None):
        params = list(params)
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        if len(self.param_groups) != 1:
            raise ValueError('FusedAdam: one parameter group')
        ps = self.param_groups[0]['params']
        if not ps or any(p.device.type != 'cuda' or p.dtype != torch.float32 for p in ps):
            raise _lib.StransferError('FusedAdam: fp32 CUDA parameters only (no CPU fallback)')
        self.L = _lib.lib()
        dev = ps[0].device
        self.bucket = bucket if bucket is not None else GradBucket(ps)
        if [id(p) for p in self.bucket.params] != [id(p) for p in ps]:
            raise ValueError('FusedAdam: the gradient bucket must hold the same parameters in the same order')
        numel = [p.numel() for p in ps]
        self.flat_param = torch.empty(sum(numel), dtype=torch.float32, device=dev)
        off = 0
        for p, n in zip(ps, numel):
            view = self.flat_param[off:off + n].view_as(p)
            view.copy_(p.data)
            p.data = view
            off += n
        self.exp_avg = torch.zeros_like(self.flat_param)
        self.exp_avg_sq = torch.zeros_like(self.flat_param)
        self.steps = torch.zeros(len(ps), dtype=torch.int32, device=dev)
        self.nseg = len(ps)
        seg = (ctypes.c_longlong * self.nseg)(*numel)
        self.nchunks = int(self.L.st_adam_chunk_count(seg, self.nseg))
        host = torch.empty(self.nchunks * 16, dtype=torch.uint8)
        _lib.check(self.L.st_adam_fill_chunks(seg, self.nseg, ctypes.c_void_p(host.data_ptr())), 'st_adam_fill_chunks')
        self.chunks = host.to(dev)
        self._active_key = None
        self.active = torch.zeros(self.nseg, dtype=torch.int32, device=dev)
        self.grad_scale = 1.0           # 1/world when the bucket holds the SUM of the ranks' gradients (dist.GradBucket)
        self._packs = None              # device table of the packed weight copies this optimizer keeps up to date
        self._pack_cache = None

    def set_packs(self, cache, named_params):
        """The Adam kernel rewrites the packed (split 16-bit, K-major) copies of the convolution weights that the
        training closure's kernels read (``transform._PackCache`` entries), so the ~28 weight-packing launches of a
        step disappear.  ``cache`` is frozen: its entries are owned by this optimizer."""
        import struct
        ps = self.param_groups[0]['params']
        index = {id(p): i for i, p in enumerate(ps)}
        dsts = [[] for _ in ps]
        for (name, mode, _prec, _fmt), (_tag, pw) in cache.entries.items():
            p = named_params[name + '.weight']
            dsts[index[id(p)]].append(pw)
        recs, off = [], 0
        for i, p in enumerate(ps):
            pws = dsts[i]
            if len(pws) > 2:
                raise _lib.StransferError('FusedAdam.set_packs: at most two packed copies per weight')
            if pws:
                cout, cin, ks, _ = p.shape
                rec = struct.pack('<qiiii', off, cout, cin, ks, len(pws))
                for pw in pws:
                    rec += struct.pack('<QQiiii', pw.hi, pw.lo or 0, pw.rows, pw.cols, pw.mode, pw.code)
                rec += b'\0' * 32 * (2 - len(pws))
            else:
                rec = struct.pack('<qiiii', off, 0, 0, 0, 0) + b'\0' * 64
            assert len(rec) == 88
            recs.append(rec)
            off += p.numel()
        table = torch.frombuffer(bytearray(b''.join(recs)), dtype=torch.uint8).clone()
        self._packs = table.to(self.flat_param.device)
        self._pack_keep = [pw for d in dsts for pw in d]
        self._pack_cache = cache
        cache.frozen = True
        cache.mark_synced(ps)

    def _refresh_active(self, ps):
        key = tuple(p.grad is not None for p in ps)
        if key != self._active_key:
            self.active.copy_(torch.tensor(key, dtype=torch.int32), non_blocking=False)
            self._active_key = key

    @torch.no_grad()
    def launch(self):
        """The Adam kernel on the current contents of the gradient bucket (stream-ordered, capturable)."""
        g = self.param_groups[0]
        ps = g['params']
        flat = self.bucket.flat
        off = 0
        for p in ps:                       # a gradient that is not a view of the bucket (eager zero_grad + backward) is copied in
            n = p.numel()
            if p.grad is not None and p.grad.data_ptr() != flat.data_ptr() + 4 * off:
                flat[off:off + n].copy_(p.grad.reshape(-1))
            off += n
        self._refresh_active(ps)
        b1, b2 = g['betas']
        _lib.check(self.L.st_adam_step(self.flat_param.data_ptr(), flat.data_ptr(), self.exp_avg.data_ptr(),
                                       self.exp_avg_sq.data_ptr(), self.chunks.data_ptr(), self.nchunks,
                                       self.steps.data_ptr(), self.active.data_ptr(), self.nseg,
                                       float(g['lr']), float(b1), float(b2), float(g['eps']), float(g['weight_decay']),
                                       float(self.grad_scale), None if self._packs is None else self._packs.data_ptr(),
                                       _stream()), 'st_adam_step')

    @torch.no_grad()
    def step(self, closure=None):
        # a closure that already CONTAINS this optimizer's update (the whole step captured in one CUDA graph:
        # closure -> all-reduce -> Adam, network.training_closure(optimizer=...)) is simply replayed
        if closure is not None and getattr(closure, 'includes_optimizer', None) is self:
            with torch.enable_grad():
                loss = closure()
            self._after_update()
            return loss
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        self.launch()
        self._after_update()
        return loss

    def _after_update(self):
        ps = self.param_groups[0]['params']
        _bump_versions(ps)
        if self._pack_cache is not None:
            self._pack_cache.mark_synced(ps)


class DeviceLBFGS(torch.optim.Optimizer):
    """torch.optim.LBFGS(lr=1, max_iter=20, history_size=100, line_search_fn=None) for ONE
    contiguous fp32 CUDA tensor (the Gatys image, network.py:403-409).  ``step(closure)`` runs
    ``max_iter`` rounds of closure -> st_lbfgs_iteration; torch's early exits (gradient /
    step / loss-change tolerances, non-descent direction) set a flag on the device that turns the
    remaining rounds of the call into no-ops; the host reads the flag RUN_AHEAD rounds late (off the
    critical path: the device queue stays that deep) and leaves the loop then.  The closure must write
    ``param.grad`` (any storage; its address is read after every call) and return the loss as a
    device scalar."""

    def __init__(self, params, lr=1.0, max_iter=20, history_size=100, tolerance_grad=1e-7, tolerance_change=1e-9):
        params = list(params)
        super().__init__(params, dict(lr=lr, max_iter=max_iter, history_size=history_size,
                                      tolerance_grad=tolerance_grad, tolerance_change=tolerance_change))
        ps = self.param_groups[0]['params']
        if len(self.param_groups) != 1 or len(ps) != 1:
            raise ValueError('DeviceLBFGS optimises exactly one tensor (the image)')
        p = ps[0]
        if p.device.type != 'cuda' or p.dtype != torch.float32 or not p.is_contiguous():
            raise _lib.StransferError('DeviceLBFGS: a contiguous fp32 CUDA tensor is required (no CPU fallback)')
        self.L = _lib.lib()
        self.p = p
        self.n = p.numel()
        self.m = int(history_size)
        self.ld = (self.n + 3) // 4 * 4
        nbytes = int(self.L.st_lbfgs_state_bytes(self.m, self.n))
        if nbytes == 0:
            raise ValueError('DeviceLBFGS: history_size must be in 1..100')
        dev = p.device
        self.state_buf = torch.empty(nbytes // 8, dtype=torch.float64, device=dev)
        self.S = torch.zeros(self.m, self.ld, dtype=torch.float32, device=dev)
        self.Y = torch.zeros(self.m, self.ld, dtype=torch.float32, device=dev)
        self.g_prev = torch.zeros(self.n, dtype=torch.float32, device=dev)
        self.dir = torch.zeros(self.n, dtype=torch.float32, device=dev)
        _lib.check(self.L.st_lbfgs_init(self.state_buf.data_ptr(), self.m, self.n, float(lr), float(tolerance_grad),
                                        float(tolerance_change), _stream()), 'st_lbfgs_init')
        # the halt flag of iteration j is copied to pinned memory; the host looks at it RUN_AHEAD iterations
        # later (the device queue stays that deep), so a stopped step costs at most RUN_AHEAD spare closures
        self._flag = [torch.zeros(1, dtype=torch.float64).pin_memory() for _ in range(self.RUN_AHEAD + 1)]
        self._flag_ev = [torch.cuda.Event() for _ in range(self.RUN_AHEAD + 1)]

    RUN_AHEAD = 2

    FIELDS = ('k', 'head', 'n_iter', 't', 'H_diag', 'd_dot_g', 'g_dot_g', 'halt', 'skip', 'loss_prev', 'pushed', 'new',
              'cg', 't_old', 'run_b', 'lr', 'tol_grad', 'tol_change', 'm', 'gtd', 'gmax', 'dmax', 'loss')

    def info(self):
        """Diagnostics (synchronises): the header of the device state."""
        h = self.state_buf[:32].cpu().tolist()
        return dict(zip(self.FIELDS, h))

    @torch.no_grad()
    def step(self, closure):
        g = self.param_groups[0]
        first_loss = None
        ring = self.RUN_AHEAD + 1
        for it in range(int(g['max_iter'])):
            if it >= self.RUN_AHEAD:
                self._flag_ev[(it - self.RUN_AHEAD) % ring].synchronize()
                if float(self._flag[(it - self.RUN_AHEAD) % ring]) != 0.0:
                    break                      # torch left its loop here; the remaining rounds would be no-ops
            with torch.enable_grad():
                loss = closure()
            if first_loss is None:
                first_loss = loss.detach().clone()       # torch returns orig_loss; a graphed closure reuses the buffer
            grad = self.p.grad
            if grad is None or not grad.is_contiguous() or grad.dtype != torch.float32:
                raise _lib.StransferError('DeviceLBFGS: the closure must leave a contiguous fp32 .grad on the image')
            loss_t = loss.detach().reshape(-1)
            _lib.check(self.L.st_lbfgs_iteration(self.p.data_ptr(), grad.data_ptr(), loss_t.data_ptr(),
                                                 self.g_prev.data_ptr(), self.dir.data_ptr(), self.S.data_ptr(),
                                                 self.Y.data_ptr(), self.n, self.ld, self.m, self.state_buf.data_ptr(),
                                                 1 if it == 0 else 0, _stream()), 'st_lbfgs_iteration')
            _bump_versions([self.p])
            self._flag[it % ring].copy_(self.state_buf[7:8], non_blocking=True)        # H_HALT
            self._flag_ev[it % ring].record()
        return first_loss
