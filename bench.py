#!/usr/bin/env python
"""Benchmark of the stransfer.network hot path on B200 (contract: see the task brief).

Headline metric (BASELINE.json, configs[1]): Gatys VGG-19 forward+backward closures per
second at 512x512, one content/style pair per GPU (N GPUs = N independent pairs, no
collective, weak scaling).  A "step" is one closure of train_gatys
(reference network.py:438-454): zero_grad -> StyleNetwork.forward -> 1e5*style + content
-> backward to the image.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

The same JSON line carries, as named objects, the other BASELINE configs (fast_st training at batch 16 of 256x256,
video_st training on 512x512 clips with and without optical flow, convert-image at 1080p, the train_gatys loop end to
end), each with its own ``roofline`` and -- at N = 1 -- ``cpu_baseline`` (the unmodified reference on the host cores,
oracle/_ref, run in a child process with the GPUs hidden).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

_REAL_STDOUT = sys.stdout
SIZE = 512
STYLE_WEIGHT, CONTENT_WEIGHT = 100_000.0, 1.0
# algorithmic work of one closure at SxS, batch 1
GF_PREFIX_FWD = 224_640 * SIZE * SIZE / 1e9          # 58.89 GF @512
GF_GRAM = 40_960 * SIZE * SIZE / 1e9                 # 10.74 GF @512 (fwd; bwd the same)
GF_CLOSURE = 2 * GF_PREFIX_FWD + 2 * GF_GRAM         # 139.25 GF @512
METRIC = 'gatys_vgg19_fwd_bwd_closures_per_s_512'
# config.workload: ONE string per workload, printed verbatim by both arms (ours and --impl reference)
WORKLOADS = {
    'gatys512': 'gatys_st 512x512, one content/style pair per GPU, train_gatys closure fwd+bwd (BASELINE configs[1])',
    'gatys256': 'gatys_st 256x256, one content/style pair, train_gatys closure fwd+bwd (BASELINE configs[0])',
    'fast_st': 'fast_st train 256x256, global batch 16, static_train closure + Adam step (BASELINE configs[2])',
    'convert_image': 'fast_st convert-image: ImageTransformNet forward on 1080x1920 frames (BASELINE configs[3])',
    'video_st': 'video_st train on 512x512 clips, global batch 8 clips, video_train closure + Adam step (BASELINE configs[4])',
}
DTYPES = {'f16x3': 'f16', 'bf16x6': 'bf16', 'bf16x3': 'bf16', 'bf16': 'bf16'}
DTYPE_NOTES = {
    'f16x3': 'operands as two fp16 planes (hi + lo, 22 mantissa bits), 3 tcgen05 kind::f16 passes per product, fp32 accumulate in TMEM',
    'bf16x6': 'operands as three bf16 planes forward (6 passes) / two backward (3 passes), fp32 accumulate in TMEM',
    'bf16x3': 'operands as two bf16 planes, 3 passes, fp32 accumulate', 'bf16': 'one bf16 plane, 1 pass, fp32 accumulate'}
MMA_PASSES = {'f16x3': 3, 'bf16x6': 4.5, 'bf16x3': 3, 'bf16': 1}      # bf16x6: forward 6 / backward 3


def seeded(shape, seed):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed))


def peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm=p['hbm_gbs'], tf_burst=p['bf16_tflops'], tf_sustained=p['bf16_tflops_sustained'], src='measured')
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, src='fallback')


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
        try:
            self.p = subprocess.Popen(['nvidia-smi', '-i', str(index), '--query-gpu=' + self.Q,
                                       '--format=csv,noheader,nounits', '-lms', '100'], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def lines(self):
        try:
            with open(self.f.name) as fh:
                return sum(1 for ln in fh if ln.count(',') >= 5)
        except OSError:
            return 0

    def busy_until(self, n, work, timeout_s):
        """Keeps the GPU under the benchmark's own load until n samples exist (the timed region of a
        short run is shorter than one nvidia-smi period; samples taken right before and right after
        it under the identical load bracket it)."""
        if self.p is None:
            return
        t0 = time.perf_counter()
        while self.lines() < n and time.perf_counter() - t0 < timeout_s:
            for _ in range(20):
                work()
            torch.cuda.synchronize()

    def stop(self):
        if self.p is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.p.terminate()
        self.p.wait()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f:
            parts = [x.strip() for x in line.split(',')]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, val in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), parts[2:6]):
                if val.lower().startswith('active'):
                    reasons.add(name)
        self.f.close()
        os.unlink(self.f.name)
        return {'sm_mhz': statistics.median(sm) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


# ---------------------------------------------------------------------------------------------------------------
# Live kernel timing for the roofline objects: CUDA events around the launches of one kernel family while the step
# runs EAGERLY (the timed `value` comes from the graph replay; this pass only attributes time to kernels).
# ---------------------------------------------------------------------------------------------------------------
class KernelTrace:
    """Wraps ``ops.conv2d`` / ``ops.instnorm_fwd`` / ``ops.instnorm_bwd`` for the duration of ``run``: every call is
    bracketed by CUDA events on the launching stream (torch's current stream = the stream the C ABI receives), with
    a short device-side spin in front so that the events bracket the kernels and not the host's launch latency."""

    def __init__(self, ops, L):
        self.ops, self.L = ops, L
        self.conv, self.norm = [], []
        # what the bracket itself costs: two event records back to back behind the same spin, no kernel in between
        # (2.6 us on B200); subtracted from every bracket.  A bracket around a ~2 us kernel reads 6.1 us, so this is a
        # lower bound of the overhead and the reported launch durations stay on the conservative side.
        ts = []
        for _ in range(20):
            _, e0, e1 = self._bracket(lambda: None, (), {})
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        self.overhead_ms = statistics.median(ts)

    def _bracket(self, fn, a, k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(200000)
        e0.record()
        out = fn(*a, **k)
        e1.record()
        return out, e0, e1

    def run(self, step, steps):
        ops, L = self.ops, self.L
        real = ops.conv2d, ops.instnorm_fwd, ops.instnorm_bwd, ops.conv2d_routed, ops.conv2d_pool

        def conv2d(xs, w, *a, **k):
            out, e0, e1 = self._bracket(real[0], (xs, w) + a, k)
            kid = L.st_last_conv_kernel()
            if kid == 4:                                           # conv_imgrad_tc: image gradient, project-then-gather form
                self.conv.append((f'conv_imgrad_tc_kernel {xs.C}->{k.get("out_nchw_channels", 3)} @{xs.H}x{xs.W}x{xs.N}', 0,
                                  2.0 * xs.N * xs.H * xs.W * xs.C * k.get('out_nchw_channels', 3) * 9, e0, e1))
            if kid == 1:                                           # conv_tc: tcgen05, stride 1
                ks = w.ks if w is not None else k['w_ptrs'][3]
                # algorithmic output channels (the packed row count is padded, e.g. 3 -> 16 for the image gradient)
                cout = (w.cout if w.mode == 0 else w.cin) if w is not None else k['w_img_rows']
                self.conv.append((f'conv_tc_kernel<{ks}x{ks}> {xs.C}->{cout} @{xs.H}x{xs.W}x{xs.N}', ks,
                                  2.0 * xs.N * xs.H * xs.W * xs.C * cout * ks * ks, e0, e1))
            return out

        def conv2d_routed(xs, w_ptrs, w_img_rows, *a, **k):
            # the Gram-backward launch that also routes the max-pool gradient: a 1x1 conv_tc launch when it returns a tensor
            out, e0, e1 = self._bracket(real[3], (xs, w_ptrs, w_img_rows) + a, k)
            if out is not None:
                self.conv.append((f'conv_tc_kernel<1x1, routed pool gradient> {xs.C}->{w_img_rows} @{xs.H}x{xs.W}x{xs.N}', 1,
                                  2.0 * xs.N * xs.H * xs.W * xs.C * w_img_rows, e0, e1))
            return out

        def conv2d_pool(xs, w, *a, **k):
            # forward 3x3 conv_tc launch whose epilogue also writes the ReLU + max-pool output
            out, e0, e1 = self._bracket(real[4], (xs, w) + a, k)
            if out is not None:
                self.conv.append((f'conv_tc_kernel<3x3, fused relu + max-pool> {xs.C}->{w.cout} @{xs.H}x{xs.W}x{xs.N}', 3,
                                  2.0 * xs.N * xs.H * xs.W * xs.C * w.cout * 9, e0, e1))
            return out

        def instnorm_fwd(x, *a, **k):
            out, e0, e1 = self._bracket(real[1], (x,) + a, k)
            y = out[0]
            # statistics pass reads x, apply pass reads x and writes y (x4 pixels when the upsample is fused)
            self.norm.append(('instnorm_fwd (in_stats + in_finalize + in_apply_fwd)', 2.0 * x.buf.numel() * 2 + y.buf.numel() * 2, e0, e1))
            return out

        def instnorm_bwd(x, gy, *a, **k):
            out, e0, e1 = self._bracket(real[2], (x, gy) + a, k)
            dx = out[0]
            rd = (x.buf[0].numel() * min(x.planes, ops.bwd_planes()) + gy.buf.numel()) * 2
            self.norm.append(('instnorm_bwd (in_stats + in_finalize_bwd + in_apply_bwd)', 2.0 * rd + dx.buf.numel() * 2, e0, e1))
            return out
        ops.conv2d, ops.instnorm_fwd, ops.instnorm_bwd, ops.conv2d_routed, ops.conv2d_pool = conv2d, instnorm_fwd, instnorm_bwd, conv2d_routed, conv2d_pool
        try:
            torch.cuda.synchronize()
            for _ in range(steps):
                step()
            torch.cuda.synchronize()
        finally:
            ops.conv2d, ops.instnorm_fwd, ops.instnorm_bwd, ops.conv2d_routed, ops.conv2d_pool = real
        self.steps = steps

    @staticmethod
    def _medians(records, steps, t0, t1):
        """Launch i of every step is the same kernel on the same shapes: its duration is the MEDIAN over the steps
        (a bracket that caught a host stall is dropped).  -> [(name, *payload, ms)] for one step."""
        per = len(records) // max(steps, 1)
        return [records[i][:t0] + (statistics.median(r[t0].elapsed_time(r[t1]) for r in records[i::per]),) for i in range(per)]

    def _net(self, one):
        """Subtracts the empty-bracket overhead from the last field (ms) of every record."""
        return [r[:-1] + (max(r[-1] - self.overhead_ms, 1e-4),) for r in one]

    def tensor_roofline(self, precision, step_ms, what):
        """Roofline of the dominant kernel family: the 3x3 conv_tc launches."""
        pk = peaks()
        one = self._net(self._medians(self.conv, self.steps, 3, 4))
        k3 = [r for r in one if r[1] == 3]
        fl, ms = sum(r[2] for r in k3), sum(r[3] for r in k3)
        achieved = fl / (ms * 1e-3) / 1e12 if ms > 0 else 0.0
        passes = MMA_PASSES[precision]
        groups = {}
        for name, _, f, t in one:
            g = groups.setdefault(name, [0.0, 0.0, 0])
            g[0] += f; g[1] += t; g[2] += 1
        roofline = {'bound': 'tensor', 'kernel': f'conv_tc_kernel<3x3> ({len(k3)} launches per {what})',
                    'achieved': achieved, 'peak': pk['tf_sustained'], 'unit': 'TFLOP/s', 'frac': achieved / pk['tf_sustained'],
                    'peak_source': pk['src'] + ' bf16_tflops_sustained (kernel timed inside a long step)',
                    'avg_launch_ms': ms / max(len(k3), 1), 'algorithmic_gflop_per_launch': fl / max(len(k3), 1) / 1e9,
                    'mma_passes_per_product': passes, 'issued_frac': achieved * passes / pk['tf_sustained'],
                    'bracket_overhead_us': self.overhead_ms * 1e3,
                    'share_of_step': ms / step_ms if step_ms else None}
        kernels = {k: {'tflops': g[0] / (g[1] * 1e-3) / 1e12, 'ms_per_launch': g[1] / g[2], 'launches_per_step': g[2]}
                   for k, g in sorted(groups.items())}
        return roofline, kernels

    def hbm_roofline(self, step_ms):
        """The InstanceNorm kernels (HBM-bound): algorithmic bytes = every plane read by the statistics and apply
        passes + every plane written, per call."""
        if not self.norm:
            return None
        pk = peaks()
        one = self._net(self._medians(self.norm, self.steps, 2, 3))
        by, ms = sum(r[1] for r in one), sum(r[2] for r in one)
        achieved = by / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        return {'bound': 'hbm', 'kernel': f'InstanceNorm2d forward/backward kernel groups ({len(one)} calls per step)',
                'achieved': achieved, 'peak': pk['hbm'], 'unit': 'GB/s', 'frac': achieved / pk['hbm'], 'peak_source': pk['src'] + ' hbm_gbs',
                'avg_call_ms': ms / len(one), 'algorithmic_mbytes_per_call': by / len(one) / 1e6,
                'share_of_step': ms / step_ms if step_ms else None, 'traffic': None}


def ncu_traffic(workload, precision):
    """dram__bytes_read.sum + dram__bytes_write.sum per 3x3 conv_tc launch from the committed `ncu --set full` capture of
    THIS build (profiles/conv_tc_traffic.json, written by tools/summarize_metrics.py --traffic together with the digest
    of the kernel sources it profiled).  A capture of another build is reported as stale and not used."""
    from styletransfer_b200 import build
    try:
        with open(os.path.join(ROOT, 'profiles', 'conv_tc_traffic.json')) as f:
            table = json.load(f)
        entry = table[f'{workload}/{precision}']
    except (OSError, ValueError, KeyError):
        return None, 'no ncu --set full capture committed for this workload / precision'
    if entry.get('sources_digest') != build.kernel_sources_digest():
        return None, f"stale: {entry.get('source')} profiled another build of csrc/"
    return entry['bytes_per_launch'], entry['source']


class HostFedStep:
    """One end-to-end step with HOST buffers, pipelined the way an input pipeline feeds a training
    loop: while the device runs step i, a copy stream uploads the input of step i+1 from pinned
    host memory into a staging buffer; step i+1 starts with a device-to-device copy into the
    tensor the step reads.  The loss of every step is copied to pinned host memory and the host
    waits for it one step later (it never runs more than one step ahead of the device)."""

    def __init__(self, host_in, dev_in, run, get_loss):
        self.host_in, self.dev_in, self.run, self.get_loss = host_in, dev_in, run, get_loss
        self.copy_stream = torch.cuda.Stream()
        self.stage = [torch.empty_like(dev_in), torch.empty_like(dev_in)]
        self.uploaded = [torch.cuda.Event(), torch.cuda.Event()]
        self.consumed = [torch.cuda.Event(), torch.cuda.Event()]
        self.done = [torch.cuda.Event(), torch.cuda.Event()]
        self.host_loss = [torch.zeros(1).pin_memory(), torch.zeros(1).pin_memory()]
        self.i = 0
        self.n_uploaded = 0
        for ev in self.consumed:
            ev.record()
        self._upload(0)

    def _upload(self, slot):
        src = self.host_in
        if isinstance(src, (list, tuple)):                           # a sequence of host inputs, one per step, cycled
            src = src[self.n_uploaded % len(src)]
        self.n_uploaded += 1
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.consumed[slot])        # the staging buffer's previous content was read
            self.stage[slot].copy_(src, non_blocking=True)
            self.uploaded[slot].record()

    def __call__(self):
        slot = self.i & 1
        self._upload(slot ^ 1)                                       # input of the next step
        cur = torch.cuda.current_stream()
        cur.wait_event(self.uploaded[slot])
        with torch.no_grad():
            self.dev_in.copy_(self.stage[slot])
        self.consumed[slot].record()
        self.run()
        self.host_loss[slot].copy_(self.get_loss().detach().reshape(1), non_blocking=True)
        self.done[slot].record()
        if self.i > 0:
            self.done[slot ^ 1].synchronize()                        # the previous step's loss is on the host
        self.i += 1

    def loss(self):
        torch.cuda.synchronize()
        return float(self.host_loss[(self.i - 1) & 1])


# torch.optim.LBFGS only stores a curvature pair when y.s > 1e-10.  On 512x512 synthetic noise with the CLI's
# default weights the curvature y.s of the (mean-type) losses stays below that threshold and NO pair ever passes the
# filter (both optimizers then do plain gradient steps).  The loop benchmark therefore uses the CLI's
# -sw / -cw options at 1e5 x the defaults (measured: at 1e3 x no pair passes yet, from 1e4 x the history fills).
LBFGS_LOSS_SCALE = 1.0e5


def run_lbfgs_loop(net, content, dev, L):
    """The loop train_gatys runs (network.py:435-456): optimizer.step(closure) with L-BFGS at torch's defaults
    (20 closures per step, history 100), closures per second with a FULL history, for the device-resident
    optimizer (optim.DeviceLBFGS, the default) and for torch.optim.LBFGS driving the same graph closure."""
    from styletransfer_b200 import optim as st_optim
    out = {}
    for name, make in (('device_lbfgs', st_optim.DeviceLBFGS), ('torch_lbfgs', torch.optim.LBFGS)):
        x = content.clone().requires_grad_(True)
        graph_closure = net.gatys_closure(x, content, LBFGS_LOSS_SCALE * STYLE_WEIGHT, LBFGS_LOSS_SCALE * CONTENT_WEIGHT)
        calls = [0]

        def evaluate():
            calls[0] += 1
            return graph_closure()
        opt = make([x])
        first = float(graph_closure().detach())
        for _ in range(6):                             # 120 iterations: the history ring is full
            opt.step(evaluate)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        steps, c0 = 3, calls[0]
        e0.record()
        for _ in range(steps):
            opt.step(evaluate)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        n = calls[0] - c0
        out[name] = {'closures_per_s': n / (ms * 1e-3), 'ms_per_iteration': ms / max(n, 1), 'timed_closures': n,
                     'loss_before': first, 'loss_after': float(graph_closure().detach())}
        if name == 'device_lbfgs':
            # the three optimizer kernels alone (dots sweep, solve, update sweep) on the full history
            info = opt.info()
            k, reps = int(info['k']), 20
            graph_closure()
            loss_t = graph_closure.out.detach().reshape(-1)
            e0.record()
            for _ in range(reps):
                L.st_lbfgs_iteration(x.data_ptr(), x.grad.data_ptr(), loss_t.data_ptr(), opt.g_prev.data_ptr(), opt.dir.data_ptr(),
                                     opt.S.data_ptr(), opt.Y.data_ptr(), opt.n, opt.ld, opt.m, opt.state_buf.data_ptr(), 1,
                                     torch.cuda.current_stream().cuda_stream)
            e1.record()
            torch.cuda.synchronize()
            t_ms = e0.elapsed_time(e1) / reps
            sweep_bytes = 2.0 * (2 * k + 3) * opt.n * 4            # two sweeps over 2k history vectors + g, g_prev, d
            pk = peaks()
            out[name]['optimizer_kernels'] = {'ms_per_iteration': t_ms, 'history_pairs': k, 'algorithmic_bytes': sweep_bytes,
                                              'achieved_gbs': sweep_bytes / (t_ms * 1e-3) / 1e9, 'peak_gbs': pk['hbm'],
                                              'frac': sweep_bytes / (t_ms * 1e-3) / 1e9 / pk['hbm'], 'bound': 'hbm'}
        del opt, evaluate, graph_closure, x
        torch.cuda.empty_cache()
    out['config'] = ('train_gatys loop at 512x512 (network.py:435-456): 3 optimizer.step calls (up to 60 closures + 60 L-BFGS '
                     'iterations) after 120 iterations, history 100 full; same CUDA-graph closure under both optimizers; '
                     'style image = 16x nearest-upsampled noise; style / content weights 1e10 / 1e5 (-sw / -cw at 1e5 x the defaults) so that '
                     'curvature pairs pass torch\'s y.s > 1e-10 filter and the history fills')
    return out


FAST_B, FAST_S = 16, 256
GF_FAST_PER_IMG = 106.6          # SURVEY.md section 8a row a13 (single-pass, no VGG wgrad)


def _step_tflops_roofline(value, gflop_per_unit, unit):
    """Step-level view: algorithmic TFLOP/s of the whole step against the sustained tensor peak."""
    pk = peaks()
    tf = value * gflop_per_unit / 1e3
    return {'tflops_algorithmic': tf, 'frac_of_peak': tf / pk['tf_sustained'], 'peak': pk['tf_sustained'],
            'algorithmic_gflop_per_' + unit: gflop_per_unit}


def run_fast_st(args, rank, world, dev, L, timed, trace=True):
    """BASELINE configs[2]: fast_st training images/s at 256x256, global batch 16 sharded over the ranks (strong
    scaling), one step = the static_train closure (network.py:690-731) + flat-bucket all-reduce + Adam step
    (network.py:765), all inside one replayed CUDA graph."""
    from oracle import oracle                                  # seeded initial weights only, outside every timed region
    from styletransfer_b200 import dist as st_dist, network, ops
    per = FAST_B // world
    style = seeded((1, 3, FAST_S, FAST_S), 201).to(dev)
    net = network.ImageTransformNet(style, per).to(dev)
    net.load_state_dict({k: v.clone() for k, v in oracle.make_transform_params(2, 3).items()})
    loss_net = network.StyleNetwork(style, torch.rand([1, 3, FAST_S, FAST_S]).to(dev),
                                    vgg_weights=network.load_vgg19_prefix_weights(seed=1))
    loss_net.compute_feature_loss = False
    opt = net.get_optimizer()
    bucket = getattr(opt, 'bucket', None) or st_dist.GradBucket(net.parameters())   # FusedAdam owns the flat gradient buffer
    global_batch = seeded((FAST_B, 3, FAST_S, FAST_S), 202)
    batch = global_batch[rank * per:(rank + 1) * per].to(dev)
    last = [None]
    train_closure = net.training_closure(batch, loss_net, bucket, graph=not args.no_graph, optimizer=opt)

    def closure():
        total = train_closure()
        last[0] = total.detach()
        return total

    def step():
        opt.step(closure)

    for _ in range(max(args.warmup, 3)):
        step()
    n0 = L.st_launch_count()
    ms = timed(step, args.steps)
    launches = L.st_launch_count() - n0
    if hasattr(train_closure, 'launches'):
        launches = train_closure.launches * args.steps       # kernels inside the replayed graph (Adam included)
    host_batch = batch.detach().cpu().pin_memory()
    e2e_step = HostFedStep(host_batch, batch, step, lambda: last[0])
    for _ in range(2):
        e2e_step()
    ms_e2e = timed(e2e_step, args.steps)
    value = FAST_B * args.steps / (ms * 1e-3)
    out = {'metric': 'fast_st_train_images_per_s_256', 'value': value, 'unit': 'img/s', 'ms_per_step': ms / args.steps,
           'scaling': 'strong', 'global_batch': FAST_B, 'per_gpu_batch': per,
           'config': {'workload': WORKLOADS['fast_st'], 'parallelism': f'dp{world}',
                      'allreduce_bytes': bucket.nbytes() if world > 1 else 0,
                      'launch': 'eager' if args.no_graph else 'one cuda graph per step (closure + all-reduce + Adam)'},
           'e2e': {'value': FAST_B * args.steps / (ms_e2e * 1e-3), 'unit': 'img/s',
                   'h2d_bytes_per_step': host_batch.numel() * 4 * world, 'd2h_bytes_per_step': 4 * world,
                   'ms_per_step': ms_e2e / args.steps},
           'gpu_launches': int(launches), 'loss_after': float(last[0]),
           'step_level': _step_tflops_roofline(value, GF_FAST_PER_IMG, 'image')}
    if trace:
        eager = net.training_closure(batch, loss_net, bucket, graph=False, optimizer=opt)
        for _ in range(2):
            eager()
        tr = KernelTrace(ops, L)
        tr.run(eager, min(args.steps, 5))
        out['roofline'], out['kernels'] = tr.tensor_roofline(args.precision, ms / args.steps, 'training step (transform net + VGG loss network, fwd + dgrad)')
        out['roofline']['traffic'], out['roofline']['traffic_source'] = ncu_traffic('fast_st', args.precision)
        out['roofline_hbm'] = tr.hbm_roofline(ms / args.steps)
    return out


INFER_B, INFER_H, INFER_W = 32, 1080, 1920
GF_INFER_PER_FRAME = 637.8


def run_convert_image(args, dev, L, timed):
    """BASELINE configs[3]: fast_st convert-image, ImageTransformNet inference on a batch of 32 synthetic 1080x1920
    frames (network.py:798-832 applied to a batch), frames per second; `value` with the frames resident in HBM, `e2e`
    host-to-host through ImageTransformNet.stylize_host_batch."""
    from oracle import oracle
    from styletransfer_b200 import network, ops
    style = seeded((1, 3, FAST_S, FAST_S), 201).to(dev)
    net = network.ImageTransformNet(style, INFER_B).to(dev)
    net.load_state_dict({k: v.clone() for k, v in oracle.make_transform_params(2, 3).items()})
    host = torch.empty((INFER_B, 3, INFER_H, INFER_W), dtype=torch.float32).pin_memory()
    four = seeded((4, 3, INFER_H, INFER_W), 301)
    for i in range(0, INFER_B, 4):
        host[i:i + 4] = four
    host_out = torch.empty_like(host).pin_memory()
    x = host.to(dev)
    steps = max(2, min(args.steps, 4))

    def step():
        with torch.no_grad():
            net(x)
    for _ in range(2):
        step()
    n0 = L.st_launch_count()
    ms = timed(step, steps)
    launches = L.st_launch_count() - n0

    def e2e_step():
        net.stylize_host_batch(host, host_out)
    e2e_step()
    ms_e2e = timed(e2e_step, steps)
    value = INFER_B * steps / (ms * 1e-3)
    tr = KernelTrace(ops, L)
    tr.run(step, 2)
    roofline, kernels = tr.tensor_roofline(args.precision, ms / steps, 'batch of 32 frames (8 chunks of 4)')
    roofline['traffic'], roofline['traffic_source'] = None, 'not captured for this workload'
    return {'metric': 'fast_st_convert_image_frames_per_s_1080p', 'value': value, 'unit': 'frames/s', 'ms_per_step': ms / steps,
            'steps': steps, 'config': {'workload': WORKLOADS['convert_image'], 'batch': INFER_B, 'chunk_frames': 4, 'parallelism': 'replicas only'},
            'e2e': {'value': INFER_B * steps / (ms_e2e * 1e-3), 'unit': 'frames/s', 'h2d_bytes_per_step': host.numel() * 4,
                    'd2h_bytes_per_step': host_out.numel() * 4, 'ms_per_step': ms_e2e / steps},
            'gpu_launches': int(launches), 'step_level': _step_tflops_roofline(value, GF_INFER_PER_FRAME, 'frame'),
            'roofline': roofline, 'roofline_hbm': tr.hbm_roofline(ms / steps), 'kernels': kernels}


VIDEO_B, VIDEO_S, VIDEO_FRAMES = 8, 512, 8
GF_VIDEO_PER_FRAME = 434.0       # SURVEY.md section 8d (video_st training at 512x512, 6-channel input)


def run_video_st(args, rank, world, dev, L, timed, with_flow=False, trace=True):
    """BASELINE configs[4]: video_st training frames/s on 8-frame synthetic 512x512 clips, global batch of 8 clips
    sharded over the ranks; one step = the video_train closure on one frame batch (network.py:984-1037: recurrent
    6-channel input, style + content + TV + temporal loss, backward) + all-reduce + Adam.  ``with_flow``: the temporal
    term compares against the flow-warped previous frames (extension; the reference's has no optical flow)."""
    from oracle import oracle
    from styletransfer_b200 import dist as st_dist, network, ops
    per = VIDEO_B // world
    style = seeded((1, 3, FAST_S, FAST_S), 201).to(dev)
    net = network.VideoTransformNet(style, per).to(dev)
    net.load_state_dict({k: v.clone() for k, v in oracle.make_transform_params(2, 6).items()})
    loss_net = network.StyleNetwork(style, torch.rand([1, 3, FAST_S, FAST_S]).to(dev),
                                    vgg_weights=network.load_vgg19_prefix_weights(seed=1))
    loss_net.compute_feature_loss = False
    opt = net.get_optimizer()
    bucket = getattr(opt, 'bucket', None) or st_dist.GradBucket(net.parameters())
    frames = [seeded((VIDEO_B, 3, VIDEO_S, VIDEO_S), 400)]
    for t in range(1, VIDEO_FRAMES):
        frames.append(frames[-1] + 0.05 * seeded((VIDEO_B, 3, VIDEO_S, VIDEO_S), 400 + t))
    host_frames = [f[rank * per:(rank + 1) * per].contiguous().pin_memory() for f in frames]
    dev_frames = [f.to(dev) for f in host_frames]
    # fixed buffers of the video_train loop (network.py:905-1069): this frame, the previous frame's content and
    # stylised images as they were before the iteration, and the slot receiving this frame's stylised output
    frame_buf, old_content, old_stylized, next_stylized = (torch.empty_like(dev_frames[0]) for _ in range(4))
    flow = (1.5 * seeded((per, 2, VIDEO_S, VIDEO_S), 450 + rank)).to(dev) if with_flow else None
    kw = dict(temporal_weight=0.8, optimizer=opt, flow=flow)
    train_closure = net.video_training_closure(frame_buf, old_content, old_stylized, next_stylized, loss_net, bucket,
                                               graph=not args.no_graph, **kw)
    state = {'t': 0, 'last': None, 'resident': True}

    def closure():
        total = train_closure()
        state['last'] = total.detach()
        return total

    def step():
        t = state['t'] % VIDEO_FRAMES
        if state['resident']:
            frame_buf.copy_(dev_frames[t])
        if t == 0:                                     # new clip: the previous frame is the frame itself
            old_content.copy_(frame_buf)
            old_stylized.copy_(frame_buf)
        opt.step(closure)
        old_content.copy_(frame_buf)
        old_stylized.copy_(next_stylized)
        state['t'] += 1
    for _ in range(max(args.warmup, 3)):
        step()
    state['t'] = 0
    n0 = L.st_launch_count()
    ms = timed(step, args.steps)
    launches = L.st_launch_count() - n0
    if hasattr(train_closure, 'launches'):
        launches = train_closure.launches * args.steps           # kernels inside the replayed graph (Adam included)
    state['t'], state['resident'] = 0, False
    e2e_step = HostFedStep(host_frames, frame_buf, step, lambda: state['last'])
    for _ in range(2):
        e2e_step()
    state['t'] = 2
    ms_e2e = timed(e2e_step, args.steps)
    value = VIDEO_B * args.steps / (ms * 1e-3)
    out = {'metric': 'video_st_train_frames_per_s_512', 'value': value, 'unit': 'frames/s', 'ms_per_step': ms / args.steps,
           'scaling': 'strong', 'global_batch_clips': VIDEO_B, 'per_gpu_clips': per,
           'config': {'workload': WORKLOADS['video_st'], 'clip_frames': VIDEO_FRAMES,
                      'temporal_loss': 'flow-warped previous frames (flow= extension)' if with_flow else 'as in the reference: no flow warp',
                      'parallelism': f'dp{world}', 'launch': 'eager' if args.no_graph else 'one cuda graph per step (closure + all-reduce + Adam)'},
           'e2e': {'value': VIDEO_B * args.steps / (ms_e2e * 1e-3), 'unit': 'frames/s',
                   'h2d_bytes_per_step': host_frames[0].numel() * 4 * world, 'd2h_bytes_per_step': 4 * world,
                   'ms_per_step': ms_e2e / args.steps},
           'gpu_launches': int(launches), 'loss_after': float(state['last']),
           'step_level': _step_tflops_roofline(value, GF_VIDEO_PER_FRAME, 'frame')}
    if trace:
        state['resident'] = True
        eager = net.video_training_closure(frame_buf, old_content, old_stylized, next_stylized, loss_net, bucket, graph=False, **kw)
        for _ in range(2):
            eager()
        tr = KernelTrace(ops, L)
        tr.run(eager, min(args.steps, 4))
        out['roofline'], out['kernels'] = tr.tensor_roofline(args.precision, ms / args.steps, 'training step on one frame batch')
        out['roofline']['traffic'], out['roofline']['traffic_source'] = None, 'not captured for this workload'
        out['roofline_hbm'] = tr.hbm_roofline(ms / args.steps)
    return out


def run_gatys_train_e2e(dev, steps=300):
    """The public call of gatys_st (network.py:411-458), end to end at 512x512: host images in -> StyleNetwork -> train_gatys
    (steps=300, L-BFGS at torch's defaults) -> stylised image back on the host; wall clock.  Twice: with the CLI's default
    loss weights (on synthetic noise no curvature pair passes torch's y.s > 1e-10 filter, the loop then takes plain gradient
    steps) and with -sw / -cw at 1e5 x the defaults (the history fills)."""
    from styletransfer_b200 import network
    vw = network.load_vgg19_prefix_weights(seed=1)
    content_h = seeded((1, 3, SIZE, SIZE), 100).pin_memory()
    style_h = torch.nn.functional.interpolate(seeded((1, 3, SIZE // 16, SIZE // 16), 111) * 2.0, scale_factor=16, mode='nearest').pin_memory()
    out = {'config': f'StyleNetwork(style, content).train_gatys(style, content, steps={steps}) at 512x512, host tensors in, host tensor out, wall clock '
                     '(includes style / content targets, graph capture, every L-BFGS iteration); style image = 16x nearest-upsampled noise'}
    for label, scale in (('default_weights', 1.0), ('weights_x1e5', LBFGS_LOSS_SCALE)):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        content, style = content_h.to(dev, non_blocking=True), style_h.to(dev, non_blocking=True)
        net = network.StyleNetwork(style, content, vgg_weights=vw)
        image = net.train_gatys(style, content, steps=steps, style_weight=scale * STYLE_WEIGHT, content_weight=scale * CONTENT_WEIGHT)
        host = image.detach().cpu()
        dt = time.perf_counter() - t0
        stats = dict(getattr(net, 'last_train_stats', {}))
        n = stats.get('closures', 0)
        out[label] = {'wall_s': dt, 'steps': steps, 'closures': n, 'closures_per_s': n / dt, 'history_pairs': stats.get('history_pairs'),
                      'style_weight': scale * STYLE_WEIGHT, 'content_weight': scale * CONTENT_WEIGHT,
                      'image_moved_rel_l2': float((host - content_h).norm() / content_h.norm()),
                      'h2d_bytes': 2 * content_h.numel() * 4, 'd2h_bytes': host.numel() * 4}
        del net, image
        torch.cuda.empty_cache()
    return out


# ---------------------------------------------------------------------------------------------------------------
# The reference arm and the CPU baselines (the only place that runs oracle/ code: oracle/ref_bench.py)
# ---------------------------------------------------------------------------------------------------------------
def hide_gpus():
    """The reference moves itself to CUDA when it sees one (constants.py:20): its CPU path is what is timed here."""
    os.environ['CUDA_VISIBLE_DEVICES'] = ''


def cpu_baselines_child(timeout_s=420):
    """Runs ``bench.py --impl reference --cpu-baselines`` in a child process with the GPUs hidden (the unmodified
    reference decides its device at import time) and returns its {workload: cpu_baseline} object."""
    env = dict(os.environ, CUDA_VISIBLE_DEVICES='')
    for k in ('RANK', 'WORLD_SIZE', 'LOCAL_RANK', 'MASTER_ADDR', 'MASTER_PORT'):
        env.pop(k, None)
    try:
        res = subprocess.run([sys.executable, os.path.abspath(__file__), '--impl', 'reference', '--cpu-baselines'],
                             capture_output=True, text=True, timeout=timeout_s, env=env, cwd=ROOT)
        lines = [ln for ln in res.stdout.splitlines() if ln.startswith('{')]
        return json.loads(lines[-1])
    except (subprocess.TimeoutExpired, IndexError, ValueError, OSError) as e:
        return {'error': f'{type(e).__name__}: {e}'[:300]}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the headline path on the host cores (rank 0 only)."""
    if rank != 0:
        return
    hide_gpus()
    from oracle import ref_bench
    torch.set_num_threads(os.cpu_count() or 1)
    if args.cpu_baselines:
        print(json.dumps(ref_bench.all_baselines()), file=_REAL_STDOUT, flush=True)
        return
    ref = ref_bench.load()
    base, n = ref_bench.gatys(ref, SIZE, budget_s=150.0, max_steps=args.steps, warmup=min(max(args.warmup, 1), 2))
    rate = base['value']
    line = {'impl': 'reference', 'metric': METRIC, 'value': rate, 'unit': 'it/s',
            'n_gpus': args.gpus, 'steps': n, 'warmup': min(max(args.warmup, 1), 2), 'ms_per_step': 1000.0 / rate,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': WORKLOADS['gatys512']},
            'requested_steps': args.steps, 'note': 'steps = closures actually timed (bounded to ~150 s of CPU work)',
            'cpu_baseline': base,
            'e2e': {'value': rate, 'unit': 'it/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)


def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device -- the product path has no CPU fallback')
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    from styletransfer_b200 import _lib, dist as st_dist, network, ops
    L = _lib.lib()
    if args.precision is None:
        args.precision = ops.PRECISION
    ops.set_precision(args.precision)

    # one independent content/style pair per rank (BASELINE configs[1]: no communication)
    content = seeded((1, 3, SIZE, SIZE), 100 + 10 * rank).to(dev)
    style = seeded((1, 3, SIZE, SIZE), 101 + 10 * rank).to(dev)
    net = network.StyleNetwork(style, content, vgg_weights=network.load_vgg19_prefix_weights(seed=1))
    net.compute_feature_loss = False           # train_gatys never reads the feature node
    x = content.clone().requires_grad_(True)

    # the closure train_gatys runs (network.py:438-451): captured once into a CUDA graph and
    # replayed per step unless --no-graph
    closure = net.gatys_closure(x, content, STYLE_WEIGHT, CONTENT_WEIGHT, graph=not args.no_graph)
    eager_closure = net.gatys_closure(x, content, STYLE_WEIGHT, CONTENT_WEIGHT, graph=False)

    def barrier():
        if dist.is_initialized() and not st_dist.forced_single():
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if dist.is_initialized() and not st_dist.forced_single():
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t)
        return ms

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1))

    for _ in range(max(args.warmup, 3)):
        closure()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.busy_until(2, closure, 4.0)           # the same closure, untimed, until nvidia-smi reports
    n0 = L.st_launch_count()
    ms = timed(closure, args.steps)
    launches = L.st_launch_count() - n0
    if not args.no_graph:
        launches = closure.launches * args.steps      # kernels inside the replayed graph
    if sampler:
        sampler.busy_until(sampler.lines() + 2, closure, 2.0)
    clocks = sampler.stop() if sampler else None
    if clocks is not None:
        clocks['sampling'] = 'nvidia-smi every 100 ms under the continuous closure load around the timed steps'

    # ---- end to end: host image in pinned memory -> closure -> loss back on the host ----
    host_img = content.detach().cpu().pin_memory()
    e2e_last = [None]

    def e2e_run():
        e2e_last[0] = closure()
    e2e_step = HostFedStep(host_img, x, e2e_run, lambda: e2e_last[0])
    for _ in range(3):
        e2e_step()
    ms_e2e = timed(e2e_step, args.steps)

    # ---- roofline of the dominant kernel: CUDA events around every tcgen05 conv launch of the eager closure ----
    for _ in range(2):
        eager_closure()          # untimed: the eager path's buffers come out of cudaMalloc on first use
    tr = KernelTrace(ops, L)
    tr.run(eager_closure, args.steps)
    roofline, per_kernel = tr.tensor_roofline(args.precision, ms / args.steps, 'closure: 4 fwd + 4 dgrad')
    roofline['traffic'], roofline['traffic_source'] = ncu_traffic('gatys512', args.precision)

    lbfgs = train_e2e = None
    if world == 1 and not args.no_secondary:
        blocky = torch.nn.functional.interpolate(seeded((1, 3, SIZE // 16, SIZE // 16), 111) * 2.0, scale_factor=16, mode='nearest').to(dev)
        lbfgs_net = network.StyleNetwork(blocky, content, vgg_weights=network.load_vgg19_prefix_weights(seed=1))
        lbfgs_net.compute_feature_loss = False
        lbfgs = run_lbfgs_loop(lbfgs_net, content, dev, L)
        del lbfgs_net
    # free the Gatys working set, then the other configs (all ranks take part in the data-parallel ones)
    del net, closure, eager_closure
    torch.cuda.empty_cache()
    if world == 1 and not args.no_secondary:
        train_e2e = run_gatys_train_e2e(dev)
    fast = video = video_flow = convert = None
    single = {}
    if not args.no_fast_st:
        fast = run_fast_st(args, rank, world, dev, L, timed)
    if not args.no_secondary:
        torch.cuda.empty_cache()
        video = run_video_st(args, rank, world, dev, L, timed)
        torch.cuda.empty_cache()
        video_flow = run_video_st(args, rank, world, dev, L, timed, with_flow=True, trace=False)
        if world == 1:
            torch.cuda.empty_cache()
            convert = run_convert_image(args, dev, L, timed)
    if world > 1:
        # strong-scaling efficiency against ONE GPU running the same global batch: rank 0 measures it here, in this
        # process, with the collective plumbing switched off; the other ranks wait
        dist.barrier()
        if rank == 0:
            with st_dist.single_process():
                torch.cuda.empty_cache()
                if fast is not None:
                    single['fast_st'] = run_fast_st(args, 0, 1, dev, L, timed, trace=False)
                if video is not None:
                    torch.cuda.empty_cache()
                    single['video_st'] = run_video_st(args, 0, 1, dev, L, timed, trace=False)
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    for name, obj in (('fast_st', fast), ('video_st', video)):
        if obj is not None and name in single:
            one = single[name]
            obj['single_gpu_same_global_batch'] = {'value': one['value'], 'e2e': one['e2e']['value'], 'unit': obj['unit']}
            obj['efficiency_vs_single_gpu'] = {'n_gpus': world, 'value': obj['value'] / (world * one['value']),
                                               'e2e': obj['e2e']['value'] / (world * one['e2e']['value'])}
    value = world * args.steps / (ms * 1e-3)
    e2e_value = world * args.steps / (ms_e2e * 1e-3)
    line = {'metric': METRIC, 'value': value, 'unit': 'it/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': ms / args.steps,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': DTYPES[args.precision], 'precision_mode': args.precision, 'dtype_note': DTYPE_NOTES[args.precision],
            'data': 'synthetic',
            'config': {'workload': WORKLOADS['gatys512'],
                       'size': SIZE, 'pairs': world, 'parallelism': f'{world} independent pairs, no collective',
                       'l2': 'working set ~1 GB per closure >> 126 MB L2 (no flush needed)',
                       'launch': 'eager' if args.no_graph else 'cuda graph replay',
                       'algorithmic_gflop_per_step': GF_CLOSURE},
            'tflops_algorithmic': value / world * GF_CLOSURE / 1e3,
            'e2e': {'value': e2e_value, 'unit': 'it/s', 'h2d_bytes_per_step': host_img.numel() * 4, 'd2h_bytes_per_step': 4,
                    'ms_per_step': ms_e2e / args.steps},
            'gpu_launches': int(launches), 'clocks': clocks, 'roofline': roofline, 'kernels': per_kernel}
    for key, obj in (('fast_st', fast), ('video_st', video), ('video_st_flow', video_flow), ('convert_image', convert),
                     ('gatys_lbfgs_loop', lbfgs), ('gatys_train_e2e', train_e2e)):
        if obj is not None:
            line[key] = obj
    if world == 1 and not args.no_cpu_baseline:
        torch.cuda.synchronize()
        base = cpu_baselines_child()
        if 'gatys512' in base:
            line['cpu_baseline'] = base['gatys512']
            line['cpu_baseline_gatys256'] = dict(base['gatys256'], workload=WORKLOADS['gatys256'])
            for key in ('fast_st', 'video_st', 'convert_image'):
                if line.get(key) is not None:
                    line[key]['cpu_baseline'] = base[key]
        else:
            line['cpu_baseline'] = {'value': None, 'unit': 'it/s', 'cores': 0, 'kind': 'reference', 'sample': base.get('error', 'failed')}
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--precision', default=None, choices=['f16x3', 'bf16x6', 'bf16x3', 'bf16'],
                    help='default: styletransfer_b200.ops.PRECISION')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-graph', action='store_true', help='launch the closure eagerly instead of replaying a CUDA graph')
    ap.add_argument('--no-secondary', action='store_true', help='skip the video_st / convert-image / L-BFGS loop metrics')
    ap.add_argument('--no-fast-st', action='store_true', help='skip the secondary fast_st training metric')
    ap.add_argument('--cpu-baselines', action='store_true',
                    help='with --impl reference: print the bounded CPU baselines of every config as one JSON object')
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local_rank = int(os.environ.get('LOCAL_RANK', 0))
    # stdout carries exactly ONE JSON line: whatever libraries print there (NCCL's version banner under
    # NCCL_DEBUG=VERSION, ...) goes to stderr; print() below writes to the saved descriptor
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), 'w')
    os.dup2(2, 1)
    if args.impl == 'reference':
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == '__main__':
    main()
