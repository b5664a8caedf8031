#!/usr/bin/env python
"""Benchmark of the stransfer.network hot path on B200 (contract: see the task brief).

Headline metric (BASELINE.json, configs[1]): Gatys VGG-19 forward+backward closures per
second at 512x512, one content/style pair per GPU (N GPUs = N independent pairs, no
collective, weak scaling).  A "step" is one closure of train_gatys
(reference network.py:438-454): zero_grad -> StyleNetwork.forward -> 1e5*style + content
-> backward to the image.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
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


def cpu_reference_closure_rate(budget_s, max_steps, warmup=1):
    """The reference's CPU path (as-executed port, oracle.gatys_closure_as_executed) on the
    host cores: closures/s at 512x512 on a bounded number of closures."""
    from oracle import oracle
    torch.set_num_threads(os.cpu_count() or 1)
    vw = oracle.make_vgg_weights(1)
    content, style = seeded((1, 3, SIZE, SIZE), 100), seeded((1, 3, SIZE, SIZE), 101)
    targets = oracle.style_targets(style, vw)
    x = content.clone()
    t0 = time.perf_counter()
    for _ in range(warmup):
        oracle.gatys_closure_as_executed(x, content, targets, vw, STYLE_WEIGHT, CONTENT_WEIGHT)
    per = (time.perf_counter() - t0) / max(warmup, 1)
    n = int(max(1, min(max_steps, budget_s / max(per, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(n):
        oracle.gatys_closure_as_executed(x, content, targets, vw, STYLE_WEIGHT, CONTENT_WEIGHT)
    dt = time.perf_counter() - t0
    return n / dt, n, torch.get_num_threads()


FAST_B, FAST_S = 16, 256
GF_FAST_PER_IMG = 106.6          # SURVEY.md section 8a row a13 (single-pass, no VGG wgrad)


def cpu_reference_fast_st_rate(budget_s=25.0, b=2):
    """fast_st training closure (oracle.fast_st_closure, torch fp32 CPU) on a bounded sample:
    images per second at 256x256."""
    from oracle import oracle
    torch.set_num_threads(os.cpu_count() or 1)
    vw, params = oracle.make_vgg_weights(1), oracle.make_transform_params(2, 3)
    style, batch = seeded((1, 3, FAST_S, FAST_S), 201), seeded((b, 3, FAST_S, FAST_S), 202)
    targets = oracle.style_targets(style, vw)
    t0 = time.perf_counter()
    oracle.fast_st_closure(batch, params, targets, vw)
    per = time.perf_counter() - t0
    n = int(max(1, min(4, budget_s / max(per, 1e-3) - 1)))
    t0 = time.perf_counter()
    for _ in range(n):
        oracle.fast_st_closure(batch, params, targets, vw)
    dt = time.perf_counter() - t0
    return n * b / dt, n, b, torch.get_num_threads()


def run_fast_st(args, rank, world, dev, L, timed, with_cpu):
    """Secondary metric (BASELINE configs[2]): fast_st training images/s at 256x256, global
    batch 16 sharded over the ranks (strong scaling), one step = the static_train closure
    (network.py:690-731) + flat-bucket all-reduce + Adam step (network.py:765)."""
    from oracle import oracle
    from styletransfer_b200 import dist as st_dist, network
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
    # the closure static_train hands to optimizer.step: a replayed CUDA graph (default) + all-reduce
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
        launches = train_closure.launches * args.steps       # kernels inside the replayed graph
    host_batch = batch.detach().cpu().pin_memory()
    e2e_step = HostFedStep(host_batch, batch, step, lambda: last[0])
    for _ in range(2):
        e2e_step()
    ms_e2e = timed(e2e_step, args.steps)
    value = FAST_B * args.steps / (ms * 1e-3)
    out = {'metric': 'fast_st_train_images_per_s_256', 'value': value, 'unit': 'img/s', 'ms_per_step': ms / args.steps,
           'scaling': 'strong', 'global_batch': FAST_B, 'per_gpu_batch': per,
           'config': {'workload': 'fast_st train 256x256, global batch 16, closure + all-reduce + Adam (BASELINE configs[2])',
                      'parallelism': f'dp{world}', 'allreduce_bytes': bucket.nbytes() if world > 1 else 0},
           'tflops_algorithmic': value * GF_FAST_PER_IMG / 1e3, 'algorithmic_gflop_per_image': GF_FAST_PER_IMG,
           'e2e': {'value': FAST_B * args.steps / (ms_e2e * 1e-3), 'unit': 'img/s',
                   'h2d_bytes_per_step': host_batch.numel() * 4 * world, 'd2h_bytes_per_step': 4 * world,
                   'ms_per_step': ms_e2e / args.steps},
           'gpu_launches': int(launches), 'loss_after': float(last[0])}
    if with_cpu:
        rate, n, b, cores = cpu_reference_fast_st_rate()
        out['cpu_baseline'] = {'value': rate, 'unit': 'img/s', 'cores': cores, 'kind': 'port',
                               'sample': f'{n} closures of batch {b} at 256x256 (oracle.fast_st_closure, torch fp32 CPU, no Adam step)'}
    return out


INFER_B, INFER_H, INFER_W = 32, 1080, 1920
GF_INFER_PER_FRAME = 637.8


def run_convert_image(args, dev, L, timed):
    """Secondary metric (BASELINE configs[3]): fast_st convert-image, ImageTransformNet inference on a batch
    of 32 synthetic 1080x1920 frames (network.py:798-832 applied to a batch), frames per second; `value` with
    the frames resident in HBM, `e2e` host-to-host through ImageTransformNet.stylize_host_batch."""
    from oracle import oracle
    from styletransfer_b200 import network
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
    return {'metric': 'fast_st_convert_image_frames_per_s_1080p', 'value': value, 'unit': 'frames/s', 'ms_per_step': ms / steps,
            'steps': steps, 'config': {'workload': 'fast_st convert-image: ImageTransformNet forward, batch 32 of 1080x1920 '
                                                   '(BASELINE configs[3]), chunks of 4 frames', 'parallelism': 'replicas only'},
            'tflops_algorithmic': value * GF_INFER_PER_FRAME / 1e3, 'algorithmic_gflop_per_frame': GF_INFER_PER_FRAME,
            'e2e': {'value': INFER_B * steps / (ms_e2e * 1e-3), 'unit': 'frames/s', 'h2d_bytes_per_step': host.numel() * 4,
                    'd2h_bytes_per_step': host_out.numel() * 4, 'ms_per_step': ms_e2e / steps},
            'gpu_launches': int(launches)}


VIDEO_B, VIDEO_S, VIDEO_FRAMES = 8, 512, 8
GF_VIDEO_PER_FRAME = 434.0       # SURVEY.md section 8d (video_st training at 512x512, 6-channel input)


def run_video_st(args, rank, world, dev, L, timed):
    """Secondary metric (BASELINE configs[4]): video_st training frames/s on 8-frame synthetic 512x512 clips,
    global batch of 8 clips sharded over the ranks; one step = the video_train closure on one frame batch
    (network.py:984-1037: recurrent 6-channel input, style + content + TV + temporal loss, backward) +
    all-reduce + Adam.  The reference's temporal loss has no optical flow."""
    from oracle import oracle
    from styletransfer_b200 import network
    per = VIDEO_B // world
    style = seeded((1, 3, FAST_S, FAST_S), 201).to(dev)
    net = network.VideoTransformNet(style, per).to(dev)
    net.load_state_dict({k: v.clone() for k, v in oracle.make_transform_params(2, 6).items()})
    loss_net = network.StyleNetwork(style, torch.rand([1, 3, FAST_S, FAST_S]).to(dev),
                                    vgg_weights=network.load_vgg19_prefix_weights(seed=1))
    loss_net.compute_feature_loss = False
    opt = net.get_optimizer()
    from styletransfer_b200 import dist as st_dist
    bucket = getattr(opt, 'bucket', None) or st_dist.GradBucket(net.parameters())
    frames = [seeded((VIDEO_B, 3, VIDEO_S, VIDEO_S), 400)]
    for t in range(1, VIDEO_FRAMES):
        frames.append(frames[-1] + 0.05 * seeded((VIDEO_B, 3, VIDEO_S, VIDEO_S), 400 + t))
    host_frames = [f[rank * per:(rank + 1) * per].contiguous().pin_memory() for f in frames]
    dev_frames = [f.to(dev) for f in host_frames]
    # fixed buffers of the video_train loop (network.py:905-1069): this frame, the previous frame's content and
    # stylised images as they were before the iteration, and the slot receiving this frame's stylised output
    frame_buf, old_content, old_stylized, next_stylized = (torch.empty_like(dev_frames[0]) for _ in range(4))
    train_closure = net.video_training_closure(frame_buf, old_content, old_stylized, next_stylized, loss_net, bucket,
                                               temporal_weight=0.8, graph=not args.no_graph, optimizer=opt)
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
    return {'metric': 'video_st_train_frames_per_s_512', 'value': value, 'unit': 'frames/s', 'ms_per_step': ms / args.steps,
            'scaling': 'strong', 'global_batch_clips': VIDEO_B, 'per_gpu_clips': per,
            'config': {'workload': 'video_st train: 8-frame synthetic 512x512 clips, global batch 8 clips, one frame batch per step, '
                                   'closure + all-reduce + Adam (BASELINE configs[4]; temporal loss without flow warp as in the reference)',
                       'parallelism': f'dp{world}', 'launch': 'eager' if args.no_graph else 'cuda graph replay'},
            'tflops_algorithmic': value * GF_VIDEO_PER_FRAME / 1e3, 'algorithmic_gflop_per_frame': GF_VIDEO_PER_FRAME,
            'e2e': {'value': VIDEO_B * args.steps / (ms_e2e * 1e-3), 'unit': 'frames/s',
                    'h2d_bytes_per_step': host_frames[0].numel() * 4 * world, 'd2h_bytes_per_step': 4 * world,
                    'ms_per_step': ms_e2e / args.steps},
            'gpu_launches': int(launches), 'loss_after': float(state['last'])}


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


def run_reference(args, rank, world):
    if rank != 0:
        return
    rate, n, cores = cpu_reference_closure_rate(budget_s=150.0, max_steps=args.steps, warmup=min(args.warmup, 2) or 1)
    sample = (f'{n} closures of the 512x512 Gatys closure, as executed by the reference (7+2 VGG prefix passes, '
              f'VGG wgrad), fp32 torch CPU port of network.py:438-454 (the Python reference cannot travel to the GPU box)')
    line = {'impl': 'reference', 'metric': 'gatys_vgg19_fwd_bwd_closures_per_s_512', 'value': rate, 'unit': 'it/s',
            'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1000.0 / rate,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': 'gatys_st 512x512 single pair, closure fwd+bwd (configs[1])', 'size': SIZE,
                       'timed_closures': n},
            'cpu_baseline': {'value': rate, 'unit': 'it/s', 'cores': cores, 'kind': 'port', 'sample': sample},
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
    from styletransfer_b200 import _lib, network, ops
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
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
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

    # ---- roofline of the dominant kernel: CUDA events around every tcgen05 conv launch ----
    records = []
    real_conv2d = ops.conv2d

    def traced_conv2d(xs, w, *a, **k):
        ks = w.ks if w is not None else k['w_ptrs'][3]
        cin = xs.C
        # algorithmic output channels (the packed row count is padded, e.g. 3 -> 16 for the image gradient)
        cout = (w.cout if w.mode == 0 else w.cin) if w is not None else k['w_img_rows']
        tc = ks in (1, 3) and k.get('stride', 1) == 1 and not k.get('up') and not k.get('transposed') and cin % 64 == 0
        if not tc:
            return real_conv2d(xs, w, *a, **k)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        # keep the GPU busy while the host queues the launch, so that the events bracket the kernel
        # alone and not the host's launch latency
        torch.cuda._sleep(300000)
        e0.record()
        out = real_conv2d(xs, w, *a, **k)
        e1.record()
        records.append((ks, cin, cout, 2.0 * xs.N * xs.H * xs.W * cin * cout * ks * ks, e0, e1))
        return out
    for _ in range(2):
        eager_closure()          # untimed: the eager path's buffers come out of cudaMalloc on first use
    ops.conv2d = traced_conv2d
    torch.cuda.synchronize()
    for _ in range(args.steps):
        eager_closure()
    torch.cuda.synchronize()
    ops.conv2d = real_conv2d
    # the launches of one closure come in a fixed order: launch i of every step is the same kernel on the
    # same shapes; its duration is the MEDIAN over the steps (a bracket that caught a host stall is dropped)
    per_step = len(records) // max(args.steps, 1)
    groups = {}
    for i in range(per_step):
        ks, cin, cout, fl = records[i][:4]
        t = statistics.median(r[4].elapsed_time(r[5]) for r in records[i::per_step])
        g = groups.setdefault(f'conv_tc_kernel<{ks}x{ks}> {cin}->{cout}', [0.0, 0.0, 0])
        g[0] += fl * args.steps; g[1] += t * args.steps; g[2] += args.steps
    pk = peaks()
    fl3 = sum(g[0] for k, g in groups.items() if '<3x3>' in k)
    ms3 = sum(g[1] for k, g in groups.items() if '<3x3>' in k)
    n3 = sum(g[2] for k, g in groups.items() if '<3x3>' in k)
    passes = {'f16x3': 3, 'bf16x6': 4.5, 'bf16x3': 3, 'bf16': 1}[args.precision]   # fwd 6 / bwd 3 passes in bf16x6
    achieved = fl3 / (ms3 * 1e-3) / 1e12 if ms3 > 0 else 0.0
    traffic, traffic_src = ncu_dram_traffic_per_launch()
    roofline = {'bound': 'tensor', 'kernel': 'conv_tc_kernel<3x3> (9 launches per closure: 4 fwd + 5 dgrad)',
                'achieved': achieved, 'peak': pk['tf_sustained'], 'unit': 'TFLOP/s', 'frac': achieved / pk['tf_sustained'],
                'peak_source': pk['src'] + ' bf16_tflops_sustained', 'traffic': traffic, 'traffic_source': traffic_src,
                'avg_launch_ms': ms3 / max(n3, 1), 'algorithmic_gflop_per_launch': fl3 / max(n3, 1) / 1e9,
                'mma_passes_per_product': passes, 'issued_frac': achieved * passes / pk['tf_sustained'],
                'share_of_step': (ms3 / args.steps) / (ms / args.steps)}
    per_kernel = {k: {'tflops': g[0] / (g[1] * 1e-3) / 1e12, 'ms_per_launch': g[1] / g[2], 'launches_per_step': g[2] / args.steps}
                  for k, g in sorted(groups.items())}

    lbfgs = None
    if world == 1 and not args.no_secondary:
        blocky = torch.nn.functional.interpolate(seeded((1, 3, SIZE // 16, SIZE // 16), 111) * 2.0, scale_factor=16, mode='nearest').to(dev)
        lbfgs_net = network.StyleNetwork(blocky, content, vgg_weights=network.load_vgg19_prefix_weights(seed=1))
        lbfgs_net.compute_feature_loss = False
        lbfgs = run_lbfgs_loop(lbfgs_net, content, dev, L)
        del lbfgs_net
    # free the Gatys working set, then the secondary fast_st training metric (all ranks take part)
    del net, closure, eager_closure
    torch.cuda.empty_cache()
    fast = video = convert = None
    if not args.no_fast_st:
        fast = run_fast_st(args, rank, world, dev, L, timed, with_cpu=(rank == 0 and world == 1 and not args.no_cpu_baseline))
    if not args.no_secondary:
        torch.cuda.empty_cache()
        video = run_video_st(args, rank, world, dev, L, timed)
        if world == 1:
            torch.cuda.empty_cache()
            convert = run_convert_image(args, dev, L, timed)
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    value = world * args.steps / (ms * 1e-3)
    e2e_value = world * args.steps / (ms_e2e * 1e-3)
    line = {'metric': 'gatys_vgg19_fwd_bwd_closures_per_s_512', 'value': value, 'unit': 'it/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': ms / args.steps,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
            'dtype': {'bf16x6': 'bf16x6 (split-bf16 operands: forward 3 planes / 6 tcgen05 passes, backward 2 planes / 3 passes, fp32 accumulate)',
                      'f16x3': 'f16x3 (forward: 2 fp16 planes / 3 passes; backward: 2 planes / 3 passes)',
                      'bf16x3': 'bf16x3 (2 planes / 3 passes)', 'bf16': 'bf16'}[args.precision],
            'data': 'synthetic',
            'config': {'workload': 'gatys_st 512x512 single pair per GPU, closure fwd+bwd (BASELINE configs[1])',
                       'size': SIZE, 'pairs': world, 'parallelism': f'{world} independent pairs, no collective',
                       'l2': 'working set ~1.5 GB per closure >> 126 MB L2 (no flush needed)',
                       'launch': 'eager' if args.no_graph else 'cuda graph replay',
                       'algorithmic_gflop_per_step': GF_CLOSURE},
            'tflops_algorithmic': value / world * GF_CLOSURE / 1e3,
            'e2e': {'value': e2e_value, 'unit': 'it/s', 'h2d_bytes_per_step': host_img.numel() * 4, 'd2h_bytes_per_step': 4,
                    'ms_per_step': ms_e2e / args.steps},
            'gpu_launches': int(launches), 'clocks': clocks, 'roofline': roofline, 'kernels': per_kernel}
    if fast is not None:
        line['fast_st'] = fast
    if video is not None:
        line['video_st'] = video
    if convert is not None:
        line['convert_image'] = convert
    if lbfgs is not None:
        line['gatys_lbfgs_loop'] = lbfgs
    if world == 1 and not args.no_cpu_baseline:
        rate, n, cores = cpu_reference_closure_rate(budget_s=20.0, max_steps=8, warmup=1)
        line['cpu_baseline'] = {'value': rate, 'unit': 'it/s', 'cores': cores, 'kind': 'port',
                                'sample': f'{n} closures at 512x512 of the as-executed reference closure '
                                          f'(oracle.gatys_closure_as_executed, torch fp32 CPU)'}
    print(json.dumps(line), file=_REAL_STDOUT, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


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


def ncu_dram_traffic_per_launch():
    """dram__bytes_read.sum + dram__bytes_write.sum per 3x3 conv_tc launch, averaged over the launches
    of one Gatys closure, from the committed `ncu --set full` capture (profiles/)."""
    import csv
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'profiles', 'r01g_conv_tc_ncu_full_summary.csv')
    try:
        rows = list(csv.reader(open(path)))
        hdr = rows[0]
        ir, iw = hdr.index('dram__bytes_read.sum'), hdr.index('dram__bytes_write.sum')
        unit = {'Mbyte': 1e6, 'Gbyte': 1e9, 'Kbyte': 1e3, 'byte': 1.0}[rows[1][ir]]
        vals = [(float(r[ir]) + float(r[iw])) * unit for r in rows[2:] if r[0].startswith('conv_tc_kernel<3')]
        return (sum(vals) / len(vals), 'profiles/r01g_conv_tc_ncu_full_summary.csv') if vals else (None, None)
    except (OSError, ValueError, KeyError, IndexError):
        return None, None


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
