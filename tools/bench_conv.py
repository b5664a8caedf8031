#!/usr/bin/env python
"""Times single st_conv2d launches (CUDA events, L2 flushed between launches) for the Gatys / fast_st
layer shapes: python tools/bench_conv.py [--reps 20]"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from styletransfer_b200 import ops  # noqa: E402


def seeded(shape, seed, scale=1.0):
    return (torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale).cuda()


INNER = 8


def time_it(fn, reps, flush):
    """median time of one launch, replayed from a CUDA graph (no host overhead inside the events)"""
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            keep = [fn() for _ in range(INNER)]        # event resolution is ~2 us: time several launches per replay
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3 / INNER)
    ts.sort()
    del keep
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--reps', type=int, default=20)
    ap.add_argument('--dbg', type=int, default=0, help='conv_tc experiment switches: 1 no global stores, 2 no MMAs, 4 no activation loads')
    ap.add_argument('--fwd-only', action='store_true')
    args = ap.parse_args()
    from styletransfer_b200 import _lib
    _lib.lib().st_set_conv_path(args.dbg << 8)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
    cases = [('conv1_2', 1, 64, 64, 512), ('conv2_1', 1, 64, 128, 256), ('conv2_2', 1, 128, 128, 256), ('conv3_1', 1, 128, 256, 128),
             ('res', 16, 128, 128, 64)]
    for name, N, cin, cout, S in cases:
        x = ops.from_nchw(seeded((N, cin, S, S), 1))
        w = seeded((cout, cin, 3, 3), 2, 0.05)
        b = seeded((cout,), 3, 0.1)
        pw = ops.PackedWeight(w, 0, cols_pad=cin)
        gf = 2 * 9 * cin * cout * N * S * S / 1e12      # TFLOP
        t1 = time_it(lambda: ops.conv2d(x, pw, b, want_out=True, want_relu=False), args.reps, flush)
        t2 = time_it(lambda: ops.conv2d(x, pw, b, want_out=True, want_relu=True), args.reps, flush)
        if args.fwd_only:
            print(f'{name:8s} N={N} {cin}->{cout} @{S}: fwd out {t1:7.1f} us ({gf / t1 * 1e6:6.1f} TF/s)  fwd out+relu {t2:7.1f} us ({gf / t2 * 1e6:6.1f})', flush=True)
            continue
        # backward-data (bf16 gradient planes) plain and with the ReLU mask of the tap + a two-plane addend
        gy = ops.from_nchw(seeded((N, cout, S, S), 4), planes=2)
        pd = ops.PackedWeight(w, 2, rows_pad=cin)
        tap, add = ops.from_nchw(seeded((N, cin, S, S), 5)), ops.from_nchw(seeded((N, cin, S, S), 6), planes=2)
        t3 = time_it(lambda: ops.conv2d(gy, pd), args.reps, flush)
        t4 = time_it(lambda: ops.conv2d(gy, pd, mask=tap, add=add), args.reps, flush)
        print(f'{name:8s} N={N} {cin}->{cout} @{S}: fwd out {t1:7.1f} us ({gf / t1 * 1e6:6.1f} TF/s)  fwd out+relu {t2:7.1f} us ({gf / t2 * 1e6:6.1f})'
              f'  dgrad plain {t3:7.1f} us ({gf / t3 * 1e6:6.1f})  dgrad mask+add {t4:7.1f} us ({gf / t4 * 1e6:6.1f})', flush=True)


if __name__ == '__main__':
    main()
