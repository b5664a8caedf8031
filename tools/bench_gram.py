#!/usr/bin/env python
"""Sweep of the Gram kernel's pixel split-K (st_set_gram_max_splits) on the five taps of a 512x512 Gatys closure
and of a batch-16 256x256 fast_st step: microseconds per ops.gram call (gram_tc_kernel + finish)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from styletransfer_b200 import _lib, ops
    L = _lib.lib()
    shapes = [(1, 64, 512, 512), (1, 128, 256, 256), (1, 256, 128, 128), (16, 64, 256, 256), (16, 128, 128, 128), (16, 256, 64, 64)]
    for B, C, H, W in shapes:
        f = ops.from_nchw(torch.randn(B, C, H, W, device='cuda'))          # the forward format of the current precision mode
        row = []
        for cap in (0, 148, 96, 64, 48, 32, 24, 16, 8):
            L.st_set_gram_max_splits(cap)
            for _ in range(5):
                ops.gram(f)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(100):
                ops.gram(f)
            e1.record()
            torch.cuda.synchronize()
            row.append(f'{cap}:{e0.elapsed_time(e1) * 10:.1f}')
        L.st_set_gram_max_splits(0)
        print(f'B{B} C{C} {H}x{W}  us per call  ' + '  '.join(row), flush=True)


if __name__ == '__main__':
    main()
