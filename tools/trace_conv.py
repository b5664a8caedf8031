#!/usr/bin/env python
"""Per-tile timeline of CTA 0 of conv_tc_kernel (resident-weights schedule) on the 64 -> 64 @512x512 forward layer.
Needs a library built with STRANSFER_CONV_TRACE=1 (python -m styletransfer_b200.build); prints cycles relative to the
first event."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from styletransfer_b200 import _lib, ops  # noqa: E402


def main():
    L = _lib.lib()
    dbg = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    L.st_set_conv_path(dbg << 8)
    g = lambda shape, seed, s=1.0: (torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * s).cuda()
    case = sys.argv[2] if len(sys.argv) > 2 else 'fwd64'
    if case == 'fwd64':
        x = ops.from_nchw(g((1, 64, 512, 512), 1))
        pw = ops.PackedWeight(g((64, 64, 3, 3), 2, 0.05), 0, cols_pad=64)
        b = g((64,), 3, 0.1)
        run = lambda: ops.conv2d(x, pw, b, want_out=True, want_relu=True)
    else:
        # Gram backward dF = F * D of a 128-channel tap at 256x256 with an addend (general schedule, 1x1, fused epilogue)
        # cases gram64 / gram128 / gram256; a trailing 'n' drops the addend (plain epilogue), a trailing 'r' makes it the
        # routed gradient of the 2x2 max-pool behind the tap (ops.conv2d_routed)
        C = int(case[4:].rstrip('nr'))
        S = 512 * 64 // C
        f = ops.from_nchw(g((1, C, S, S), 1))
        diff = g((1, C, C), 2, 1e-3)
        amax = diff.abs().max().reshape(1)
        gw = torch.ones(1, device='cuda')
        keep, wp = ops.style_seed(diff, amax, S * S, gw, f.fmt)
        add = ops.from_nchw(g((1, C, S, S), 3), planes=2)
        run = lambda: ops.conv2d(f, None, w_ptrs=wp, w_img_rows=C, add=None if case.endswith('n') else add)
        if case.endswith('r'):
            _, idx = ops.relu_maxpool_fwd(f, want_idx=True)
            gy = ops.from_nchw(g((1, C, S // 2, S // 2), 4), planes=2, fmt=0)
            run = lambda: ops.conv2d_routed(f, wp, C, gy, idx)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    raw = ctypes.CDLL(_lib.LIB_PATH if hasattr(_lib, 'LIB_PATH') else os.path.join(os.path.dirname(_lib.__file__), 'libstransfer_b200.so'))
    buf = (ctypes.c_longlong * (3 * 64 * 8))()
    assert raw.st_conv_trace_read(buf) == 0
    t = torch.tensor(list(buf)).view(3, 64, 8)
    t0 = int(t[t > 0].min())
    names = {0: ['slot for hi', 'slot for lo'], 1: ['stage free', 'hi landed', 'hi issued', 'lo landed', 'lo issued'], 2: ['acc ready', 'acc released', 'stores issued']}
    for tile in range(int((t[1, :, 0] > 0).sum()) or 14):
        line = f'tile {tile:2d}:'
        for role in range(3):
            line += '  |' + ' '.join(f'{names[role][e]} {int(t[role, tile, e]) - t0:6d}' for e in range(len(names[role])))
        print(line)


if __name__ == '__main__':
    main()
