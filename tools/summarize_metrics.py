#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,
sm__pipe_tensor_cycles_active...,gpu__dram_throughput...` CSV (long format) per kernel name:
launches, total time, DRAM GB/s achieved (read + write bytes / time), tensor-pipe activity."""
import collections
import csv
import re
import sys


def main(path, out=None):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10 and r[0].isdigit()]
    per = collections.defaultdict(dict)          # launch id -> metric -> value
    names = {}
    for r in rows:
        lid = int(r[0])
        names[lid] = re.sub(r'\(.*', '', r[4]).replace('<unnamed>::', '').replace('void ', '')
        unit, val = r[-2], float(r[-1].replace(',', ''))
        scale = {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0, 'ms': 1e6, 'us': 1e3, 'ns': 1.0, 's': 1e9}.get(unit, 1.0)
        per[lid][r[-3]] = val * scale
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0, 0.0])   # n, ns, bytes, tensor*ns, l2 bytes
    for lid, m in per.items():
        a = agg[names[lid]]
        ns = m.get('gpu__time_duration.sum', 0.0)
        a[0] += 1; a[1] += ns
        a[2] += m.get('dram__bytes_read.sum', 0.0) + m.get('dram__bytes_write.sum', 0.0)
        a[3] += m.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 0.0) * ns
        a[4] += m.get('lts__t_bytes.sum', 0.0)
    total = sum(a[1] for a in agg.values())
    lines = ['kernel,launches,time_us,share_pct,dram_GBps,l2_GBps,tensor_pipe_active_pct']
    for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append(f'{n},{a[0]},{a[1] / 1e3:.1f},{100 * a[1] / total:.1f},{a[2] / a[1]:.0f},{a[4] / a[1]:.0f},{a[3] / a[1]:.1f}')
    text = '\n'.join(lines) + '\n'
    if out:
        open(out, 'w').write(text)
    print(text)


if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None)
