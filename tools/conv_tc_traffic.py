#!/usr/bin/env python
"""DRAM traffic per launch of the dominant kernel from an `ncu --set full` capture, tied to the build it profiled.

    ncu -i gpurun_out/X.ncu-rep --page raw --csv > profiles/rNN_conv_tc_ncu_full_raw.csv
    python tools/conv_tc_traffic.py profiles/rNN_conv_tc_ncu_full_raw.csv gatys512 f16x3

Reads the raw-page CSV (one row per profiled launch, a units row under the header), averages
dram__bytes_read.sum + dram__bytes_write.sum over the 3x3 conv_tc launches and writes the entry
``<workload>/<precision>`` of profiles/conv_tc_traffic.json together with the digest of the kernel's sources
(styletransfer_b200.build.CONV_TC_SOURCES; bench.py reports the number only while the digest still matches).  Also writes a compact per-launch summary CSV next
to the input (kernel, duration, DRAM bytes, tensor-pipe activity, achieved occupancy)."""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
UNITS = {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0, 'Tbyte': 1e12}
TIME = {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 's': 1e6, 'usecond': 1.0, 'msecond': 1e3, 'nsecond': 1e-3, 'second': 1e6}
KEEP = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu.sum', 'launch__registers_per_thread', 'launch__grid_size']


def num(s):
    try:
        return float(s.replace(',', ''))
    except ValueError:
        return None


def main(path, workload, precision, pattern=r'conv_tc_kernel<\(int\)3|conv_tc_kernel<3'):
    from styletransfer_b200 import build
    rows = list(csv.reader(open(path)))
    h = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
    hdr, units, body = rows[h], rows[h + 1], rows[h + 2:]
    col = {n: i for i, n in enumerate(hdr)}
    kn = col['Kernel Name']
    ir, iw = col['dram__bytes_read.sum'], col['dram__bytes_write.sum']
    summary = [['kernel'] + [f'{m} [{units[col[m]]}]' for m in KEEP if m in col]]
    traffic = []
    for r in body:
        if len(r) <= max(ir, iw):
            continue
        name = re.sub(r'\(.*', '', r[kn]).replace('<unnamed>::', '').replace('void ', '')
        summary.append([name] + [r[col[m]] for m in KEEP if m in col])
        if re.search(pattern, r[kn]):
            rd, wr = num(r[ir]), num(r[iw])
            if rd is not None and wr is not None:
                traffic.append(rd * UNITS.get(units[ir], 1.0) + wr * UNITS.get(units[iw], 1.0))
    if not traffic:
        sys.exit('no matching launches in ' + path)
    out = os.path.join(ROOT, 'profiles', 'conv_tc_traffic.json')
    table = json.load(open(out)) if os.path.exists(out) else {}
    table[f'{workload}/{precision}'] = {'bytes_per_launch': sum(traffic) / len(traffic), 'launches': len(traffic),
                                        'source': os.path.relpath(os.path.abspath(path), ROOT),
                                        'sources_digest': build.kernel_sources_digest()}
    json.dump(table, open(out, 'w'), indent=1, sort_keys=True)
    spath = re.sub(r'(_raw)?\.csv$', '_summary.csv', path)
    with open(spath, 'w', newline='') as f:
        csv.writer(f).writerows(summary)
    print(json.dumps(table[f'{workload}/{precision}']), '->', out, 'and', spath)


if __name__ == '__main__':
    main(*sys.argv[1:4])
