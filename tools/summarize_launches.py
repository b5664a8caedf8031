#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel."""
import collections
import csv
import re
import sys


def main(path, top=22):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10 and r[0].isdigit()]
    tot = collections.defaultdict(lambda: [0, 0])
    for r in rows:
        n = re.sub(r'\(.*', '', r[4]).replace('<unnamed>::', '').replace('void ', '')
        tot[n][0] += int(r[-1]); tot[n][1] += 1
    T = sum(v[0] for v in tot.values())
    print(f'{len(rows)} launches, {T / 1e3:.1f} us of kernel time')
    for n, v in sorted(tot.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f'{n[:72]:72s} {v[1]:4d} {v[0] / 1e3:9.1f} us {100 * v[0] / T:5.1f}%')


if __name__ == '__main__':
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 22)
