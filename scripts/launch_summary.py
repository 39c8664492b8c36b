#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name.
usage: python scripts/launch_summary.py gpurun_out/launches.csv > profiles/<name>.txt"""
import collections
import csv
import re
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith('==')]
    agg, tot = collections.defaultdict(lambda: [0, 0.0]), 0.0
    for row in csv.DictReader(lines):
        if row.get('Metric Name') != 'gpu__time_duration.sum':
            continue
        v = float(row['Metric Value'].replace(',', ''))
        v *= {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 's': 1e6}.get(row['Metric Unit'], 1.0)
        n = re.sub(r'\(.*', '', re.sub(r'<.*', '', row['Kernel Name']))[:80]
        agg[n][0] += 1
        agg[n][1] += v
        tot += v
    print(f'# {path}: {sum(c for c, _ in agg.values())} launches, {tot:.1f} us total device time '
          '(ncu per-launch times are cold-cache and serialised: compare SHARES)')
    print(f'{"us":>12} {"launches":>8} {"share":>7}  kernel')
    for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f'{t:12.1f} {c:8d} {100 * t / tot:6.1f}%  {n}')


if __name__ == '__main__':
    main(sys.argv[1])
