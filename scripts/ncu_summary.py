#!/usr/bin/env python
"""Compact per-launch table from an `ncu --set full` report.
usage: python scripts/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x_summary.txt"""
import csv
import io
import subprocess
import sys

COLS = [
    ('gpu__time_duration.sum', 'time'),
    ('TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed', 'tensor%'),
    ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm%'),
    ('lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l2%'),
    ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram%'),
    ('dram__bytes_read.sum', 'dram_rd'),
    ('dram__bytes_write.sum', 'dram_wr'),
    ('lts__t_sector_hit_rate.pct', 'l2hit%'),
    ('launch__registers_per_thread', 'regs'),
    ('launch__grid_size', 'grid'),
    ('launch__block_size', 'block'),
    ('smsp__inst_executed.sum', 'inst'),
]


def main(path):
    if path.endswith('.csv'):
        raw = open(path).read()
    else:
        raw = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(io.StringIO(raw)) if len(r) > 5]
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {n: hdr.index(n) for n, _ in COLS if n in hdr}
    kn = hdr.index('Kernel Name')
    print(f'# {path}: {len(data)} launches (ncu --set full --clock-control none; serialised, cold-cache replays)')
    print('  '.join(f'{lab}[{units[idx[n]]}]' for n, lab in COLS if n in idx) + '  kernel')
    for r in data:
        cells = []
        for n, lab in COLS:
            if n in idx:
                v = r[idx[n]].replace(',', '')
                try:
                    v = f'{float(v):.4g}'
                except ValueError:
                    pass
                cells.append(v)
        print('  '.join(f'{c:>9}' for c in cells) + '  ' + r[kn][:60])


if __name__ == '__main__':
    main(sys.argv[1])
