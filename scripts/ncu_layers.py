#!/usr/bin/env python
"""Labelled per-launch table of one profiled forward: joins an `ncu --set full` report of
scripts/ncu_conv_capture.py with that program's stderr log (layer announcements + SE_GEMM_DEBUG=4 plans).

  python scripts/ncu_layers.py gpurun_out/r2_gemm_full.ncu-rep gpurun_out/r2_gemm_capture.log \
      [--traffic-json profiles/r2_conv_traffic.json --model dccrn] > profiles/r2_gemm_full_summary.txt

Per row: layer, engine plan (TW, nacc, concat / split), time, tensor-pipe %, SM %, DRAM read + write of the launch,
the layer's algorithmic minimum (fp32 input + output + packed weight once, SURVEY.md 8a "min MB", x batch) and the
ratio of the two.  --traffic-json writes the sum over the conv launches, which bench.py prints as roofline.traffic."""
import argparse
import csv
import io
import json
import re
import subprocess
import sys

M = {'time': 'gpu__time_duration.sum',
     'tensor': 'TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed',
     'hmma': 'TPC.TriageCompute.sm__pipe_tensor_subpipe_hmma_cycles_active_realtime.avg',
     'memtensor': 'sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',
     'tensor2': 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',
     'tensor3': 'sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active',
     'sm': 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
     'dram': 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
     'rd': 'dram__bytes_read.sum', 'wr': 'dram__bytes_write.sum',
     'regs': 'launch__registers_per_thread', 'grid': 'launch__grid_size',
     'smclk': 'sm__cycles_elapsed.avg.per_second', 'cyc': 'sm__cycles_elapsed.max'}
UNIT_SCALE = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'ns': 1e-3, 'us': 1.0, 'usecond': 1.0,
              'msecond': 1e3, 'ms': 1e3, 'nsecond': 1e-3, 'second': 1e6}


def read_report(path):
    if path.endswith('.csv'):                    # `ncu -i x.ncu-rep --page raw --csv` exported on the GPU box
        raw = open(path).read()
    else:
        raw = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(io.StringIO(raw)) if len(r) > 5]
    hdr, units, data = rows[0], rows[1], rows[2:]
    kn = hdr.index('Kernel Name')
    out = []
    for r in data:
        d = {'kernel': r[kn]}
        for k, name in M.items():
            if name in hdr:
                i = hdr.index(name)
                try:
                    v = float(r[i].replace(',', ''))
                except ValueError:
                    continue
                d[k] = v * UNIT_SCALE.get(units[i], 1.0)
        out.append(d)
    return out


def read_log(path):
    """expected launches of the profiled step, in order: dicts {label, kernel (regex), plan, alg_bytes}"""
    lines = open(path).read().splitlines()
    start = max(i for i, l in enumerate(lines) if l.startswith('=== PROFILED STEP'))
    head = lines[start]
    exp, cur, in_linear = [], None, False
    for l in lines[start + 1:]:
        if l.startswith('=== END'):
            break
        if l.startswith('LAYER '):
            f = dict(kv.split('=') for kv in l.split()[2:] if '=' in kv)
            name = l.split()[1]
            B = int(f['B'])
            kh, kw = (int(v) for v in f['k'].split('x'))
            H, W = (int(v) for v in f['in'].split('x'))
            Ho, Wo = (int(v) for v in f['out'].split('x'))
            cin2, cout2 = int(f['cin2']), int(f['cout2'])
            alg = 4.0 * (B * (cin2 * H * W + cout2 * Ho * Wo) + cin2 * cout2 * kh * kw)
            flop = 2.0 * B * int(f['pix']) * cin2 * kh * kw * cout2
            cur = {'label': name, 'alg': alg, 'flop': flop, 'shape': f'{f["kind"]} {cin2}->{cout2} {kh}x{kw}'}
            if f.get('direct') == '1':
                exp.append(dict(cur, kernel='cconv_first|cconv_small_cin', plan='direct fp32'))
            elif f.get('scatter') == '1':
                exp.append(dict(cur, kernel='cconv_gemm', plan=None, label=name + ' (1x1 scatter GEMM)'))
                exp.append(dict(cur, kernel='col2im', plan='gather', label=name + ' (col2im)', alg=0.0, flop=0.0))
            else:
                exp.append(dict(cur, kernel='cconv_gemm', plan=None))
        elif l.startswith('LINEAR '):
            in_linear = True
        elif l.startswith('LINEAR_END'):
            in_linear = False
        elif l.startswith('se_cconv_gemm:'):
            m = re.search(r'TH=(\d+) TW=(\d+) nacc=(\d+) concat=(\d+) nsplit=(\d+).*tiles=(\d+)', l)
            plan = (f'TH{m.group(1)} TW{m.group(2)} nacc{m.group(3)} {"concat" if m.group(4) == "1" else "split x" + m.group(5)} '
                    f'tiles{m.group(6)}') if m else l[:60]
            if in_linear:
                exp.append({'label': 'lstm/linear projection', 'kernel': 'cconv_gemm', 'plan': plan, 'alg': 0.0,
                            'flop': 0.0, 'shape': 'linear'})
            else:
                for e in exp[::-1]:
                    if e['kernel'] == 'cconv_gemm' and e['plan'] is None:
                        e['plan'] = plan
                        break
    return head, exp


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('report')
    ap.add_argument('log')
    ap.add_argument('--traffic-json')
    ap.add_argument('--model', default='dccrn')
    ap.add_argument('--cublas-ref', help='raw CSV of scripts/ncu_cublas_ref.py under the same ncu settings')
    a = ap.parse_args()
    rows = read_report(a.report)
    head, exp = read_log(a.log)
    batch = int(re.search(r'batch=(\d+)', head).group(1))
    print(f'# {a.report} joined with {a.log}')
    print(f'# {head}')
    print(f'# {len(rows)} profiled launches, {len(exp)} announced (ncu --set full --clock-control none: serialised, '
          'cold-cache replays; times are not bench times)')
    if len(rows) != len(exp):
        print(f'# WARNING: launch count mismatch, rows are matched in order as far as they go')
    ref_rate = None
    if a.cublas_ref:
        ref = read_report(a.cublas_ref)
        rates = [r['hmma'] / r['time'] for r in ref if r.get('hmma') and r.get('time')]
        ref_rate = sum(rates) / len(rates)
        pcts = ', '.join(f'{r.get("tensor", float("nan")):.1f}' for r in ref)
        print(f'# tensor-pipe reference: cuBLAS bf16 8192^3 under the same ncu settings ({a.cublas_ref}): '
              f'{sum(r["time"] for r in ref) / len(ref):.1f} us per launch = '
              f'{2 * 8192 ** 3 / (sum(r["time"] for r in ref) / len(ref)) / 1e6:.0f} TFLOP/s, hmma-subpipe active '
              f'{ref_rate:.0f} cycles/us.  The percent counter sm__pipe_tensor_cycles_active_realtime...pct_of_peak_sustained_elapsed '
              f'read {pcts} % on these identical launches, i.e. it is not reproducible; "vs cuBLAS" below is '
              'hmma-active cycles per microsecond of the launch relative to that cuBLAS run (stable to 0.4 %).')
    print(f'{"layer":28s} {"shape":22s} {"plan":38s} {"time us":>8s} {"tensor%":>7s} {"vs cuBLAS":>9s} {"sm%":>5s} {"dram%":>5s} '
          f'{"DRAM MB":>8s} {"alg MB":>8s} {"ratio":>5s} {"TFLOP/s":>7s} {"regs":>4s}  kernel')
    tot_dram = tot_alg = tot_t = tot_fl = 0.0
    for r, e in zip(rows, exp):
        if not re.search(e['kernel'], r['kernel']):
            print(f'# WARNING: expected {e["kernel"]} for {e["label"]}, report has {r["kernel"][:50]}')
        dram = r.get('rd', 0.0) + r.get('wr', 0.0)
        tensor = r.get('tensor', r.get('tensor2', r.get('tensor3', float('nan'))))
        conv = 'linear' not in e['label']
        if conv:
            tot_dram += dram
            tot_alg += e['alg']
            tot_t += r.get('time', 0.0)
            tot_fl += e['flop']
        ratio = f'{dram / e["alg"]:5.2f}' if e['alg'] else '    -'
        tf = f'{e["flop"] / r["time"] / 1e6:7.1f}' if e['flop'] and r.get('time') else '      -'
        vs = f'{r["hmma"] / r["time"] / ref_rate:9.2f}' if ref_rate and r.get('hmma') and r.get('time') else '        -'
        print(f'{e["label"]:28s} {e["shape"]:22s} {str(e["plan"]):38s} {r.get("time", 0):8.1f} {tensor:7.1f} {vs} {r.get("sm", 0):5.1f} '
              f'{r.get("dram", 0):5.1f} {dram / 1e6:8.1f} {e["alg"] / 1e6:8.1f} {ratio} {tf} {int(r.get("regs", 0)):4d}  '
              f'{r["kernel"][:40]}')
    print(f'# conv launches: DRAM {tot_dram / 1e9:.3f} GB vs algorithmic {tot_alg / 1e9:.3f} GB = {tot_dram / tot_alg:.2f}x; '
          f'{tot_t / 1e3:.3f} ms under ncu, {tot_fl / tot_t / 1e6:.1f} TFLOP/s algorithmic')
    if a.traffic_json:
        try:
            with open(a.traffic_json) as f:
                t = json.load(f)
        except Exception:
            t = {}
        t[a.model] = {'batch': batch, 'dram_bytes_per_step': tot_dram, 'algorithmic_bytes_per_step': tot_alg,
                      'launches': sum(1 for e in exp if 'linear' not in e['label']),
                      'source': f'profiles/{a.report.split("/")[-1].replace(".ncu-rep", "_summary.txt").replace("_raw.csv", "_summary.txt")} (ncu --set full, '
                                'dram__bytes_read.sum + dram__bytes_write.sum summed over the conv launches of one step)'}
        with open(a.traffic_json, 'w') as f:
            json.dump(t, f, indent=1, sort_keys=True)


if __name__ == '__main__':
    main()
