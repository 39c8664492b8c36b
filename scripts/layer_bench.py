#!/usr/bin/env python
"""Per-layer timing of the complex-conv stack at the BASELINE shapes.

  python scripts/layer_bench.py [--model dccrn|dcunet|frcrn] [--batch B] [--iters N] [--only i,j]

Prints one line per layer: CUDA-event ms per call (whole C-ABI call: operand prep + GEMM),
algorithmic TFLOP/s (2*M*K*N, stride-exact) and the fraction of the measured bf16 peak / 3
(the bf16x3 split issues three tensor passes per product).  Used under ncu for the per-kernel
launch list and the --set full capture committed under profiles/.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

K52 = dict(stride=(2, 1), padding=(2, 0))
K52T = dict(stride=(2, 1), padding=(2, 0), output_padding=(1, 0))
# (name, transposed, cin, cout, kernel, kwargs, H, W)
DCCRN = [('enc0', False, 2, 32, (5, 2), K52, 256, 643), ('enc1', False, 32, 64, (5, 2), K52, 128, 643),
         ('enc2', False, 64, 128, (5, 2), K52, 64, 643), ('enc3', False, 128, 256, (5, 2), K52, 32, 643),
         ('enc4', False, 256, 256, (5, 2), K52, 16, 643), ('enc5', False, 256, 256, (5, 2), K52, 8, 643),
         ('dec0', True, 512, 256, (5, 2), K52T, 4, 642), ('dec1', True, 512, 256, (5, 2), K52T, 8, 642),
         ('dec2', True, 512, 128, (5, 2), K52T, 16, 642), ('dec3', True, 256, 64, (5, 2), K52T, 32, 642),
         ('dec4', True, 128, 32, (5, 2), K52T, 64, 642), ('dec5', True, 64, 2, (5, 2), K52T, 128, 642)]
FR = dict(stride=(2, 1), padding=(0, 0))
FRCRN = [('enc0', False, 2, 128, (5, 2), FR, 320, 404), ('enc1', False, 128, 128, (5, 2), FR, 158, 404),
         ('enc2', False, 128, 128, (5, 2), FR, 77, 404), ('enc3', False, 128, 128, (5, 2), FR, 37, 404),
         ('dec3', True, 256, 128, (5, 2), FR, 37, 403), ('dec4', True, 256, 128, (5, 2), FR, 77, 403),
         ('dec5', True, 256, 128, (5, 2), FR, 158, 403)]
DCUNET = [('first', False, 2, 64, 3, dict(padding=1), 257, 501),
          ('enc0', False, 64, 64, (7, 1), dict(padding=(3, 0)), 257, 501),
          ('enc1', False, 64, 64, (1, 7), dict(padding=(0, 3)), 257, 501),
          ('enc2', False, 64, 128, (7, 5), dict(stride=(2, 2), padding=(3, 2)), 257, 501),
          ('enc3', False, 128, 128, (7, 5), dict(stride=(2, 1), padding=(3, 2)), 129, 251),
          ('enc4', False, 128, 128, (5, 3), dict(stride=(2, 2), padding=(2, 1)), 65, 251),
          ('dec5', True, 256, 128, (5, 3), dict(stride=(2, 2), padding=(2, 1)), 33, 126),
          ('dec6', True, 256, 128, (7, 5), dict(stride=(2, 1), padding=(3, 2)), 65, 251),
          ('dec7', True, 256, 64, (7, 5), dict(stride=(2, 2), padding=(3, 2)), 129, 251),
          ('dec8', True, 128, 64, (1, 7), dict(padding=(0, 3)), 257, 501),
          ('dec9', True, 128, 2, (7, 1), dict(padding=(3, 0)), 257, 501)]
TABLES = {'dccrn': DCCRN, 'frcrn': FRCRN, 'dcunet': DCUNET}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--model', default='dccrn', choices=sorted(TABLES))
    ap.add_argument('--batch', type=int, default=16)
    ap.add_argument('--iters', type=int, default=10)
    ap.add_argument('--only', default='')
    args = ap.parse_args()
    import speech_enhancement_b200 as se
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            peak = json.load(f)['bf16_tflops']
    except Exception:
        peak = 1590.0
    dev = torch.device('cuda:0')
    B = args.batch
    only = set(args.only.split(',')) if args.only else None
    tot_ms = tot_gf = 0.0
    for name, tr, cin, cout, k, kw, H, W in TABLES[args.model]:
        if only and name not in only:
            continue
        m = (se.ComplexConvTranspose2d if tr else se.ComplexConv2d)(cin, cout, k, bias=False, **kw).to(dev)
        x = torch.randn(B, cin, H, W, device=dev)
        with torch.no_grad():
            for _ in range(2):
                y = m(x)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.iters):
                y = m(x)
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.iters
        kh, kwid = (k, k) if isinstance(k, int) else k
        pix = (H * W) if tr else (y.shape[2] * y.shape[3])
        gflop = 2.0 * cout * cin * kh * kwid * pix * B / 1e9
        tf = gflop / ms
        tot_ms += ms
        tot_gf += gflop
        print(f'{args.model} {name:6s} B={B:3d} in=({cin},{H},{W}) out={tuple(y.shape[1:])} {ms:8.3f} ms '
              f'{tf:7.1f} TFLOP/s alg  {100 * tf * 3 / peak:5.1f}% of bf16 peak (x3 passes)', flush=True)
        del m, x, y
    if tot_ms:
        print(f'{args.model} TOTAL B={B} {tot_ms:8.3f} ms {tot_gf / tot_ms:7.1f} TFLOP/s alg '
              f'{100 * tot_gf / tot_ms * 3 / peak:5.1f}% of bf16 peak (x3 passes)')


if __name__ == '__main__':
    main()
