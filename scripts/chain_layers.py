#!/usr/bin/env python
"""Per-layer device time of the CHAINED conv stack inside a model forward (eval / no_grad, CL2 between layers).

  python scripts/chain_layers.py [--model dccrn|dcunet20|frcrn] [--batch B] [--iters N]

Wraps FusedBlock.__call__ / first_layer with CUDA events and prints, per layer: GEMM shape (pixels x K x N),
ms, algorithmic TFLOP/s (2*M*K*N with stride-exact pixel counts) and that as a fraction of the measured
bf16 peak / 3 (three tensor passes per product)."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--model', default='dccrn')
    ap.add_argument('--batch', type=int, default=64)
    ap.add_argument('--iters', type=int, default=5)
    a = ap.parse_args()
    import bench
    from speech_enhancement_b200 import chain
    from speech_enhancement_b200.synth import synth_noisy_speech
    pk = bench.peaks()
    dev = torch.device('cuda:0')
    model, what = bench.build_gpu_model(a.model, dev)
    print(f'# model: {what}')
    x = synth_noisy_speech(a.batch, bench.L, seed=1).to(dev)
    rec = []
    orig_call, orig_first = chain.FusedBlock.__call__, chain.FusedBlock.first_layer

    def timed(fn, kind):
        def wrap(self, *args, **kw):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn(self, *args, **kw)
            e1.record()
            act, y = out
            if kind == 'first':
                B, H, W = args[4], args[5], args[6]
                cin2 = self.cin2
            else:
                srcs = args[0]
                B, H, W = srcs[0].B, max(s.H for s in srcs), max(s.W for s in srcs)
                cin2 = sum(s.C for s in srcs)
            Ho, Wo = (act.H, act.W) if act is not None else (y.shape[2], y.shape[3])
            pix = H * W if self.transposed else Ho * Wo
            flop = 2.0 * pix * B * cin2 * self.kh * self.kw * self.cout2
            rec.append((kind, self.transposed, cin2, self.cout2, self.kh, self.kw, self.sh, self.sw, H, W, Ho, Wo, flop, e0, e1))
            return out
        return wrap

    chain.FusedBlock.__call__ = timed(orig_call, 'gemm')
    chain.FusedBlock.first_layer = timed(orig_first, 'first')
    with torch.no_grad():
        for _ in range(2):
            model(x)
        rec.clear()
        for _ in range(a.iters):
            model(x)
    torch.cuda.synchronize()
    n = len(rec) // a.iters
    tot_ms = tot_fl = 0.0
    print(f'# {a.model} B={a.batch}: {n} conv launches per forward; bf16 peak {pk["bf16_tflops_sustained"]:.0f} TFLOP/s sustained')
    for i in range(n):
        rows = rec[i::n]
        ms = sorted(r[-2].elapsed_time(r[-1]) for r in rows)[len(rows) // 2]
        kind, tr, cin2, cout2, kh, kw, sh, sw, H, W, Ho, Wo, flop = rows[0][:13]
        tf = flop / ms / 1e9
        tot_ms += ms
        tot_fl += flop
        print(f'{i:2d} {"convT" if tr else "conv "} {kind:5s} cin2={cin2:4d} cout2={cout2:4d} k={kh}x{kw} s={sh}x{sw} '
              f'in={H}x{W} out={Ho}x{Wo}  K={cin2 * kh * kw:5d}  {ms:7.3f} ms {tf:7.1f} TFLOP/s  '
              f'{100 * tf * 3 / pk["bf16_tflops_sustained"]:5.1f}% of peak/3')
    print(f'TOTAL {tot_ms:.3f} ms  {tot_fl / tot_ms / 1e9:.1f} TFLOP/s algorithmic')


if __name__ == '__main__':
    main()
