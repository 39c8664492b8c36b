#!/usr/bin/env python
"""The program profiled by the labelled conv capture: ONE forward of a model at the bench batch between
cudaProfilerStart / Stop, with every conv launch announced on stderr (layer name) just before it is issued and
SE_GEMM_DEBUG=4 printing the engine's plan for it.  scripts/ncu_layers.py joins this log with the ncu report.

  SE_GEMM_DEBUG=4 ncu --set full --clock-control none --import-source on --profile-from-start off \
      -k regex:'cconv_gemm_kernel|cconv_first_kernel|cconv_small_cin|col2im_small' -o gpurun_out/r2_gemm_full -f \
      python scripts/ncu_conv_capture.py --model dccrn --batch 64 2> gpurun_out/r2_gemm_capture.log
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--model', default='dccrn')
    ap.add_argument('--batch', type=int, default=64)
    a = ap.parse_args()
    import bench
    from speech_enhancement_b200 import chain
    from speech_enhancement_b200.synth import synth_noisy_speech
    dev = torch.device('cuda:0')
    model, what = bench.build_gpu_model(a.model, dev)
    x = synth_noisy_speech(a.batch, bench.L, seed=1).to(dev)
    view = getattr(model, '_se_chained', model)
    blocks = view._fused_blocks()
    names = {}
    if a.model == 'dcunet20':
        first, enc, dec = blocks
        names[id(first)] = 'first_conv'
    else:
        enc, dec = blocks
    for i, b in enumerate(enc):
        names[id(b)] = f'enc{i}'
    for i, b in enumerate(dec):
        names[id(b)] = f'dec{i}'
    orig_call, orig_first, orig_many = chain.FusedBlock.__call__, chain.FusedBlock.first_layer, chain.linear_many

    def say(msg):
        sys.stderr.write(msg + '\n')
        sys.stderr.flush()

    def call(self, srcs, *args, **kw):
        H, W = max(s.H for s in srcs), max(s.W for s in srcs)
        pix = H * W if self.transposed else None
        Ho, Wo = self.out_size(H, W, kw.get('pad_top', 0), kw.get('pad_left', 0))
        npix = (H * W) if self.transposed else (Ho * Wo)
        say(f'LAYER {names.get(id(self), "conv")} kind={"T" if self.transposed else "C"} cin2={self.cin2} cout2={self.cout2} '
            f'k={self.kh}x{self.kw} pix={npix} B={srcs[0].B} in={H}x{W} out={Ho}x{Wo} scatter={int(self.scatter_form)}')
        return orig_call(self, srcs, *args, **kw)

    def first(self, x_, xs_b, xs_c, xs_h, B, H, W, *args, **kw):
        Ho, Wo = self.out_size(H, W, kw.get('pad_top', 0), kw.get('pad_left', 0))
        say(f'LAYER {names.get(id(self), "conv")} kind=C cin2={self.cin2} cout2={self.cout2} k={self.kh}x{self.kw} '
            f'pix={Ho * Wo} B={B} in={H}x{W} out={Ho}x{Wo} scatter=0 direct=1')
        return orig_first(self, x_, xs_b, xs_c, xs_h, B, H, W, *args, **kw)

    def many(jobs):
        say(f'LINEAR jobs={len(jobs)}')
        out = orig_many(jobs)
        say('LINEAR_END')
        return out

    with torch.no_grad():
        for _ in range(3):
            model(x)
        torch.cuda.synchronize()
        chain.FusedBlock.__call__, chain.FusedBlock.first_layer, chain.linear_many = call, first, many
        import speech_enhancement_b200.modules.complex_aux as aux
        say(f'=== PROFILED STEP model={a.model} batch={a.batch} ({what})')
        torch.cuda.profiler.start()
        model(x)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        say('=== END')


if __name__ == '__main__':
    main()
