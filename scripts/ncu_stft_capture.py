#!/usr/bin/env python
"""The program profiled by the STFT / iSTFT capture: one ConvSTFT and one ConviSTFT call (FFT engine) at a batch whose
tensors exceed L2 (default 256 x 4 s: 338 MB spectrogram), between cudaProfilerStart / Stop.

  ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:fft_kernel \
      -o /tmp/stft_full -f python scripts/ncu_stft_capture.py --batch 256"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import speech_enhancement_b200 as se  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--batch', type=int, default=256)
ap.add_argument('--win', type=int, default=400)
ap.add_argument('--hop', type=int, default=100)
ap.add_argument('--fft', type=int, default=512)
a = ap.parse_args()
dev = 'cuda:0'
x = torch.randn(a.batch, 64000, device=dev) * 0.1
st, ist = se.ConvSTFT(a.win, a.hop, a.fft).to(dev), se.ConviSTFT(a.win, a.hop, a.fft).to(dev)
with torch.no_grad():
    for _ in range(3):
        spec = st(x)
        wav = ist(spec)
    torch.cuda.synchronize()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    flush.zero_()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    spec = st(x)
    wav = ist(spec)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
