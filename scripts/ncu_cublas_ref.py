#!/usr/bin/env python
"""Reference point for the tensor-pipe counter: a dense bf16 cuBLAS GEMM (8192^3, the shape MEASURED_PEAKS.json's
bf16 peak was taken on) profiled with the SAME ncu settings as the conv capture.  Whatever
sm__pipe_tensor_cycles_active...pct_of_peak_sustained_elapsed reads for it is the ceiling that counter shows for a
kernel that does nothing but issue MMAs at the rate the power cap allows.

  ncu --set full --clock-control none --profile-from-start off -o /tmp/cublas_ref python scripts/ncu_cublas_ref.py
"""
import torch

a = torch.randn(8192, 8192, device='cuda', dtype=torch.bfloat16)
b = torch.randn(8192, 8192, device='cuda', dtype=torch.bfloat16)
for _ in range(20):            # warm up AND heat up: the conv capture also runs after warm-up steps at the power cap
    c = a @ b
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(3):
    c = a @ b
torch.cuda.synchronize()
torch.cuda.profiler.stop()
