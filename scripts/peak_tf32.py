#!/usr/bin/env python
"""Dense TF32 and bf16 matmul rate by the method of MEASURED_PEAKS.json (torch.matmul 8192^3, best of 10 and 4 s back to
back), so that BASELINE.md section 2 can state the TF32 peak it promised next to the driver-written bf16 figure."""
import json
import time

import torch

N = 8192
out = {}
for name, dtype, tf32 in (('bf16', torch.bfloat16, False), ('tf32', torch.float32, True)):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    a = torch.randn(N, N, device='cuda', dtype=dtype)
    b = torch.randn(N, N, device='cuda', dtype=dtype)
    for _ in range(3):
        a @ b
    best = 1e9
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        a @ b
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    t0, n = time.time(), 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    while time.time() - t0 < 4.0:
        for _ in range(20):
            a @ b
        n += 20
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    out[name + '_tflops'] = 2 * N ** 3 / best / 1e9
    out[name + '_tflops_sustained'] = 2 * N ** 3 * n / e0.elapsed_time(e1) / 1e9
print(json.dumps(out))
