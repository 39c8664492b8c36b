"""eager vs CUDA-graph replay of the DCCRN-CL forward at small per-GPU batches"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from speech_enhancement_b200 import callers
from speech_enhancement_b200.graphs import GraphedInference
from speech_enhancement_b200.synth import seeded_state_dict, synth_noisy_speech
dev = 'cuda:0'
model = callers.DCCRN('dccrn-CL', 400, 100, 512).eval()
model.load_state_dict(seeded_state_dict(model, seed=0)); model.to(dev)
g = GraphedInference(model)
def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for B in (1, 4, 16, 64):
    x = synth_noisy_speech(B, 64000, seed=1).to(dev)
    with torch.no_grad():
        t_e = timeit(lambda: model(x))
    t_g = timeit(lambda: g(x))
    print(f'B={B:3d}  eager {t_e:7.3f} ms ({B*4/t_e*1e3:8.0f} audio-s/s)   graph {t_g:7.3f} ms ({B*4/t_g*1e3:8.0f} audio-s/s)', flush=True)
