#!/usr/bin/env python
"""The two drop-in paths through the reference's UNMODIFIED model class (oracle/_ref), timed side by side:
module by module (dropin.install() alone) and chained (dropin.accelerate(model)).

  python scripts/dropin_paths.py [--model dccrn|dcunet20|frcrn] [--batch 64] [--iters 10]"""
import argparse
import os
import sys
import warnings

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--model', default='dccrn')
    ap.add_argument('--batch', type=int, default=64)
    ap.add_argument('--iters', type=int, default=10)
    a = ap.parse_args()
    import bench
    from oracle import ref_loader
    from speech_enhancement_b200 import _abi, dropin
    from speech_enhancement_b200.synth import seeded_state_dict, synth_noisy_speech
    if not ref_loader.available():
        raise SystemExit(ref_loader.why_unavailable())
    m = bench.MODELS[a.model]
    dev = torch.device('cuda:0')
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        model = getattr(ref_loader.load_dropin(), m['cls'])(*m['args'])
        model.load_state_dict(seeded_state_dict(model, seed=0))
        model.eval().to(dev)
    x = synth_noisy_speech(a.batch, bench.L, seed=1).to(dev)
    lib = _abi.lib()

    def timed(label):
        with torch.no_grad(), warnings.catch_warnings():
            warnings.simplefilter('ignore')
            for _ in range(3):
                wav = model(x)[1]
            torch.cuda.synchronize()
            n0 = lib.se_launch_count()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.iters):
                wav = model(x)[1]
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.iters
        print(f'{label:46s} {ms:8.2f} ms / step  {a.batch * bench.CLIP_S / ms * 1e3:9.0f} audio-s/s  '
              f'{(lib.se_launch_count() - n0) // a.iters:4d} launches of libse_b200.so per step')
        return wav.clone()

    print(f'# {m["name"]} {a.batch} x 4 s, the reference\'s own class ({type(model).__module__}.{type(model).__name__}) on cuda:0')
    w_mod = timed('module by module (dropin.install())')
    dropin.accelerate(model)
    w_ch = timed('chained (dropin.install() + accelerate(model))')
    err = ((w_ch.double() - w_mod.double()).norm() / w_mod.double().norm()).item()
    print(f'# chained vs module by module: rel-L2 {err:.2e}')


if __name__ == '__main__':
    main()
