"""BASELINE config 5: ConvSTFT -> ConviSTFT round-trip microbench.
win/hop/fft in {400/100/512, 512/128/512}, clip length 1-60 s, batch sweep capped at B*L <= 2^28 samples
(SURVEY.md 8d).  Prints one line per point: ms per round trip, algorithmic GB/s (8d: 4*(L + C2*T) bytes each
way), fraction of the measured HBM copy peak, audio-s/s, and the round-trip error."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import speech_enhancement_b200 as se  # noqa: E402


def hbm_peak():
    try:
        return json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']
    except Exception:
        return 6650.0


def timeit(fn, iters, flush):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--engine', default='auto', choices=['auto', 'simt', 'umma', 'fft'])
    ap.add_argument('--iters', type=int, default=5)
    ap.add_argument('--cap', type=int, default=1 << 26, help='max B*L samples per point')
    a = ap.parse_args()
    dev = 'cuda:0'
    peak = hbm_peak()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    print(f'# engine={a.engine}  HBM copy peak {peak:.0f} GB/s  (L2 flushed between iterations)')
    print('# win/hop/fft      L(s)     B   ms/roundtrip   GB/s(alg)  frac_hbm   audio-s/s   rel-L2')
    for win, hop, fft in ((400, 100, 512), (512, 128, 512)):
        st, ist = se.ConvSTFT(win, hop, fft).to(dev), se.ConviSTFT(win, hop, fft).to(dev)
        st.engine = ist.engine = a.engine
        for L in (16000, 64000, 160000, 480000, 960000):
            for B in (1, 8, 64, 256, 1024):
                if B * L > a.cap:
                    continue
                x = torch.randn(B, L, device=dev) * 0.1
                with torch.no_grad():
                    y = ist(st(x))
                    ms = timeit(lambda: ist(st(x)), a.iters, flush)
                T = (L + fft - win) // hop + 1
                byt = 4 * (L + (fft + 2) * T) + 4 * ((fft + 2) * T + y.shape[1])
                gbs = byt * B / ms / 1e6
                err = ((y - x[:, :y.shape[1]]).norm() / x[:, :y.shape[1]].norm()).item()
                print(f'{win}/{hop}/{fft}  {L / 16000:6.0f} {B:6d} {ms:12.3f} {gbs:12.0f} {gbs / peak:9.3f} '
                      f'{B * L / 16000 / (ms / 1e3):12.0f}   {err:.1e}', flush=True)
                del x, y


if __name__ == '__main__':
    main()
