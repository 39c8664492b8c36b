"""Time ConvSTFT / ConviSTFT (both engines) on the headline framing: win 400, hop 100, fft 512, 4 s @ 16 kHz.
Reports ms per call and the algorithmic HBM rate (STFT 1.58 MB/utt, iSTFT 1.57 MB/utt)."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import speech_enhancement_b200 as se  # noqa: E402


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        ev[0].record()
        fn()
        ev[1].record()
        torch.cuda.synchronize()
        tot += ev[0].elapsed_time(ev[1])
    return tot / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=64)
    ap.add_argument('--len', type=int, default=64000)
    ap.add_argument('--win', type=int, default=400)
    ap.add_argument('--hop', type=int, default=100)
    ap.add_argument('--fft', type=int, default=512)
    a = ap.parse_args()
    dev = 'cuda:0'
    x = torch.randn(a.batch, a.len, device=dev) * 0.1
    for engine in ('simt', 'umma', 'fft'):
        st = se.ConvSTFT(a.win, a.hop, a.fft).to(dev)
        ist = se.ConviSTFT(a.win, a.hop, a.fft).to(dev)
        st.engine = ist.engine = engine
        with torch.no_grad():
            spec = st(x)
            wav = ist(spec)
            t_st = timeit(lambda: st(x))
            t_ist = timeit(lambda: ist(spec))
        b_st = x.numel() * 4 + spec.numel() * 4
        b_ist = spec.numel() * 4 + wav.numel() * 4
        err = ((wav - x[:, :wav.shape[1]]).norm() / x[:, :wav.shape[1]].norm()).item()
        print(f'{engine}: B={a.batch} L={a.len}  stft {t_st:.3f} ms ({b_st / t_st / 1e6:.0f} GB/s)  '
              f'istft {t_ist:.3f} ms ({b_ist / t_ist / 1e6:.0f} GB/s)  round-trip rel-L2 {err:.2e}', flush=True)


if __name__ == '__main__':
    main()
