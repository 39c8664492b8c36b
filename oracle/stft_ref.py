"""Oracle (test infrastructure): CPU restatement of the conv-STFT front/back end.

Follows /root/reference/models/conv_stft.py:
  init_kernels   :7-26    -> init_kernels_ref
  ConvSTFT.fwd   :48-66   -> stft_ref / stft_magphase_ref
  ConviSTFT.fwd  :89-116  -> istft_ref (ola_norm_ref is the `coff` of :104-105)
The arithmetic of the reference lives in ATen (conv1d / conv_transpose1d), so
the restatement calls the same ATen ops on CPU; dtype follows the inputs so the
same functions serve as the fp64 error-budget reference.
"""
import numpy as np
import torch
import torch.nn.functional as F


def _window(win_type, window_size):
    from scipy.signal import get_window
    return get_window(win_type, window_size, fftbins=True)          # conv_stft.py:10


def init_kernels_ref(window_size, fft_size, win_type='hann', inverse=False):
    """Windowed real-DFT basis [fft+2, 1, win] and window [1, win, 1] (conv_stft.py:7-26).

    Row c < F is w[k]*cos(2*pi*c*k/N); row F+c is -w[k]*sin(2*pi*c*k/N); the window
    covers samples 0..win-1 of the N-sample frame (left aligned, :11).
    """
    n_bins = fft_size // 2 + 1
    k = np.arange(window_size, dtype=np.float64)[None, :]
    c = np.arange(n_bins, dtype=np.float64)[:, None]
    ang = 2.0 * np.pi * c * k / fft_size
    basis = np.concatenate([np.cos(ang), -np.sin(ang)], axis=0)     # [fft+2, win]   (:11-15)
    if inverse:
        basis = np.linalg.pinv(basis).T                              # (:17-18)
    w = _window(win_type, window_size)
    basis = basis * w[None, :]                                       # (:20)
    kernel = torch.from_numpy(basis[:, None, :].astype(np.float32))
    window = torch.from_numpy(w[None, :, None].astype(np.float32))
    return kernel, window


def _as_b1l(x):
    if x.dim() == 1:                                                 # (:49-52)
        return x[None, None, :]
    if x.dim() == 2:
        return x[:, None, :]
    return x


def stft_ref(x, weight, hop, pad, center=True):
    """[B,L] -> [B, fft+2, T]: reflect pad, strided basis projection (conv_stft.py:54-56)."""
    x = _as_b1l(x)
    if center:
        x = F.pad(x, [pad, pad], mode='reflect')
    return F.conv1d(x, weight.to(x.dtype), stride=hop)


def stft_magphase_ref(x, weight, hop, pad, center=True):
    """return_mag_phase=True branch (conv_stft.py:58-64): no epsilon in the sqrt."""
    spec = stft_ref(x, weight, hop, pad, center)
    f = spec.shape[1] // 2
    re, im = spec[:, :f], spec[:, f:]
    return torch.sqrt(re * re + im * im), torch.atan2(im, re)


def ola_norm_ref(window, n_frames, hop):
    """`coff` of conv_stft.py:104-105: overlap-added squared window, [1,1,(T-1)*hop+win]."""
    win = window.reshape(-1)
    t = (win * win)[None, :, None].repeat(1, 1, n_frames)
    eye = torch.eye(win.numel(), dtype=win.dtype)[:, None, :]
    return F.conv_transpose1d(t, eye, stride=hop)


def istft_ref(spec, weight, window, hop, pad, center=True, phase=None, output_length=None):
    """[B, fft+2, T] (or mags + phase) -> [B, L'] (conv_stft.py:89-116)."""
    if phase is not None:                                            # (:96-100)
        spec = torch.cat([spec * torch.cos(phase), spec * torch.sin(phase)], dim=1)
    dt = spec.dtype
    y = F.conv_transpose1d(spec, weight.to(dt), stride=hop)          # (:101)
    coff = ola_norm_ref(window.to(dt), spec.shape[-1], hop)
    y = y / (coff + 1e-8)                                            # (:106)
    if center:                                                       # (:109-111)
        y = y[..., pad:]
        if output_length is None:
            y = y[..., :-pad]
    if output_length is not None:                                    # (:113-114)
        y = y[..., :output_length]
    return y.squeeze(1)
