"""ConvSTFT / ConviSTFT on B200: drop-in for /root/reference/models/conv_stft.py.

Same names, constructor and forward signatures, public attributes and state_dict keys
(`weight`; `weight`, `window`, `enframe`) as the reference (conv_stft.py:30, :48, :70, :89);
the arithmetic runs in libse_b200.so.  The kernels consume the module's own `weight` buffer,
so a checkpoint with a different basis is honoured.
"""
import numpy as np
import torch
import torch.nn as nn

from . import _abi


def init_kernels(window_size, fft_size, win_type='hann', inverse=False):
    """Windowed real-DFT basis and window, host side, float64 -> fp32 (conv_stft.py:7-26).

    Returns kernel [fft+2, 1, win] (rows: real bins then imaginary bins; the window sits on
    samples 0..win-1 of the fft-sample frame) and window [1, win, 1].
    """
    from scipy.signal import get_window
    window = get_window(win_type, window_size, fftbins=True)
    n = np.arange(window_size)[None, :]
    f = np.arange(fft_size // 2 + 1)[:, None]
    phase = 2.0 * np.pi * ((f * n) % fft_size) / fft_size
    kernel = np.concatenate([np.cos(phase), -np.sin(phase)], axis=0)
    if inverse:
        kernel = np.linalg.pinv(kernel).T
    kernel = (kernel * window)[:, None, :]
    return (torch.from_numpy(kernel.astype(np.float32)),
            torch.from_numpy(window[None, :, None].astype(np.float32)))


def _no_grad_only(name, *tensors):
    if torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors):
        raise RuntimeError(f'{name}: backward is not implemented in this build; '
                           'call under torch.no_grad() / with inputs that do not require grad')


class ConvSTFT(nn.Module):
    def __init__(self, window_size, hop_size, fft_size=None, win_type='hann', center=True,
                 return_mag_phase=False, fix=True):
        super().__init__()
        self.fft_size = window_size if fft_size is None else fft_size
        kernel, _ = init_kernels(window_size, self.fft_size, win_type)
        self.register_buffer('weight', kernel)
        self.hop_size = hop_size
        self.window_size = window_size
        self.center = center
        self.return_mag_phase = return_mag_phase
        self.pad = self.fft_size // 2

    def forward(self, inputs):
        x = _abi.require_cuda_f32(inputs, 'ConvSTFT.forward(inputs)')
        _no_grad_only('ConvSTFT', x)
        if x.dim() == 1:
            x = x[None, :]
        elif x.dim() == 3:
            if x.shape[1] != 1:
                raise RuntimeError(f'ConvSTFT: expected [B,1,L], got {tuple(x.shape)}')
            x = x[:, 0, :]
        elif x.dim() != 2:
            raise RuntimeError(f'ConvSTFT: expected a 1-, 2- or 3-D tensor, got {x.dim()}-D')
        x = x.contiguous()
        w = _abi.require_cuda_f32(self.weight, 'ConvSTFT.weight').contiguous()
        B, L = x.shape
        c2, _, win = w.shape
        pad = self.pad if self.center else 0
        if pad >= L and self.center:
            raise RuntimeError(f'ConvSTFT: reflect padding ({pad}) must be smaller than the input length ({L})')
        if L + 2 * pad < win:
            raise RuntimeError(f'ConvSTFT: input length {L} is shorter than one window ({win})')
        T = (L + 2 * pad - win) // self.hop_size + 1
        out = torch.empty((B, c2, T), device=x.device, dtype=torch.float32)
        with torch.cuda.device(x.device), _abi.stage('stft'):
            _abi.check(_abi.lib().se_stft_fwd(_abi.ptr(x), B, L, _abi.ptr(w), c2, win, self.hop_size,
                                              self.pad, int(self.center), _abi.ptr(out), T,
                                              _abi.stream_ptr()))
            if not self.return_mag_phase:
                return out
            F = self.fft_size // 2 + 1
            mags = torch.empty((B, F, T), device=x.device, dtype=torch.float32)
            phase = torch.empty_like(mags)
            _abi.check(_abi.lib().se_magphase_fwd(_abi.ptr(out), B, F, T, _abi.ptr(mags), _abi.ptr(phase),
                                                  _abi.stream_ptr()))
        return mags, phase


class ConviSTFT(nn.Module):
    def __init__(self, window_size, hop_size, fft_size=None, win_type='hann', center=True, fix=True):
        super().__init__()
        self.fft_size = window_size if fft_size is None else fft_size
        kernel, window = init_kernels(window_size, self.fft_size, win_type, inverse=True)
        self.register_buffer('weight', kernel)
        self.register_buffer('window', window)
        # kept for state_dict compatibility (conv_stft.py:82); the kernels never read it
        self.register_buffer('enframe', torch.eye(window_size)[:, None, :])
        self.hop_size = hop_size
        self.window_size = window_size
        self.center = center
        self.pad = self.fft_size // 2
        self._packed = None
        self._packed_key = None

    def _packed_basis(self):
        """inverse basis regrouped per hop phase (se_istft_pack_basis); rebuilt when `weight` changes."""
        w = self.weight
        key = (w.data_ptr(), w._version, w.device)
        if self._packed_key != key:
            c2, _, win = w.shape
            J = -(-win // self.hop_size)
            packed = torch.empty((self.hop_size, J * c2), device=w.device, dtype=torch.float32)
            _abi.check(_abi.lib().se_istft_pack_basis(_abi.ptr(w.contiguous()), c2, win, self.hop_size,
                                                      _abi.ptr(packed), _abi.stream_ptr()))
            self._packed, self._packed_key = packed, key
        return self._packed

    def forward(self, inputs, phase=None, output_length=None):
        """inputs: [B, N+2, T] complex spec, or [B, N//2+1, T] mags with `phase` (same shape)."""
        spec = _abi.require_cuda_f32(inputs, 'ConviSTFT.forward(inputs)')
        _no_grad_only('ConviSTFT', spec, phase)
        if spec.dim() != 3:
            raise RuntimeError(f'ConviSTFT: expected [B, C, T], got {tuple(spec.shape)}')
        spec = spec.contiguous()
        w = _abi.require_cuda_f32(self.weight, 'ConviSTFT.weight')
        c2, _, win = w.shape
        B, C, T = spec.shape
        if phase is not None:
            phase = _abi.require_cuda_f32(phase, 'ConviSTFT.forward(phase)').contiguous()
            if phase.shape != spec.shape or 2 * C != c2:
                raise RuntimeError(f'ConviSTFT: mags {tuple(spec.shape)} / phase {tuple(phase.shape)} '
                                   f'do not match a {c2}-row basis')
        elif C != c2:
            raise RuntimeError(f'ConviSTFT: expected {c2} spectrum rows, got {C}')
        full = (T - 1) * self.hop_size + win
        start, end = 0, full
        if self.center:                                  # conv_stft.py:109-111
            start = self.pad
            if output_length is None:
                end = full - self.pad
        if output_length is not None:                    # conv_stft.py:113-114
            end = min(end, start + output_length)
        out_len = max(end - start, 0)
        out = torch.empty((B, out_len), device=spec.device, dtype=torch.float32)
        if out_len == 0 or B == 0:
            return out
        with torch.cuda.device(spec.device), _abi.stage('istft'):
            packed = self._packed_basis()
            window = self.window.reshape(-1).contiguous()
            _abi.check(_abi.lib().se_istft_fwd(_abi.ptr(spec), _abi.ptr(phase), _abi.ptr(packed),
                                               _abi.ptr(window), B, c2, T, win, self.hop_size, start,
                                               out_len, _abi.ptr(out), _abi.stream_ptr()))
        return out
