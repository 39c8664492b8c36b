"""Fused mask / magnitude-phase ops on libse_b200.so.

Each function takes the four [B, F, T] planes the reference model code works on (noisy real /
imag, mask real / imag -- arbitrary batch and row strides, unit stride along T, so slices of the
STFT output and of the decoder output are read in place) and returns the [B, 2F, T] real|imag
stack that ConviSTFT consumes, i.e. the result of the reference's `_mask_processing` followed by
its `torch.cat([real, imag], 1)`.
"""
import torch

from . import _abi

_DCCRN = {'R': _abi.MASK_DCCRN_R, 'C': _abi.MASK_DCCRN_C, 'E': _abi.MASK_DCCRN_E}
_DCUNET = {'unbounded': _abi.MASK_DCUNET_UNB, 'bounded_sigmoid': _abi.MASK_DCUNET_SIGMOID,
           'bounded_tanh': _abi.MASK_DCUNET_TANH}


def _plane(t, name, B, F, T):
    t = _abi.require_cuda_f32(t, name)
    if t.dim() != 3 or t.shape[0] != B or t.shape[1] != F or t.shape[2] != T:
        raise RuntimeError(f'{name}: expected shape {(B, F, T)}, got {tuple(t.shape)}')
    if T > 1 and t.stride(2) != 1:
        t = t.contiguous()
    return t


def _mask_launch(kind, nr, ni, mr, mi):
    B, F, T = nr.shape
    out = torch.empty((B, 2 * F, T), device=nr.device, dtype=torch.float32)
    args = []
    for t in (nr, ni, mr, mi):
        args += [_abi.ptr(t), t.stride(0), t.stride(1)]
    with torch.cuda.device(nr.device), _abi.stage('mask'):
        _abi.check(_abi.lib().se_mask_fwd(kind, *args, B, F, T, _abi.ptr(out), _abi.stream_ptr()))
    return out


class _MaskFn(torch.autograd.Function):
    """se_mask_fwd / se_mask_bwd as one autograd node (the reference builds ~24 pointwise nodes)."""

    @staticmethod
    def forward(ctx, kind, nr, ni, mr, mi):
        ctx.kind = kind
        ctx.save_for_backward(nr, ni, mr, mi)
        return _mask_launch(kind, nr, ni, mr, mi)

    @staticmethod
    def backward(ctx, g):
        nr, ni, mr, mi = ctx.saved_tensors
        B, F, T = nr.shape
        g = g.contiguous().float()
        d_nr, d_ni = torch.empty((B, F, T), device=g.device), torch.empty((B, F, T), device=g.device)
        d_mr, d_mi = torch.empty(mr.shape, device=g.device), torch.empty(mr.shape, device=g.device)
        args = []
        for t in (nr, ni, mr, mi):
            args += [_abi.ptr(t), t.stride(0), t.stride(1)]
        with torch.cuda.device(g.device), _abi.stage('mask_bwd'):
            _abi.check(_abi.lib().se_mask_bwd(ctx.kind, *args, B, F, T, _abi.ptr(g), _abi.ptr(d_nr), _abi.ptr(d_ni),
                                              _abi.ptr(d_mr), _abi.ptr(d_mi), _abi.stream_ptr()))
        return None, d_nr, d_ni, d_mr, d_mi


def _mask(kind, nr, ni, mr, mi, mask_rows=None):
    nr = _abi.require_cuda_f32(nr, 'noisy_real')
    if nr.dim() != 3:
        raise RuntimeError(f'noisy_real: expected [B, F, T], got {tuple(nr.shape)}')
    B, F, T = nr.shape
    Fm = F if mask_rows is None else mask_rows
    nr, ni = _plane(nr, 'noisy_real', B, F, T), _plane(ni, 'noisy_imag', B, F, T)
    mr, mi = _plane(mr, 'mask_real', B, Fm, T), _plane(mi, 'mask_imag', B, Fm, T)
    if torch.is_grad_enabled() and any(t.requires_grad for t in (nr, ni, mr, mi)):
        return _MaskFn.apply(kind, nr, ni, mr, mi)
    return _mask_launch(kind, nr, ni, mr, mi)


def mask_dccrn(masking, noisy_real, noisy_imag, mask_real, mask_imag):
    """DCCRN 'R' / 'C' / 'E' masks (_2008_00264_dccrn.py:187-207)."""
    return _mask(_DCCRN[masking], noisy_real, noisy_imag, mask_real, mask_imag)


def mask_dcunet(method, noisy_real, noisy_imag, mask_real, mask_imag):
    """DCUNet 'unbounded' / 'bounded_sigmoid' / 'bounded_tanh' (_1903_03107_dcunet.py:158-184)."""
    return _mask(_DCUNET[method], noisy_real, noisy_imag, mask_real, mask_imag)


def mask_carn(noisy_real, noisy_imag, mask_real, mask_imag):
    """CARN complex mask with the reference's sign quirk (_2104_05267_carn.py:165-166)."""
    return _mask(_abi.MASK_CARN, noisy_real, noisy_imag, mask_real, mask_imag)


def mask_frcrn(noisy_real, noisy_imag, mask_real, mask_imag):
    """FRCRN tanh mask (_2206_07293_frcrn.py:141-152): mask planes are the final_conv output
    with F-2 rows; rows 0 and 1 of the result are zero."""
    return _mask(_abi.MASK_FRCRN, noisy_real, noisy_imag, mask_real, mask_imag,
                 mask_rows=noisy_real.shape[1] - 2)


def polar_to_rect(mags, phase):
    """mags*cos(phase) | mags*sin(phase) stacked on dim 1 (conv_stft.py:96-100)."""
    mags = _abi.require_cuda_f32(mags, 'mags').contiguous()
    phase = _abi.require_cuda_f32(phase, 'phase').contiguous()
    if mags.shape != phase.shape or mags.dim() != 3:
        raise RuntimeError(f'polar_to_rect: shapes {tuple(mags.shape)} / {tuple(phase.shape)}')
    B, F, T = mags.shape
    out = torch.empty((B, 2 * F, T), device=mags.device, dtype=torch.float32)
    with torch.cuda.device(mags.device):
        _abi.check(_abi.lib().se_polar_to_rect(_abi.ptr(mags), _abi.ptr(phase), B, F, T, _abi.ptr(out),
                                               _abi.stream_ptr()))
    return out


def pcm16_to_float(x):
    """int16 PCM_S 16 samples -> fp32 in [-1, 1) (torchaudio.load's x / 32768; audio_dataset.py:12-21)."""
    if not (torch.is_tensor(x) and x.is_cuda and x.dtype == torch.int16):
        raise RuntimeError('pcm16_to_float: expected an int16 CUDA tensor')
    x = x.contiguous()
    y = torch.empty(x.shape, device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device):
        _abi.check(_abi.lib().se_pcm16_to_f32(_abi.ptr(x), x.numel(), _abi.ptr(y), _abi.stream_ptr()))
    return y


def pcm16_from_float(x):
    """fp32 waveform -> int16 PCM_S 16 as the reference's savers write it (mix_audio.py:144-146, trainer.py:255-262):
    clamp(-1, 1), x * 32768 rounded to nearest (ties to even), saturated at 32767."""
    x = _abi.require_cuda_f32(x, 'pcm16_from_float(x)').contiguous()
    y = torch.empty(x.shape, device=x.device, dtype=torch.int16)
    with torch.cuda.device(x.device):
        _abi.check(_abi.lib().se_pcm16_from_f32(_abi.ptr(x), x.numel(), _abi.ptr(y), _abi.stream_ptr()))
    return y
