"""Oracle (test infrastructure): nn.Module wrappers of the CPU restatement + `OPS` namespace.

`OPS` plugs into speech_enhancement_b200.callers.{DCCRN,DCUNet,FRCRN}(ops=OPS) to give the
whole-model CPU oracle / timed CPU baseline.  Buffer and parameter names follow the reference
(conv_stft.py:40, :80-82; complex_nn.py:77-78, :90-91) so one state_dict drives the reference,
the oracle and the CUDA build.
"""
import types

import torch
import torch.nn as nn

from .stft_ref import init_kernels_ref, stft_ref, stft_magphase_ref, istft_ref
from .cconv_ref import cconv2d_ref, cconvT2d_ref
from .mask_ref import mask_dccrn_ref, mask_dcunet_ref, mask_frcrn_ref, mask_carn_ref


class ConvSTFTRef(nn.Module):
    def __init__(self, window_size, hop_size, fft_size=None, win_type='hann', center=True,
                 return_mag_phase=False, fix=True):
        super().__init__()
        self.fft_size = window_size if fft_size is None else fft_size
        self.register_buffer('weight', init_kernels_ref(window_size, self.fft_size, win_type)[0])
        self.hop_size, self.window_size, self.center = hop_size, window_size, center
        self.return_mag_phase, self.pad = return_mag_phase, self.fft_size // 2

    def forward(self, x):
        fn = stft_magphase_ref if self.return_mag_phase else stft_ref
        return fn(x, self.weight, self.hop_size, self.pad, self.center)


class ConviSTFTRef(nn.Module):
    def __init__(self, window_size, hop_size, fft_size=None, win_type='hann', center=True, fix=True):
        super().__init__()
        self.fft_size = window_size if fft_size is None else fft_size
        kernel, window = init_kernels_ref(window_size, self.fft_size, win_type, inverse=True)
        self.register_buffer('weight', kernel)
        self.register_buffer('window', window)
        self.register_buffer('enframe', torch.eye(window_size)[:, None, :])
        self.hop_size, self.window_size, self.center, self.pad = hop_size, window_size, center, self.fft_size // 2

    def forward(self, inputs, phase=None, output_length=None):
        return istft_ref(inputs, self.weight, self.window, self.hop_size, self.pad, self.center,
                         phase, output_length)


class _CConvRef(nn.Module):
    def forward(self, x):
        rc, ic = self.real_conv, self.imag_conv
        if isinstance(rc, nn.ConvTranspose2d):
            return cconvT2d_ref(x, rc.weight, ic.weight, rc.bias, ic.bias, rc.stride, rc.padding,
                                rc.output_padding, rc.dilation)
        return cconv2d_ref(x, rc.weight, ic.weight, rc.bias, ic.bias, rc.stride, rc.padding, rc.dilation)


class ComplexConv2dRef(_CConvRef):
    def __init__(self, in_channels, out_channels, kernel_size, **kwargs):
        super().__init__()
        assert in_channels % 2 == 0 and out_channels % 2 == 0
        self.real_conv = nn.Conv2d(in_channels // 2, out_channels // 2, kernel_size, **kwargs)
        self.imag_conv = nn.Conv2d(in_channels // 2, out_channels // 2, kernel_size, **kwargs)


class ComplexConvTranspose2dRef(_CConvRef):
    def __init__(self, in_channels, out_channels, kernel_size, **kwargs):
        super().__init__()
        assert in_channels % 2 == 0 and out_channels % 2 == 0
        self.real_conv = nn.ConvTranspose2d(in_channels // 2, out_channels // 2, kernel_size, **kwargs)
        self.imag_conv = nn.ConvTranspose2d(in_channels // 2, out_channels // 2, kernel_size, **kwargs)


OPS = types.SimpleNamespace(
    ConvSTFT=ConvSTFTRef, ConviSTFT=ConviSTFTRef,
    ComplexConv2d=ComplexConv2dRef, ComplexConvTranspose2d=ComplexConvTranspose2dRef,
    mask_dccrn=mask_dccrn_ref, mask_dcunet=mask_dcunet_ref, mask_frcrn=mask_frcrn_ref,
    mask_carn=mask_carn_ref)
