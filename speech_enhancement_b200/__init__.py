"""B200-native STFT / complex-convolution / mask hot path of shs2783/Speech-Enhancement.

Public surface mirrors the reference's (`models/__init__.py:9-10`): ConvSTFT, ConviSTFT,
ComplexConv2d, ComplexConvTranspose2d (+ the untouched neighbours), the fused mask functions in
`functional`, the caller models in `callers`, and `dropin.install()` which makes the reference's
unmodified model files bind to these classes.
"""
from .conv_stft import ConvSTFT, ConviSTFT, init_kernels
from .modules import (ComplexConv2d, ComplexConvTranspose2d, ComplexBatchNorm2d, ComplexLSTM, ComplexLinear,
                      ComplexPReLU, ComplexReLU, ComplexLeakyReLU, complex_concat, split_complex, merge_real_imag)
from . import functional
from ._abi import invalidate_caches

__all__ = ['ConvSTFT', 'ConviSTFT', 'init_kernels', 'ComplexConv2d', 'ComplexConvTranspose2d',
           'ComplexBatchNorm2d', 'ComplexLSTM', 'ComplexLinear', 'ComplexPReLU', 'ComplexReLU',
           'ComplexLeakyReLU', 'complex_concat', 'split_complex', 'merge_real_imag', 'functional', 'invalidate_caches']
