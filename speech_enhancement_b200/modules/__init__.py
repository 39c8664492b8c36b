from .complex_nn import (ComplexConv2d, ComplexConvTranspose2d, complex_concat, split_complex,
                         merge_real_imag)
from .complex_aux import (ComplexBatchNorm2d, ComplexLSTM, ComplexLinear, ComplexPReLU, ComplexReLU,
                          ComplexLeakyReLU)
