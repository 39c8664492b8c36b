"""Oracle (test infrastructure): CPU restatement of the complex 2-D convolutions.

Follows /root/reference/models/modules/complex_nn.py:
  ComplexConv.forward       :52-65  (4 real convs on the channel halves)
  ComplexConv2d             :67-78
  ComplexConvTranspose2d    :80-91
x is [B, 2*Ci, H, W] with the "real" half = first Ci channels (torch.chunk(x, 2, 1)).
"""
import torch
import torch.nn.functional as F


def _combine(x, op_r, op_i):
    xr, xi = torch.chunk(x, 2, dim=1)                       # complex_nn.py:26 / :53
    real = op_r(xr) - op_i(xi)                              # :55-56, :61
    imag = op_i(xr) + op_r(xi)                              # :58-59, :62
    return torch.cat([real, imag], dim=1)                   # :40 / :64


def cconv2d_ref(x, wr, wi, br=None, bi=None, stride=1, padding=0, dilation=1):
    """wr, wi: [Co, Ci, kh, kw] (nn.Conv2d layout).  Bias folds to [br-bi, br+bi]."""
    dt = x.dtype
    cv = lambda w, b: (lambda t: F.conv2d(t, w.to(dt), None if b is None else b.to(dt),
                                          stride=stride, padding=padding, dilation=dilation))
    return _combine(x, cv(wr, br), cv(wi, bi))


def cconvT2d_ref(x, wr, wi, br=None, bi=None, stride=1, padding=0, output_padding=0, dilation=1):
    """wr, wi: [Ci, Co, kh, kw] (nn.ConvTranspose2d layout)."""
    dt = x.dtype
    cv = lambda w, b: (lambda t: F.conv_transpose2d(t, w.to(dt), None if b is None else b.to(dt),
                                                    stride=stride, padding=padding,
                                                    output_padding=output_padding, dilation=dilation))
    return _combine(x, cv(wr, br), cv(wi, bi))


def pack_conv_weight_ref(wr, wi):
    """[[Wr,-Wi],[Wi,Wr]] as one real [2Co, 2Ci, kh, kw] conv weight."""
    top = torch.cat([wr, -wi], dim=1)
    bot = torch.cat([wi, wr], dim=1)
    return torch.cat([top, bot], dim=0)


def pack_convT_weight_ref(wr, wi):
    """[[Wr,Wi],[-Wi,Wr]] in the (in, out, kh, kw) layout: [2Ci, 2Co, kh, kw]."""
    top = torch.cat([wr, wi], dim=1)
    bot = torch.cat([-wi, wr], dim=1)
    return torch.cat([top, bot], dim=0)
