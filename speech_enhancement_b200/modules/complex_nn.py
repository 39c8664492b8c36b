"""Complex-valued layers on B200: drop-in for /root/reference/models/modules/complex_nn.py.

ComplexConv2d / ComplexConvTranspose2d keep the reference constructor signature
(complex_nn.py:68, :81), the `real_conv` / `imag_conv` parameter holders (so state_dict keys
and utils.initialize_params keep working) and forward(x) semantics (complex_nn.py:52-65), but
run as ONE packed implicit GEMM [[Wr,-Wi],[Wi,Wr]] in libse_b200.so instead of four cuDNN convs
plus chunk/sub/add/cat.  The remaining symbols of the reference module that the model files
import (ComplexBatchNorm2d, ComplexLSTM, ComplexLinear, ComplexPReLU, complex_concat, ...) are
out of the hot-path scope and are provided here on stock PyTorch
so the package is self-contained on a box without the reference checkout.
"""
import torch
import torch.nn as nn

from .. import _abi


# --------------------------------------------------------------------------- layout helpers
def split_complex(x, dim=1):
    """real / imaginary parts of a (real, imag) pair, a complex tensor or a stacked real tensor."""
    if isinstance(x, (tuple, list)):
        return x[0], x[1]
    if isinstance(x, torch.Tensor):
        if torch.is_complex(x):
            return x.real, x.imag
        return torch.chunk(x, 2, dim=dim)
    raise ValueError("Input must be a complex tensor or a tuple of real and imaginary tensors")


def merge_real_imag(x, real, imag, dim=1):
    """re-assemble in the container kind of `x` (list, complex tensor or stacked real tensor)."""
    if isinstance(x, (tuple, list)):
        return [real, imag]
    if torch.is_complex(x):
        return torch.complex(real, imag)
    return torch.cat([real, imag], dim=dim)


def complex_concat(inputs, dim=1):
    """concatenate stacked-complex tensors keeping all real halves first (complex_nn.py:4-16)."""
    halves = [torch.chunk(t, 2, dim=dim) for t in inputs]
    return torch.cat([h[0] for h in halves] + [h[1] for h in halves], dim=dim)


def _pair(v):
    return (v, v) if isinstance(v, int) else tuple(v)


# --------------------------------------------------------------------------- the two hot layers
class ComplexConv(nn.Module):
    """Shared forward of the two complex conv layers (complex_nn.py:45-65)."""
    _transposed = False
    engine = _abi.ENGINE_AUTO          # class-level knob: ENGINE_AUTO / ENGINE_SIMT / ENGINE_UMMA

    def __init__(self):
        super().__init__()
        self.real_conv = None
        self.imag_conv = None
        self._pack = None
        self._pack_key = None

    # packed operand cache: rebuilt after optimizer.step() / load_state_dict / re-init
    def _packed(self):
        rc, ic = self.real_conv, self.imag_conv
        ts = [rc.weight, ic.weight] + ([rc.bias, ic.bias] if rc.bias is not None else [])
        key = tuple((t.data_ptr(), t._version) for t in ts) + (rc.weight.device,)
        if self._pack_key != key:
            wr = _abi.require_cuda_f32(rc.weight.detach(), 'real_conv.weight').contiguous()
            wi = _abi.require_cuda_f32(ic.weight.detach(), 'imag_conv.weight').contiguous()
            kh, kw = wr.shape[2], wr.shape[3]
            if self._transposed:
                ci, co = wr.shape[0], wr.shape[1]
            else:
                co, ci = wr.shape[0], wr.shape[1]
            packed = torch.empty((2 * co, 2 * ci * kh * kw), device=wr.device, dtype=torch.float32)
            bias = None
            br = bi = None
            if rc.bias is not None:
                br, bi = rc.bias.detach().contiguous(), ic.bias.detach().contiguous()
                bias = torch.empty((2 * co,), device=wr.device, dtype=torch.float32)
            _abi.check(_abi.lib().se_cconv_pack_weight(
                _abi.ptr(wr), _abi.ptr(wi), _abi.ptr(br), _abi.ptr(bi), co, ci, kh, kw,
                int(self._transposed), _abi.ptr(packed), _abi.ptr(bias), _abi.stream_ptr()))
            self._pack, self._pack_key = (packed, bias, co, ci, kh, kw), key
        return self._pack

    def _check_supported(self):
        rc = self.real_conv
        if rc.groups != 1 or _pair(rc.dilation) != (1, 1):
            raise RuntimeError(f'{type(self).__name__}: groups != 1 / dilation != 1 are not supported '
                               'by the B200 kernels')
        if getattr(rc, 'padding_mode', 'zeros') != 'zeros' or isinstance(rc.padding, str):
            raise RuntimeError(f'{type(self).__name__}: only numeric zero padding is supported')

    def forward(self, x):
        real, imag = split_complex(x)
        if isinstance(x, torch.Tensor) and not torch.is_complex(x):
            xin = x
        else:
            xin = torch.cat([real, imag], dim=1)       # tuple / complex forms: convert at the edge
        y = self._forward_stacked(xin)
        if xin is x:
            return y
        yr, yi = torch.chunk(y, 2, dim=1)
        return merge_real_imag(x, yr, yi)

    def _forward_stacked(self, x):
        x = _abi.require_cuda_f32(x, f'{type(self).__name__}.forward(x)')
        if torch.is_grad_enabled() and (x.requires_grad or self.real_conv.weight.requires_grad):
            raise RuntimeError(f'{type(self).__name__}: backward is not implemented in this build; '
                               'call under torch.no_grad()')
        if x.dim() != 4:
            raise RuntimeError(f'{type(self).__name__}: expected [B, C, H, W], got {tuple(x.shape)}')
        self._check_supported()
        x = x.contiguous()
        with torch.cuda.device(x.device), _abi.stage('cconv'):
            packed, bias, co, ci, kh, kw = self._packed()
            B, c2, H, W = x.shape
            if c2 != 2 * ci:
                raise RuntimeError(f'{type(self).__name__}: expected {2 * ci} input channels, got {c2}')
            rc = self.real_conv
            sh, sw = _pair(rc.stride)
            ph, pw = _pair(rc.padding)
            lib = _abi.lib()
            if self._transposed:
                oph, opw = _pair(rc.output_padding)
                Ho = (H - 1) * sh - 2 * ph + kh + oph
                Wo = (W - 1) * sw - 2 * pw + kw + opw
                fn = lib.se_cconvT2d_fwd
            else:
                Ho = (H + 2 * ph - kh) // sh + 1
                Wo = (W + 2 * pw - kw) // sw + 1
                fn = lib.se_cconv2d_fwd
            if Ho <= 0 or Wo <= 0:
                raise RuntimeError(f'{type(self).__name__}: input {tuple(x.shape)} too small for the kernel')
            y = torch.empty((B, 2 * co, Ho, Wo), device=x.device, dtype=torch.float32)
            ws_bytes = lib.se_cconv_workspace_bytes(B, c2, H, W, 2 * co, kh, kw, sh, sw,
                                                    int(self._transposed)) if self.engine != _abi.ENGINE_SIMT else 0
            ws = torch.empty((ws_bytes,), device=x.device, dtype=torch.uint8) if ws_bytes > 0 else None
            _abi.check(fn(_abi.ptr(x), _abi.ptr(packed), _abi.ptr(bias), B, c2, H, W, 2 * co, kh, kw,
                          sh, sw, ph, pw, _abi.ptr(y), Ho, Wo, self.engine, _abi.ptr(ws), ws_bytes,
                          _abi.stream_ptr()))
        return y


def _halve(in_channels, out_channels):
    assert in_channels % 2 == 0, f'in_channels must be a factor of 2, current channels: {in_channels}'
    assert out_channels % 2 == 0, f'out_channels must be a factor of 2, current channels: {out_channels}'
    return in_channels // 2, out_channels // 2


class ComplexConv2d(ComplexConv):
    def __init__(self, in_channels, out_channels, kernel_size, **kwargs):
        super().__init__()
        ci, co = _halve(in_channels, out_channels)
        # parameter holders only (never called): keep state_dict keys and initialize_params() working
        self.real_conv = nn.Conv2d(ci, co, kernel_size, **kwargs)
        self.imag_conv = nn.Conv2d(ci, co, kernel_size, **kwargs)


class ComplexConvTranspose2d(ComplexConv):
    _transposed = True

    def __init__(self, in_channels, out_channels, kernel_size, **kwargs):
        super().__init__()
        ci, co = _halve(in_channels, out_channels)
        self.real_conv = nn.ConvTranspose2d(ci, co, kernel_size, **kwargs)
        self.imag_conv = nn.ConvTranspose2d(ci, co, kernel_size, **kwargs)
