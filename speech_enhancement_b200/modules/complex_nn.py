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
class ComplexConv(_abi.InvalidatesCaches, nn.Module):
    """Shared forward of the two complex conv layers (complex_nn.py:45-65)."""
    _transposed = False
    engine = _abi.ENGINE_AUTO          # class-level knob: ENGINE_AUTO / ENGINE_SIMT / ENGINE_UMMA

    def __init__(self):
        super().__init__()
        self.real_conv = None
        self.imag_conv = None
        self._pack = None
        self._pack_key = None

    # packed operand cache, keyed on (data_ptr, _version) of the weights: rebuilt after optimizer.step(),
    # load_state_dict, .to() and in-place ops on the parameters.  Writes through `weight.data` (reference
    # utils.py:50-69 initialize_params) are invisible to the version counter: call `invalidate()` after them.
    def _packed(self):
        rc, ic = self.real_conv, self.imag_conv
        ts = [rc.weight, ic.weight] + ([rc.bias, ic.bias] if rc.bias is not None else [])
        key = _abi.cache_key(ts)
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
            with torch.cuda.device(wr.device):      # the weights' device, not whatever device is current
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
        # real_conv / imag_conv are parameter HOLDERS whose forward never runs, so anything that recomputes
        # `.weight` in a forward pre-hook (torch.nn.utils.weight_norm, reference utils.py:54-55 weight_norm=True) or
        # through a parametrization would silently be ignored
        for holder in (rc, self.imag_conv):
            if holder._forward_pre_hooks or getattr(holder, 'parametrizations', None) is not None \
                    or not isinstance(holder._parameters.get('weight'), nn.Parameter):
                raise RuntimeError(f'{type(self).__name__}: weight_norm / parametrizations / forward pre-hooks on '
                                   'real_conv / imag_conv are not supported (their forward is never called); '
                                   'remove them (torch.nn.utils.remove_weight_norm) before the first forward')

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

    def _geometry(self):
        rc = self.real_conv
        return _pair(rc.stride) + _pair(rc.padding)

    def _launch(self, transposed, x, packed, bias, cout2, kh, kw, Ho, Wo):
        """one engine call: NCHW fp32 in / out, `packed` [cout2][cin2*kh*kw] in the engine's layout"""
        B, c2, H, W = x.shape
        sh, sw, ph, pw = self._geometry()
        lib = _abi.lib()
        fn = lib.se_cconvT2d_fwd if transposed else lib.se_cconv2d_fwd
        y = torch.empty((B, cout2, Ho, Wo), device=x.device, dtype=torch.float32)
        ws_bytes = lib.se_cconv_workspace_bytes(B, c2, H, W, cout2, kh, kw, sh, sw,
                                                int(transposed)) if self.engine != _abi.ENGINE_SIMT else 0
        ws = torch.empty((ws_bytes,), device=x.device, dtype=torch.uint8) if ws_bytes > 0 else None
        with torch.cuda.device(x.device):
            _abi.check(fn(_abi.ptr(x), _abi.ptr(packed), _abi.ptr(bias), B, c2, H, W, cout2, kh, kw,
                          sh, sw, ph, pw, _abi.ptr(y), Ho, Wo, self.engine, _abi.ptr(ws), ws_bytes, _abi.stream_ptr()))
        return y

    def _out_size(self, H, W, kh, kw):
        sh, sw, ph, pw = self._geometry()
        if self._transposed:
            oph, opw = _pair(self.real_conv.output_padding)
            return (H - 1) * sh - 2 * ph + kh + oph, (W - 1) * sw - 2 * pw + kw + opw
        return (H + 2 * ph - kh) // sh + 1, (W + 2 * pw - kw) // sw + 1

    def _forward_stacked(self, x):
        x = _abi.require_cuda_f32(x, f'{type(self).__name__}.forward(x)')
        if x.dim() != 4:
            raise RuntimeError(f'{type(self).__name__}: expected [B, C, H, W], got {tuple(x.shape)}')
        self._check_supported()
        x = x.contiguous()
        rc, ic = self.real_conv, self.imag_conv
        if torch.is_grad_enabled() and (x.requires_grad or rc.weight.requires_grad or ic.weight.requires_grad):
            return _CConvFn.apply(self, x, rc.weight, ic.weight, rc.bias, ic.bias)
        return self._forward_impl(x)

    def _forward_impl(self, x):
        with torch.cuda.device(x.device), _abi.stage('cconv'):
            packed, bias, co, ci, kh, kw = self._packed()
            B, c2, H, W = x.shape
            if c2 != 2 * ci:
                raise RuntimeError(f'{type(self).__name__}: expected {2 * ci} input channels, got {c2}')
            Ho, Wo = self._out_size(H, W, kh, kw)
            if Ho <= 0 or Wo <= 0:
                raise RuntimeError(f'{type(self).__name__}: input {tuple(x.shape)} too small for the kernel')
            return self._launch(self._transposed, x, packed, bias, 2 * co, kh, kw, Ho, Wo)

    # ---- backward pieces (autograd of complex_nn.py:52-65; the reference gets them from ATen) ----------
    def _dgrad(self, gy, x_shape):
        """d loss / d x: the adjoint convolution = the OTHER flavour of the engine with the packed weight's
        channel axes swapped (no tap flip: the engine's transposed form indexes taps like PyTorch's)."""
        packed, _, co, ci, kh, kw = self._packed()
        key = self._pack_key
        if self.__dict__.get('_pack_T_key') != key:
            pt = packed.view(2 * co, 2 * ci, kh * kw).permute(1, 0, 2).contiguous().view(2 * ci, -1)
            self.__dict__['_pack_T'], self.__dict__['_pack_T_key'] = pt, key
        with torch.cuda.device(gy.device), _abi.stage('cconv_dgrad'):
            return self._launch(not self._transposed, gy, self.__dict__['_pack_T'], None, 2 * ci, kh, kw,
                                x_shape[2], x_shape[3])

    def _wgrad(self, x, gy):
        """d loss / d (packed weight) [2co][2ci][taps] through se_cconv_wgrad (pixel-reduction GEMM)."""
        _, _, co, ci, kh, kw = self._packed()
        sh, sw, ph, pw = self._geometry()
        lib = _abi.lib()
        u, v = (x, gy) if self._transposed else (gy, x)       # u: un-shifted grid, v: sampled through the taps
        B = x.shape[0]
        with torch.cuda.device(x.device), _abi.stage('cconv_wgrad'):
            Cu, Hu, Wu = u.shape[1:]
            Cv, Hv, Wv = v.shape[1:]
            ub = torch.empty(lib.se_nc2_elems(B, Cu, Hu, Wu), device=x.device, dtype=torch.bfloat16)
            _abi.check(lib.se_nc2_from_nchw(_abi.ptr(u), B, Cu, Hu, Wu, _abi.ptr(ub), _abi.stream_ptr()))
            vb = torch.empty(lib.se_nc2_gather_elems(B, Cv, Hv, Wu, kw), device=x.device, dtype=torch.bfloat16)
            _abi.check(lib.se_nc2_gather_w(_abi.ptr(v), B, Cv, Hv, Wv, Wu, kw, sw, pw, _abi.ptr(vb), _abi.stream_ptr()))
            G = torch.zeros((kh * kw, Cu, Cv), device=x.device, dtype=torch.float32)
            _abi.check(lib.se_cconv_wgrad(_abi.ptr(ub), Cu, Hu, Wu, _abi.ptr(vb), Cv, Hv, B, kh, kw, sh, ph, _abi.ptr(G),
                                          _abi.stream_ptr()))
        # G[tap][cu][cv] -> engine layout [2co][2ci][tap]
        return G.permute(2, 1, 0) if self._transposed else G.permute(1, 2, 0)

    def _unpack_grads(self, dP):
        """gradient of the packed operand [[Wr, -Wi], [Wi, Wr]] -> (d real_conv.weight, d imag_conv.weight)"""
        _, _, co, ci, kh, kw = self._packed()
        dP = dP.reshape(2 * co, 2 * ci, kh, kw)
        d_wr = dP[:co, :ci] + dP[co:, ci:]
        d_wi = dP[co:, :ci] - dP[:co, ci:]
        if self._transposed:                                  # module layout is [ci][co][kh][kw]
            d_wr, d_wi = d_wr.transpose(0, 1), d_wi.transpose(0, 1)
        return d_wr.contiguous(), d_wi.contiguous()


class _CConvFn(torch.autograd.Function):
    """ComplexConv2d / ComplexConvTranspose2d as ONE autograd node: forward, dgrad and wgrad each one
    tensor-core launch on the packed real operand."""

    @staticmethod
    def forward(ctx, mod, x, wr, wi, br, bi):
        ctx.mod = mod
        ctx.save_for_backward(x)
        return mod._forward_impl(x)

    @staticmethod
    def backward(ctx, gy):
        mod = ctx.mod
        (x,) = ctx.saved_tensors
        gy = gy.contiguous().float()
        need = ctx.needs_input_grad
        dx = mod._dgrad(gy, x.shape) if need[1] else None
        d_wr = d_wi = d_br = d_bi = None
        if need[2] or need[3]:
            d_wr, d_wi = mod._unpack_grads(mod._wgrad(x, gy))
        if need[4] or need[5]:
            co = gy.shape[1] // 2
            s = gy.sum(dim=(0, 2, 3))
            d_br, d_bi = s[:co] + s[co:], s[co:] - s[:co]     # packed bias = [br - bi, br + bi]
        return None, dx, d_wr, d_wi, d_br, d_bi


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
