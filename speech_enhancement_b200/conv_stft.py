"""ConvSTFT / ConviSTFT on B200: drop-in for /root/reference/models/conv_stft.py.

Same names, constructor and forward signatures, public attributes and state_dict keys
(`weight`; `weight`, `window`, `enframe`) as the reference (conv_stft.py:30, :48, :70, :89);
the arithmetic runs in libse_b200.so.  The kernels consume the module's own `weight` buffer,
so a checkpoint with a different basis is honoured.
"""
import ctypes as C

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _abi

# engine selection (class attribute `engine` on both modules):
#   'auto'  the FFT kernels (stft_fft.cu) when `weight` is the windowed real-DFT basis of init_kernels and the
#           framing is covered (fft 512); else tensor cores when the framing fits the conv engine; else SIMT
#   'fft'   FFT kernels or raise;  'umma'  tensor cores or raise;  'simt' the exact-fp32 SIMT kernels
#   (stft_simt.cu).  Every engine consumes the module's own buffers: a checkpoint whose basis is NOT the closed
#   form fails the check below and is projected densely.
_MAX_TAPS = 16          # ceil(win / hop) taps per launch (se_cconv_gemm allows 49 in total)


def _round_up(a, m):
    return (a + m - 1) // m * m


def _tc_ok(win, hop):
    return hop >= 8 and hop <= 256 and -(-win // hop) <= _MAX_TAPS


def _prep_wq(packed, cout2, cin2, taps, Cp):
    """packed fp32 [cout2][cin2][taps] -> bf16 hi|lo GEMM operand (se_cconv_prep_weight)."""
    n = _abi.lib().se_cconv_wq_elems(cout2, Cp, taps)
    wq = torch.empty(n, device=packed.device, dtype=torch.bfloat16)
    with torch.cuda.device(packed.device):          # launches go on the current stream of the TENSOR's device
        _abi.check(_abi.lib().se_cconv_prep_weight(_abi.ptr(packed), cout2, cin2, taps, None, Cp, _abi.ptr(wq),
                                                   _abi.stream_ptr()))
    return wq


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


def _dft_rows(win, fft, device):
    """[cos; -sin](2 pi f n / fft), f = 0..fft/2, n = 0..win-1, float64 [fft+2, win] (conv_stft.py:11-15)"""
    n = torch.arange(win, device=device, dtype=torch.int64)[None, :]
    f = torch.arange(fft // 2 + 1, device=device, dtype=torch.int64)[:, None]
    phase = (2.0 * np.pi / fft) * ((f * n) % fft).double()
    return torch.cat([torch.cos(phase), -torch.sin(phase)], 0)


def _is_dft_basis(weight, window, fft, inverse):
    """True when `weight` [fft+2, 1, win] equals the closed form the FFT kernels implement, to fp32 rounding:
    forward  window[n] * [cos; -sin](2 pi f n / fft);  inverse  window[n] * pinv(K)^T with
    pinv(K)^T = K (K^T K)^-1 and K^T K = (fft/2) I + e e^T + o o^T (e / o = even / odd sample indicators)."""
    c2, _, win = weight.shape
    if c2 != fft + 2 or win > fft or window.numel() != win:
        return False
    w = weight.detach()[:, 0, :].double()
    K = _dft_rows(win, fft, weight.device)
    if inverse:
        a = fft / 2.0
        par = (torch.arange(win, device=weight.device) % 2).double()
        e, o = 1.0 - par, par
        K = (K - torch.outer(K @ e, e) / (a + e.sum()) - torch.outer(K @ o, o) / (a + o.sum())) / a
    ref = K * window.detach().reshape(1, win).double()
    return bool(((w - ref).abs().max() <= 2e-6 * ref.abs().max()).item())


def _frames_operand(w, hop):
    """basis [C2, 1, win] regrouped for the framing projection (a 1 x J convolution over hop-sample pixels,
    `hop` input channels): [C2][hop][J], output rows in chunks of <= 256."""
    c2, _, win = w.shape
    J = -(-win // hop)
    Cp = _round_up(hop, 64)
    wp = F.pad(w[:, 0, :], (0, J * hop - win)).view(c2, J, hop).permute(0, 2, 1).contiguous()
    chunks = []
    for c0 in range(0, c2, 256):
        n = min(256, c2 - c0)
        chunks.append((c0, n, _prep_wq(wp[c0:c0 + n].contiguous(), n, hop, J, Cp)))
    return J, Cp, chunks


def _ola_operand(w, hop):
    """basis [C2, 1, win] regrouped for overlap-add (a 1 x J convolution over frames with C2 input channels
    and one output channel per hop phase): [hop][C2][J];  y[q*hop + r] = sum_j sum_c W[c, j*hop + r] spec[c, q-j]."""
    c2, _, win = w.shape
    J = -(-win // hop)
    Cp = _round_up(c2, 64)
    wp = F.pad(w[:, 0, :], (0, J * hop - win)).view(c2, J, hop).flip(1).permute(2, 0, 1).contiguous()
    return J, Cp, _prep_wq(wp, hop, c2, J, Cp)


def _frames_project(x, operand, hop, pad, T, out):
    """out[b, c, t] = sum_k W[c, k] * reflect_pad(x, pad)[b, t*hop + k]  (conv_stft.py:60-63)"""
    B, L = x.shape
    c2 = out.shape[1]
    J, Cp, chunks = operand
    Q = T + J - 1
    frames = torch.empty((1, 2, B, Q, Cp), device=x.device, dtype=torch.bfloat16)
    _abi.check(_abi.lib().se_stft_frames_cl2(_abi.ptr(x), B, L, hop, pad, Q, Cp, _abi.ptr(frames), _abi.stream_ptr()))
    for c0, n, wq in chunks:
        a = _abi.ConvArgs()
        a.transposed, a.B, a.nsrc = 0, 1, 1
        a.src[0] = frames.data_ptr()
        a.src_Cp[0], a.src_Hpitch[0], a.src_Wpitch[0] = Cp, B, Q
        a.H, a.W, a.Cout2, a.kh, a.kw, a.sh, a.sw, a.Ho, a.Wo = B, Q, n, 1, J, 1, 1, B, T
        a.wq, a.bias, a.slope = wq.data_ptr(), None, 1.0
        a.y_nchw = out.data_ptr() + 4 * c0 * T
        a.y_cs, a.y_hs, a.y_ws = T, c2 * T, 1            # out[b=row, c0 + co, t]
        _abi.check(_abi.lib().se_cconv_gemm(C.byref(a), _abi.stream_ptr()))


def _overlap_add(spec, operand, hop, scale, out, start, out_len):
    """out[b, n - start] = scale[n] * sum_t sum_c W[c, n - t*hop] * spec[b, c, t]  (conv_stft.py:99-114)"""
    B, c2, T = spec.shape
    J, Cp, wq = operand
    img = torch.empty((1, 2, B, T, Cp), device=spec.device, dtype=torch.bfloat16)
    _abi.check(_abi.lib().se_cl2_from_strided(_abi.ptr(spec), 0, T, c2 * T, 1, c2, B, T, _abi.ptr(img), Cp,
                                              _abi.stream_ptr()))
    a = _abi.ConvArgs()
    a.transposed, a.B, a.nsrc = 0, 1, 1
    a.src[0] = img.data_ptr()
    a.src_Cp[0], a.src_Hpitch[0], a.src_Wpitch[0] = Cp, B, T
    a.H, a.W, a.Cout2, a.kh, a.kw, a.sh, a.sw, a.ph, a.pw = B, T, hop, 1, J, 1, 1, 0, J - 1
    a.Ho, a.Wo = B, T + J - 1
    a.wq, a.bias, a.slope = wq.data_ptr(), None, 1.0
    a.y_nchw = out.data_ptr()
    a.y_cs, a.y_hs, a.y_ws, a.y_off, a.y_lim = 1, out_len, hop, -start, out_len
    a.y_scale = scale.data_ptr() if scale is not None else None
    _abi.check(_abi.lib().se_cconv_gemm(C.byref(a), _abi.stream_ptr()))


class _StftFn(torch.autograd.Function):
    """ConvSTFT as one autograd node.  The op is linear, so the backward is the adjoint: overlap-add of the
    gradient frames with the SAME basis, folded back through the reflect padding."""

    @staticmethod
    def forward(ctx, mod, x):
        ctx.mod, ctx.L = mod, x.shape[1]
        return mod._project(x)

    @staticmethod
    def backward(ctx, g):
        return None, ctx.mod._project_adjoint(g.contiguous().float(), ctx.L)


class _IstftFn(torch.autograd.Function):
    """ConviSTFT as one autograd node; backward = framing projection of the normalised, un-cropped gradient
    with the inverse basis."""

    @staticmethod
    def forward(ctx, mod, spec, start, out_len):
        ctx.mod, ctx.T, ctx.start = mod, spec.shape[2], start
        return mod._synthesize(spec, start, out_len)

    @staticmethod
    def backward(ctx, g):
        return None, ctx.mod._synthesize_adjoint(g.contiguous().float(), ctx.T, ctx.start), None, None


class ConvSTFT(_abi.InvalidatesCaches, nn.Module):
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
        self._tc = None
        self._tc_key = None

    engine = 'auto'

    def _fft_ready(self, window, inverse):
        """the FFT kernels apply: framing covered and `weight` is the closed-form basis (checked once per
        weight version, on the device)"""
        w = self.weight
        key = _abi.cache_key((w, window))
        c = self.__dict__.get('_fft_check')
        if c is None or c[0] != key:
            ok = (_abi.lib().se_stft_fft_supported(self.window_size, self.hop_size, self.fft_size) == 1
                  and _is_dft_basis(w, window, self.fft_size, inverse))
            c = self.__dict__['_fft_check'] = (key, ok)
        return c[1]

    def _pick_engine(self, name, window, inverse, nonempty=True):
        win, hop = self.window_size, self.hop_size
        if self.engine in ('auto', 'fft') and nonempty and self._fft_ready(window, inverse):
            return 'fft'
        if self.engine == 'fft':
            raise RuntimeError(f'{name}: window {win} / hop {hop} / fft {self.fft_size} (or a `weight` that is not the '
                               'windowed DFT basis) is outside the FFT engine')
        if self.engine != 'simt' and _tc_ok(win, hop) and nonempty:
            return 'umma'
        if self.engine == 'umma':
            raise RuntimeError(f'{name}: window {win} / hop {hop} does not fit the tensor-core framing')
        return 'simt'

    def _cached(self, slot, build):
        w = self.weight
        key = _abi.cache_key((w,), slot)
        c = self.__dict__.setdefault('_tc_cache', {})
        if c.get(slot, (None,))[0] != key:
            with torch.no_grad():
                c[slot] = (key, build(w.detach().contiguous(), self.hop_size))
        return c[slot][1]

    def _project(self, x):
        """[B, L] -> [B, C2, T] (+ nothing else): the tensor-core framing projection or the SIMT kernel"""
        w = self.weight
        B, L = x.shape
        c2, _, win = w.shape
        pad = self.pad if self.center else 0
        T = (L + 2 * pad - win) // self.hop_size + 1
        out = torch.empty((B, c2, T), device=x.device, dtype=torch.float32)
        if B == 0:                                        # empty batch: nothing to launch (F.conv1d also returns [0, C2, T])
            return out
        window = w[0, 0]                                  # row f = 0 of the real block is the analysis window
        eng = self._pick_engine('ConvSTFT', window, False, B > 0)
        with torch.cuda.device(x.device), _abi.stage('stft'):
            if x.dtype == torch.int16 and eng != 'fft':
                # framings outside the FFT engine: one conversion pass, then the fp32 engines
                xf = torch.empty(x.shape, device=x.device, dtype=torch.float32)
                _abi.check(_abi.lib().se_pcm16_to_f32(_abi.ptr(x), x.numel(), _abi.ptr(xf), _abi.stream_ptr()))
                x = xf
            if eng == 'fft':
                fn = _abi.lib().se_stft_fft_fwd_pcm16 if x.dtype == torch.int16 else _abi.lib().se_stft_fft_fwd
                _abi.check(fn(_abi.ptr(x), B, L, _abi.ptr(window), win, self.hop_size,
                              self.fft_size, pad, _abi.ptr(out), T, _abi.stream_ptr()))
            elif eng == 'umma':
                _frames_project(x, self._cached('fwd', _frames_operand), self.hop_size, pad, T, out)
            else:
                _abi.check(_abi.lib().se_stft_fwd(_abi.ptr(x), B, L, _abi.ptr(w.contiguous()), c2, win, self.hop_size,
                                                  self.pad, int(self.center), _abi.ptr(out), T, _abi.stream_ptr()))
        return out

    def _project_adjoint(self, g, L):
        """d loss / d x from d loss / d spec [B, C2, T]."""
        c2, _, win = self.weight.shape
        hop = self.hop_size
        if not _tc_ok(win, hop):
            raise RuntimeError('ConvSTFT: backward needs the tensor-core framing (hop >= 8, ceil(win/hop) <= 16)')
        B, _, T = g.shape
        pad = self.pad if self.center else 0
        full = (T - 1) * hop + win
        gp = torch.zeros((B, L + 2 * pad), device=g.device, dtype=torch.float32)
        with torch.cuda.device(g.device), _abi.stage('stft_bwd'):
            buf = gp if full == gp.shape[1] else torch.empty((B, full), device=g.device, dtype=torch.float32)
            _overlap_add(g, self._cached('adj', _ola_operand), hop, None, buf, 0, full)
            if buf is not gp:
                gp[:, :full] = buf
        if pad == 0:
            return gp
        gx = gp[:, pad:pad + L].clone()
        # adjoint of F.pad(mode='reflect') (conv_stft.py:60-62): the mirrored samples add onto their sources
        gx[:, 1:pad + 1] += gp[:, :pad].flip(1)
        gx[:, L - 1 - pad:L - 1] += gp[:, pad + L:].flip(1)
        return gx

    def forward(self, inputs):
        """inputs: fp32 waveform [L] / [B, L] / [B, 1, L] (conv_stft.py:48-52), or the same shapes in int16: PCM_S 16
        samples as they sit in the wav file (audio_dataset.py:12-21), scaled by 1 / 32768 in the kernel's stage load."""
        if torch.is_tensor(inputs) and inputs.dtype == torch.int16:
            if not inputs.is_cuda:
                raise RuntimeError('ConvSTFT.forward(inputs): speech_enhancement_b200 runs on CUDA (sm_100a) only; got a '
                                   'cpu tensor and there is no CPU fallback')
            x = inputs
        else:
            x = _abi.require_cuda_f32(inputs, 'ConvSTFT.forward(inputs)')
        if x.dim() == 1:
            x = x[None, :]
        elif x.dim() == 3:
            if x.shape[1] != 1:
                raise RuntimeError(f'ConvSTFT: expected [B,1,L], got {tuple(x.shape)}')
            x = x[:, 0, :]
        elif x.dim() != 2:
            raise RuntimeError(f'ConvSTFT: expected a 1-, 2- or 3-D tensor, got {x.dim()}-D')
        x = x.contiguous()
        w = _abi.require_cuda_f32(self.weight, 'ConvSTFT.weight')
        B, L = x.shape
        c2, _, win = w.shape
        pad = self.pad if self.center else 0
        if pad >= L and self.center:
            raise RuntimeError(f'ConvSTFT: reflect padding ({pad}) must be smaller than the input length ({L})')
        if L + 2 * pad < win:
            raise RuntimeError(f'ConvSTFT: input length {L} is shorter than one window ({win})')
        if torch.is_grad_enabled() and x.requires_grad:
            out = _StftFn.apply(self, x)
        else:
            out = self._project(x)
        if not self.return_mag_phase:
            return out
        F_ = self.fft_size // 2 + 1
        if torch.is_grad_enabled() and out.requires_grad:
            # CRN's polar branch under autograd (conv_stft.py:58-64): plain pointwise ops
            real, imag = out[:, :F_], out[:, F_:]
            return torch.sqrt(real ** 2 + imag ** 2), torch.atan2(imag, real)
        T = out.shape[2]
        mags = torch.empty((B, F_, T), device=x.device, dtype=torch.float32)
        phase = torch.empty_like(mags)
        with torch.cuda.device(x.device), _abi.stage('stft'):
            _abi.check(_abi.lib().se_magphase_fwd(_abi.ptr(out), B, F_, T, _abi.ptr(mags), _abi.ptr(phase),
                                                  _abi.stream_ptr()))
        return mags, phase


class ConviSTFT(_abi.InvalidatesCaches, nn.Module):
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
        self._scale = {}

    engine = 'auto'
    # True: forward returns int16 PCM_S 16 samples -- clamp(-1, 1) (the models' final clamp_, _2008_00264_dccrn.py:184),
    # round to nearest, saturate -- written by the overlap-add epilogue itself (inference only).  The chained forwards
    # (callers.py / dropin.accelerate) skip their clamp_ when they see an int16 waveform.
    pcm16_out = False

    _cached = ConvSTFT._cached
    _fft_ready = ConvSTFT._fft_ready
    _pick_engine = ConvSTFT._pick_engine

    def _ola_scale(self, T, J):
        """1 / (sum of squared windows + 1e-8) per output sample (conv_stft.py:101-106), [Q*hop]."""
        window = self.window
        key = _abi.cache_key((window,), T)
        sc = self._scale.get(key)
        if sc is None:
            hop, win = self.hop_size, self.window_size
            w2 = F.pad(window.detach().reshape(-1).double() ** 2, (0, J * hop - win)).view(J, hop)
            coff = torch.zeros((T + J - 1, hop), device=window.device, dtype=torch.float64)
            for j in range(J):
                coff[j:j + T] += w2[j]
            sc = (1.0 / (coff.float() + 1e-8)).reshape(-1).contiguous()
            self._scale = {key: sc}
        return sc

    def _synthesize(self, spec, start, out_len):
        """[B, C2, T] -> [B, out_len]: overlap-add, normaliser, crop -- tensor-core framing or SIMT kernel"""
        w = self.weight
        c2, _, win = w.shape
        B, _, T = spec.shape
        eng = self._pick_engine('ConviSTFT', self.window, True)
        fused_pcm = self.pcm16_out and eng == 'fft'
        out = torch.empty((B, out_len), device=spec.device, dtype=torch.int16 if fused_pcm else torch.float32)
        with torch.cuda.device(spec.device), _abi.stage('istft'):
            if eng == 'fft':
                J = -(-win // self.hop_size)
                fn = _abi.lib().se_istft_fft_fwd_pcm16 if fused_pcm else _abi.lib().se_istft_fft_fwd
                _abi.check(fn(_abi.ptr(spec), B, T, _abi.ptr(self.window), win, self.hop_size,
                              self.fft_size, _abi.ptr(self._ola_scale(T, J)), start, out_len,
                              _abi.ptr(out), _abi.stream_ptr()))
            elif eng == 'umma':
                operand = self._cached('fwd', _ola_operand)
                _overlap_add(spec, operand, self.hop_size, self._ola_scale(T, operand[0]), out, start, out_len)
            else:
                packed = self._packed_basis()
                window = self.window.reshape(-1).contiguous()
                _abi.check(_abi.lib().se_istft_fwd(_abi.ptr(spec), None, _abi.ptr(packed), _abi.ptr(window), B, c2, T,
                                                   win, self.hop_size, start, out_len, _abi.ptr(out),
                                                   _abi.stream_ptr()))
            if self.pcm16_out and not fused_pcm:
                pcm = torch.empty((B, out_len), device=spec.device, dtype=torch.int16)
                _abi.check(_abi.lib().se_pcm16_from_f32(_abi.ptr(out), out.numel(), _abi.ptr(pcm), _abi.stream_ptr()))
                out = pcm
        return out

    def _synthesize_adjoint(self, g, T, start):
        """d loss / d spec [B, C2, T] from d loss / d wav [B, out_len]."""
        c2, _, win = self.weight.shape
        hop = self.hop_size
        if not _tc_ok(win, hop):
            raise RuntimeError('ConviSTFT: backward needs the tensor-core framing (hop >= 8, ceil(win/hop) <= 16)')
        B, out_len = g.shape
        full = (T - 1) * hop + win
        operand = self._cached('adj', _frames_operand)
        gfull = torch.zeros((B, full), device=g.device, dtype=torch.float32)
        gfull[:, start:start + out_len] = g * self._ola_scale(T, operand[0])[start:start + out_len]
        out = torch.empty((B, c2, T), device=g.device, dtype=torch.float32)
        with torch.cuda.device(g.device), _abi.stage('istft_bwd'):
            _frames_project(gfull, operand, hop, 0, T, out)
        return out

    def _packed_basis(self):
        """inverse basis regrouped per hop phase (se_istft_pack_basis); rebuilt when `weight` changes."""
        w = self.weight
        key = _abi.cache_key((w,))
        if self._packed_key != key:
            c2, _, win = w.shape
            J = -(-win // self.hop_size)
            packed = torch.empty((self.hop_size, J * c2), device=w.device, dtype=torch.float32)
            with torch.cuda.device(w.device):
                _abi.check(_abi.lib().se_istft_pack_basis(_abi.ptr(w.contiguous()), c2, win, self.hop_size,
                                                          _abi.ptr(packed), _abi.stream_ptr()))
            self._packed, self._packed_key = packed, key
        return self._packed

    def forward(self, inputs, phase=None, output_length=None):
        """inputs: [B, N+2, T] complex spec, or [B, N//2+1, T] mags with `phase` (same shape)."""
        spec = _abi.require_cuda_f32(inputs, 'ConviSTFT.forward(inputs)')
        if spec.dim() != 3:
            raise RuntimeError(f'ConviSTFT: expected [B, C, T], got {tuple(spec.shape)}')
        spec = spec.contiguous()
        w = _abi.require_cuda_f32(self.weight, 'ConviSTFT.weight')
        c2, _, win = w.shape
        B, C, T = spec.shape
        needs_grad = torch.is_grad_enabled() and (spec.requires_grad or (phase is not None and phase.requires_grad))
        if phase is not None:
            phase = _abi.require_cuda_f32(phase, 'ConviSTFT.forward(phase)').contiguous()
            if phase.shape != spec.shape or 2 * C != c2:
                raise RuntimeError(f'ConviSTFT: mags {tuple(spec.shape)} / phase {tuple(phase.shape)} '
                                   f'do not match a {c2}-row basis')
            if needs_grad:
                # polar branch under autograd (conv_stft.py:96-100): plain pointwise ops in front of the node
                spec = torch.cat([spec * torch.cos(phase), spec * torch.sin(phase)], 1)
            else:
                rect = torch.empty((B, c2, T), device=spec.device, dtype=torch.float32)
                with torch.cuda.device(spec.device), _abi.stage('istft'):
                    _abi.check(_abi.lib().se_polar_to_rect(_abi.ptr(spec), _abi.ptr(phase), B, C, T, _abi.ptr(rect),
                                                           _abi.stream_ptr()))
                spec = rect
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
        if out_len == 0 or B == 0:
            return torch.empty((B, out_len), device=spec.device,
                               dtype=torch.int16 if self.pcm16_out and not needs_grad else torch.float32)
        if needs_grad:
            if self.pcm16_out:
                raise RuntimeError('ConviSTFT: pcm16_out is an inference option (int16 samples carry no gradient)')
            return _IstftFn.apply(self, spec, start, out_len)
        return self._synthesize(spec, start, out_len)
