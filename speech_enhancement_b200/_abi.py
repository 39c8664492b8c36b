"""ctypes binding of libse_b200.so (C ABI declared in include/se_b200.h).

There is deliberately NO fallback: if the shared library is missing or a call fails the
caller gets a RuntimeError.  Tensors are passed as raw device pointers plus sizes; launches
go on ``torch.cuda.current_stream()`` so they order with surrounding PyTorch ops.
"""
import ctypes as C
import os
import re

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'csrc', 'libse_b200.so')
HEADER_PATH = os.path.join(os.path.dirname(_HERE), 'include', 'se_b200.h')

SE_OK, SE_ERR_BAD_ARG, SE_ERR_UNSUPPORTED, SE_ERR_CUDA, SE_ERR_WORKSPACE = 0, -1, -2, -3, -4
MASK_DCCRN_R, MASK_DCCRN_C, MASK_DCCRN_E = 0, 1, 2
MASK_DCUNET_UNB, MASK_DCUNET_SIGMOID, MASK_DCUNET_TANH, MASK_CARN, MASK_FRCRN = 3, 4, 5, 6, 7
ENGINE_AUTO, ENGINE_SIMT, ENGINE_UMMA = 0, 1, 2

_P, _I, _L = C.c_void_p, C.c_int, C.c_int64
_SIGS = {
    'se_version': (C.c_int, []),
    'se_last_error': (C.c_char_p, []),
    'se_device_ok': (C.c_int, []),
    'se_launch_count': (C.c_int64, []),
    'se_stft_fwd': (C.c_int, [_P, _L, _L, _P, _I, _I, _I, _I, _I, _P, _L, _P]),
    'se_magphase_fwd': (C.c_int, [_P, _L, _I, _L, _P, _P, _P]),
    'se_istft_pack_basis': (C.c_int, [_P, _I, _I, _I, _P, _P]),
    'se_istft_fwd': (C.c_int, [_P, _P, _P, _P, _L, _I, _L, _I, _I, _L, _L, _P, _P]),
    'se_cconv_pack_weight': (C.c_int, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P, _P, _P]),
    'se_cconv2d_fwd': (C.c_int, [_P, _P, _P, _L] + [_I] * 10 + [_P, _I, _I, _I, _P, _L, _P]),
    'se_cconvT2d_fwd': (C.c_int, [_P, _P, _P, _L] + [_I] * 10 + [_P, _I, _I, _I, _P, _L, _P]),
    'se_cconv2d_fwd_static': (C.c_int, [_P, _P, _P, _L] + [_I] * 10 + [_P, _I, _I, _P]),
    'se_cconv_workspace_bytes': (C.c_int64, [_L] + [_I] * 9),
    'se_cl2_from_nchw': (C.c_int, [_P, _L, _I, _I, _I, _P, _I, _P]),
    'se_cl2_from_strided': (C.c_int, [_P, _L, _L, _L, _L, _I, _I, _I, _P, _I, _P]),
    'se_stft_frames_cl2': (C.c_int, [_P, _L, _L, _I, _I, _I, _I, _P, _P]),
    'se_cl2_to_nchw': (C.c_int, [_P, _L, _I, _I, _I, _I, _P, _P]),
    'se_cl2_from_lstm_rows': (C.c_int, [_P, _L, _L, _I, _I, _I, _P, _I, _P]),
    'se_cconv_wq_elems': (C.c_int64, [_I, _I, _I]),
    'se_cconv_prep_weight': (C.c_int, [_P, _I, _I, _I, _P, _I, _P, _P]),
    'se_cconv_gemm': (C.c_int, [_P, _P]),
    'se_cconv_small_cin': (C.c_int, [_P, _L, _L, _L, _L, _I, _I, _I, _I, _I, _P, _P, C.c_float] + [_I] * 7
                           + [_P, _P, _I, _I, _I, _P]),
    'se_lstm_rec_supported': (C.c_int, [_I]),
    'se_lstm_rec_fwd': (C.c_int, [_P, _L, _P, _L, _L, _I, _P, _P]),
    'se_lstm_rec_pair_fwd': (C.c_int, [_P, _L, _P, _P, _L, _L, _I, _P, _P, _P]),
    'se_cl2_from_rows': (C.c_int, [_P, _L, _I, _L, _P, _I, _P]),
    'se_cconvT_col2im': (C.c_int, [_P, _L] + [_I] * 12 + [_P, C.c_float, _P, _P]),
    'se_nc2_elems': (C.c_int64, [_L, _I, _I, _I]),
    'se_nc2_from_nchw': (C.c_int, [_P, _L, _I, _I, _I, _P, _P]),
    'se_nc2_gather_elems': (C.c_int64, [_L, _I, _I, _I, _I]),
    'se_nc2_gather_w': (C.c_int, [_P, _L, _I, _I, _I, _I, _I, _I, _I, _P, _P]),
    'se_cconv_wgrad': (C.c_int, [_P, _I, _I, _I, _P, _I, _I, _L, _I, _I, _I, _I, _P, _P]),
    'se_ccbam_num_partials': (C.c_int, [_L]),
    'se_ccbam_channel_stats': (C.c_int, [_P, _L, _L, _I, _I, _P, _P]),
    'se_ccbam_spatial_pool': (C.c_int, [_P, _P, _L, _L, _I, _I, _P, _P]),
    'se_ccbam_apply': (C.c_int, [_P, _P, _P, _L, _L, _I, _I, _P, _P]),
    'se_ccbam_channel_attn': (C.c_int, [_P, _I, _L, _P, _P, _L, _I, _I, _I, _P, _P]),
    'se_ccbam_spatial_attn': (C.c_int, [_P, _P, _P, _L, _I, _I, _P, _P]),
    'se_cbn_splits': (C.c_int, [_L, _I, _I, _I]),
    'se_cbn_stats': (C.c_int, [_P, _L, _I, _I, _I, _P, _P]),
    'se_cbn_apply': (C.c_int, [_P, _P, _L, _I, _I, _I, _P, _P]),
    'se_cbn_bwd_stats': (C.c_int, [_P, _P, _L, _I, _I, _I, _P, _P]),
    'se_cbn_bwd_apply': (C.c_int, [_P, _P, _P, _L, _I, _I, _I, _P, _P]),
    'se_mask_fwd': (C.c_int, [_I] + [_P, _L, _L] * 4 + [_L, _I, _L, _P, _P]),
    'se_mask_bwd': (C.c_int, [_I] + [_P, _L, _L] * 4 + [_L, _I, _L, _P, _P, _P, _P, _P, _P]),
    'se_polar_to_rect': (C.c_int, [_P, _P, _L, _I, _L, _P, _P]),
    'se_stft_fft_supported': (C.c_int, [_I, _I, _I]),
    'se_stft_fft_fwd': (C.c_int, [_P, _L, _L, _P, _I, _I, _I, _I, _P, _L, _P]),
    'se_istft_fft_fwd': (C.c_int, [_P, _L, _L, _P, _I, _I, _I, _P, _L, _L, _P, _P]),
    'se_stft_fft_fwd_pcm16': (C.c_int, [_P, _L, _L, _P, _I, _I, _I, _I, _P, _L, _P]),
    'se_istft_fft_fwd_pcm16': (C.c_int, [_P, _L, _L, _P, _I, _I, _I, _P, _L, _L, _P, _P]),
    'se_pcm16_to_f32': (C.c_int, [_P, _L, _P, _P]),
    'se_pcm16_from_f32': (C.c_int, [_P, _L, _P, _P]),
}



class ConvArgs(C.Structure):
    """mirror of `se_cconv_args` (include/se_b200.h)"""
    _fields_ = [('transposed', C.c_int), ('B', C.c_int64), ('nsrc', C.c_int), ('src', C.c_void_p * 2),
                ('src_Cp', C.c_int * 2), ('src_Hpitch', C.c_int * 2), ('src_Wpitch', C.c_int * 2),
                ('H', C.c_int), ('W', C.c_int), ('pad_top', C.c_int), ('pad_left', C.c_int),
                ('Cout2', C.c_int), ('kh', C.c_int), ('kw', C.c_int), ('sh', C.c_int), ('sw', C.c_int),
                ('ph', C.c_int), ('pw', C.c_int), ('Ho', C.c_int), ('Wo', C.c_int),
                ('wq', C.c_void_p), ('bias', C.c_void_p), ('slope', C.c_float), ('y_nchw', C.c_void_p),
                ('y_cl2', C.c_void_p), ('y_Cp', C.c_int),
                ('y_bs', C.c_int64), ('y_cs', C.c_int64), ('y_hs', C.c_int64), ('y_ws', C.c_int64),
                ('y_off', C.c_int64), ('y_lim', C.c_int64), ('y_scale', C.c_void_p),
                ('src_H', C.c_int * 2), ('src_W', C.c_int * 2), ('src_C', C.c_int * 2),
                ('fold_h', C.c_int), ('fold_w', C.c_int), ('fold_cg', C.c_int), ('fold_cout2', C.c_int),
                ('out_h_mul', C.c_int), ('out_h_add', C.c_int), ('out_w_mul', C.c_int), ('out_w_add', C.c_int)]


_lib = None


def declared_symbols():
    """Names of every function prototyped in include/se_b200.h."""
    with open(HEADER_PATH) as f:
        return sorted(set(re.findall(r'SE_API\s+[\w\s\*]+?\b(se_\w+)\s*\(', f.read())))


def lib():
    """Load (once) and return the ctypes handle; raises RuntimeError when the .so is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f'{LIB_PATH} not found: build it with `python -c "import __graft_entry__ as g; g.build()"` '
                '(speech_enhancement_b200 has no CPU or PyTorch fallback)')
        h = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(h, name)
            fn.restype, fn.argtypes = res, args
        _lib = h
    return _lib


_EXC = {SE_ERR_BAD_ARG: ValueError, SE_ERR_UNSUPPORTED: RuntimeError, SE_ERR_CUDA: RuntimeError,
        SE_ERR_WORKSPACE: RuntimeError}


def check(rc):
    if rc != SE_OK:
        msg = lib().se_last_error().decode(errors='replace')
        raise _EXC.get(rc, RuntimeError)(f'libse_b200: {msg} (code {rc})')


# ---- derived-operand caches (packed conv weights, folded BN operands, linear operands, STFT bases) ------------
# Every cache is keyed on the (data_ptr, _version, device) of the tensors it was derived from.  Autograd's version
# counter sees optimizer steps, load_state_dict and in-place ops on the parameter itself, but NOT writes through
# `param.data` (reference utils.py:50-69 initialize_params, complex_nn.py:205-209, EMA swaps): `.data` is a detached
# alias with its own counter.  After such an edit call `invalidate_caches()` (or `module.invalidate()`); nn.Module
# `_apply` (.to / .cuda / .float) and `_load_from_state_dict` of this package's modules do it themselves.
# SE_VERIFY_CACHES=1 adds a checksum of every source tensor to the key (one device sync per lookup: debugging only).
_cache_epoch = 0
_VERIFY = os.environ.get('SE_VERIFY_CACHES', '0') not in ('', '0')


def invalidate_caches():
    """drop every derived operand of this package (they are rebuilt on next use)"""
    global _cache_epoch
    _cache_epoch += 1


def cache_key(tensors, *extra):
    key = [_cache_epoch]
    for t in tensors:
        if t is None:
            key.append(None)
            continue
        key.append((t.data_ptr(), t._version, t.device))
        if _VERIFY:
            d = t.detach().double()
            key.append((float(d.sum()), float(d.abs().sum())))
    return tuple(key) + tuple(extra)


class InvalidatesCaches:
    """nn.Module mixin: moving / casting the module or loading a state_dict drops the derived operands"""

    def invalidate(self):
        invalidate_caches()

    def _apply(self, fn, *args, **kwargs):
        invalidate_caches()
        return super()._apply(fn, *args, **kwargs)

    def _load_from_state_dict(self, *args, **kwargs):
        invalidate_caches()
        return super()._load_from_state_dict(*args, **kwargs)


# ---- optional per-stage timing (bench.py installs a timer; no-op otherwise) ----------------------
_timer = None


class stage:
    """`with _abi.stage('cconv'):` brackets the launches of one hot-path stage for bench.py's CUDA-event
    profiling pass; costs one attribute test when no timer is installed."""
    __slots__ = ('kind',)

    def __init__(self, kind):
        self.kind = kind

    def __enter__(self):
        if _timer is not None:
            _timer.start(self.kind)

    def __exit__(self, *exc):
        if _timer is not None:
            _timer.stop(self.kind)
        return False


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def require_cuda_f32(t, name):
    """The kernels take fp32 CUDA tensors only (SURVEY.md 8b error conventions)."""
    if not isinstance(t, torch.Tensor):
        raise ValueError(f'{name}: expected a torch.Tensor, got {type(t).__name__}')
    if not t.is_cuda:
        raise RuntimeError(f'{name}: speech_enhancement_b200 runs on CUDA (sm_100a) only; got a '
                           f'{t.device.type} tensor and there is no CPU fallback')
    if t.dtype != torch.float32:
        raise RuntimeError(f'{name}: only float32 is supported, got {t.dtype}')
    return t
