"""se_cconv2d_fwd_static (allocation-free entry, scratch in a static __device__ array) against
se_cconv2d_fwd with a caller-provided workspace, through the C ABI, on two calls in a row.
Named test_zz_* so that pytest collects it last."""
import pytest
import torch

from speech_enhancement_b200 import _abi
from conftest import rand, rel_l2

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
TOL_UMMA = 2e-4          # module-level tolerance of the bf16x3 tensor engine (tests/test_gpu_parity.py)
STATIC_WS_BYTES = 32 * 1024 * 1024


def _conv(lib, fn_static, x, packed, bias, cout2, k, s, p, y=None):
    B, c2, H, W = x.shape
    Ho, Wo = (H + 2 * p[0] - k[0]) // s[0] + 1, (W + 2 * p[1] - k[1]) // s[1] + 1
    if y is None:
        y = torch.full((B, cout2, Ho, Wo), float('nan'), device=x.device)
    common = (_abi.ptr(x), _abi.ptr(packed), _abi.ptr(bias), B, c2, H, W, cout2, k[0], k[1], s[0], s[1], p[0], p[1],
              _abi.ptr(y), Ho, Wo)
    if fn_static:
        rc = lib.se_cconv2d_fwd_static(*common, _abi.stream_ptr())
    else:
        n = lib.se_cconv_workspace_bytes(B, c2, H, W, cout2, k[0], k[1], s[0], s[1], 0)
        ws = torch.empty((n,), device=x.device, dtype=torch.uint8)
        rc = lib.se_cconv2d_fwd(*common, _abi.ENGINE_UMMA, _abi.ptr(ws), n, _abi.stream_ptr())
    return rc, y


def test_static_workspace_entry_matches_workspace_entry_twice(built_lib):
    lib = built_lib
    assert lib.se_device_ok() == 1
    # two different problems back to back: the second is smaller, so stale bytes of the first stay in the array
    cases = [((3, 64, 33, 70), 128, (5, 2), (2, 1), (2, 0), True),
             ((2, 32, 17, 41), 48, (3, 3), (1, 1), (1, 1), False)]
    for rep in range(2):
        for i, (xs, cout2, k, s, p, has_bias) in enumerate(cases):
            x = rand(xs, 900 + 10 * rep + i).to(DEV)
            packed = (rand((cout2, xs[1] * k[0] * k[1]), 950 + i) * 0.05).to(DEV)
            bias = rand((cout2,), 970 + i).to(DEV) if has_bias else None
            rc0, y0 = _conv(lib, False, x, packed, bias, cout2, k, s, p)
            _abi.check(rc0)
            y1 = torch.full_like(y0, float('nan'))
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info()[0]
            rc1, y1 = _conv(lib, True, x, packed, bias, cout2, k, s, p, y=y1)
            _abi.check(rc1)
            torch.cuda.synchronize()
            free1 = torch.cuda.mem_get_info()[0]
            assert torch.isfinite(y1).all()
            assert rel_l2(y1.cpu(), y0.cpu()) <= TOL_UMMA
            assert torch.equal(y1, y0)            # same kernels, same staging: bit-identical in practice
            if rep == 1:                           # library module loaded and warm: device memory must not move
                assert free1 == free0


def test_static_workspace_entry_rejects_oversize_and_unsupported(built_lib):
    lib = built_lib
    x = torch.zeros((1, 64, 8, 8), device=DEV)
    packed = torch.zeros((64, 64 * 9), device=DEV)
    y = torch.zeros((1, 64, 8, 8), device=DEV)
    # B * H * W large enough that the CL2 staging alone exceeds the 32 MiB array: nothing is launched or read
    B, H, W = 64, 128, 128
    need = lib.se_cconv_workspace_bytes(B, 64, H, W, 64, 3, 3, 1, 1, 0)
    assert need > STATIC_WS_BYTES
    n0 = lib.se_launch_count()
    rc = lib.se_cconv2d_fwd_static(_abi.ptr(x), _abi.ptr(packed), None, B, 64, H, W, 64, 3, 3, 1, 1, 1, 1,
                                   _abi.ptr(y), H, W, _abi.stream_ptr())
    assert rc == _abi.SE_ERR_WORKSPACE and b'SE_STATIC_WS_BYTES' in lib.se_last_error()
    # 2-channel input is outside the tensor engine
    rc = lib.se_cconv2d_fwd_static(_abi.ptr(x), _abi.ptr(packed), None, 1, 2, 8, 8, 64, 3, 3, 1, 1, 1, 1,
                                   _abi.ptr(y), 8, 8, _abi.stream_ptr())
    assert rc == _abi.SE_ERR_UNSUPPORTED
    assert lib.se_launch_count() == n0
