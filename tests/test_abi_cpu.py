"""CPU-side checks of the C-ABI boundary: the library loads without a GPU driver, exports every
symbol include/se_b200.h declares, and argument validation fails loudly (no compute calls)."""
import ctypes as C
import os

import pytest
import torch

from speech_enhancement_b200 import _abi


def test_library_exports_every_declared_symbol(built_lib):
    names = _abi.declared_symbols()
    assert len(names) >= 13 and 'se_cconv2d_fwd' in names and 'se_stft_fwd' in names
    for n in names:
        assert hasattr(built_lib, n), n
    assert set(names) == set(_abi._SIGS), 'ctypes signatures out of sync with the header'
    assert built_lib.se_version() >= 100


def test_bad_arguments_return_codes(built_lib):
    null = C.c_void_p(0)
    assert built_lib.se_stft_fwd(null, 1, 100, null, 514, 400, 100, 256, 1, null, 1, null) == _abi.SE_ERR_BAD_ARG
    assert b'null pointer' in built_lib.se_last_error()
    with pytest.raises(ValueError):
        _abi.check(built_lib.se_mask_fwd(99, *([null, 0, 0] * 4), 1, 1, 1, null, null))
    one = C.c_void_p(8)   # never dereferenced: validation rejects first
    assert built_lib.se_cconv2d_fwd(one, one, null, 1, 3, 8, 8, 4, 3, 3, 1, 1, 1, 1, one, 8, 8, 0, null, 0,
                                    null) == _abi.SE_ERR_BAD_ARG
    assert b'even' in built_lib.se_last_error()
    assert built_lib.se_cconv2d_fwd(one, one, null, 1, 4, 8, 8, 4, 3, 3, 1, 1, 1, 1, one, 7, 8, 0, null, 0,
                                    null) == _abi.SE_ERR_BAD_ARG
    assert built_lib.se_stft_fwd(one, 1, 100, one, 514, 400, 100, 256, 1, one, 3, null) == _abi.SE_ERR_BAD_ARG


def test_modules_refuse_cpu_tensors(built_lib):
    import speech_enhancement_b200 as se
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        se.ConvSTFT(400, 100, 512)(torch.zeros(1, 4000))
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        se.ComplexConv2d(4, 4, 3)(torch.zeros(1, 4, 8, 8))
    with pytest.raises(AssertionError, match='in_channels must be a factor of 2, current channels: 3'):
        se.ComplexConv2d(3, 4, 3)
    with pytest.raises(ValueError):
        se.split_complex(3.0)


def test_state_dict_keys_match_reference_layout():
    import speech_enhancement_b200 as se
    assert list(se.ConvSTFT(400, 100, 512).state_dict()) == ['weight']
    assert list(se.ConviSTFT(400, 100, 512).state_dict()) == ['weight', 'window', 'enframe']
    assert sorted(se.ComplexConvTranspose2d(4, 4, 3, bias=True).state_dict()) == [
        'imag_conv.bias', 'imag_conv.weight', 'real_conv.bias', 'real_conv.weight']
    st = se.ConvSTFT(400, 100)
    assert (st.fft_size, st.hop_size, st.window_size, st.center, st.return_mag_phase, st.pad) == (400, 100, 400, True, False, 200)


def test_init_kernels_matches_oracle():
    import oracle
    import speech_enhancement_b200 as se
    for win, fft in ((400, 512), (320, 640), (256, 256)):
        for inv in (False, True):
            k, w = se.init_kernels(win, fft, inverse=inv)
            kr, wr = oracle.init_kernels_ref(win, fft, inverse=inv)
            assert k.shape == (fft + 2, 1, win) and w.shape == (1, win, 1)
            assert (k - kr).abs().max() < 2e-7 and torch.equal(w, wr)


def test_conv_args_mirror_matches_the_c_struct(tmp_path):
    """the ctypes mirror of `se_cconv_args` has the C struct's size and field offsets (compiled from include/se_b200.h
    with the host C compiler: the header is plain C)"""
    import ctypes as C
    import shutil
    import subprocess
    from speech_enhancement_b200 import _abi
    cc = shutil.which('gcc') or shutil.which('cc')
    if cc is None:
        pytest.skip('no host C compiler')
    fields = [name for name, _ in _abi.ConvArgs._fields_]
    src = tmp_path / 'layout.c'
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "se_b200.h"\nint main(void) {\n'
                   '  printf("%zu\\n", sizeof(se_cconv_args));\n' +
                   ''.join(f'  printf("%zu\\n", offsetof(se_cconv_args, {f}));\n' for f in fields) +
                   '  return 0;\n}\n')
    exe = tmp_path / 'layout'
    subprocess.run([cc, '-I', os.path.dirname(_abi.HEADER_PATH), str(src), '-o', str(exe)], check=True)
    out = [int(v) for v in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    assert out[0] == C.sizeof(_abi.ConvArgs)
    assert out[1:] == [getattr(_abi.ConvArgs, f).offset for f in fields]
