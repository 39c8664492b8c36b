"""Two devices in ONE process (model.to('cuda:1'), nn.DataParallel-style use): the > 48 KB dynamic shared-memory opt-in,
the SM count and the static scratch pointer are per-device state inside libse_b200.so.  Skipped on a single-GPU box."""
import pytest
import torch

import speech_enhancement_b200 as se
from speech_enhancement_b200 import callers
from speech_enhancement_b200.synth import seeded_state_dict, synth_noisy_speech
from conftest import rel_l2

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two CUDA devices in one process')]


def test_same_model_on_two_devices_in_one_process():
    x = synth_noisy_speech(2, 16000, seed=5)
    outs = []
    for dev in ('cuda:0', 'cuda:1', 'cuda:0'):
        model = callers.DCCRN('dccrn-CL', 400, 100, 512).eval()
        model.load_state_dict(seeded_state_dict(model, seed=9))
        model.to(dev)
        with torch.no_grad():
            spec, wav = model(x.to(dev))                       # chained engine: GEMM, first layer, LSTM, FFT kernels
            model.fused = False
            _, wav_mod = model(x.to(dev))                      # module path: workspace GEMM, SIMT / direct kernels
        torch.cuda.synchronize(dev)
        assert wav.device == torch.device(dev)
        assert rel_l2(wav.cpu(), wav_mod.cpu()) < 1e-4
        outs.append(wav.cpu())
    assert torch.equal(outs[0], outs[2])
    assert rel_l2(outs[1], outs[0]) < 1e-6                     # same kernels, same inputs on the second device


def test_static_workspace_entry_on_the_second_device(built_lib):
    from speech_enhancement_b200 import _abi
    lib = built_lib
    res = []
    for dev in ('cuda:0', 'cuda:1'):
        with torch.cuda.device(dev):
            x = torch.randn(2, 64, 9, 40, generator=torch.Generator().manual_seed(1)).to(dev)
            packed = (torch.randn(64, 64 * 9, generator=torch.Generator().manual_seed(2)) * 0.05).to(dev)
            y = torch.empty((2, 64, 9, 40), device=dev)
            rc = lib.se_cconv2d_fwd_static(_abi.ptr(x), _abi.ptr(packed), None, 2, 64, 9, 40, 64, 3, 3, 1, 1, 1, 1,
                                           _abi.ptr(y), 9, 40, _abi.stream_ptr())
            _abi.check(rc)
            torch.cuda.synchronize(dev)
            res.append(y.cpu())
    assert torch.isfinite(res[1]).all() and rel_l2(res[1], res[0]) < 1e-6
