"""The reference's UNMODIFIED model classes (staged in oracle/_ref by oracle/stage_ref.py) on the GPU.

For every caller of the hot path the same seeded state_dict goes into
  (a) the reference exactly as it ships, on the CPU (oracle.ref_loader.load_original: its own conv_stft.py /
      complex_nn.py, fp32 PyTorch) -- the parity oracle,
  (b) the same model file bound to this package by dropin.install() on cuda:0, module by module, and
  (c) (b) after dropin.accelerate(model): the chained engine behind model.forward,
on 4 s (BASELINE.json shape) inputs.  Bar (north star): enhanced-waveform rel-L2 <= 1e-3 and SI-SNR delta
<= 0.01 dB (reference losses.SI_SNR_loss, losses.py:62-84, against the clean synthetic source); (b) and (c)
agree to 1e-4.  Skipped WITH A REASON when oracle/_ref is absent (it is git-ignored; build() stages it).
"""
import warnings

import pytest
import torch

from oracle import ref_loader
from speech_enhancement_b200 import dropin
from speech_enhancement_b200.synth import seeded_state_dict, synth_noisy_speech
from conftest import rel_l2

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not ref_loader.available(), reason=ref_loader.why_unavailable())]
DEV = 'cuda:0'
TOL_WAV, TOL_SISNR_DB, TOL_PATHS = 1e-3, 0.01, 1e-4

#        id                 class     ctor args / kwargs                                   B  L      chained?
CASES = [('dccrn_cl', 'DCCRN', (('dccrn-CL', 400, 100, 512), {}), 2, 64000, True),
         ('dccrn_e', 'DCCRN', (('dccrn-E', 400, 100, 512), {}), 1, 64000, True),
         ('dccrn_cl_bidirectional', 'DCCRN', (('dccrn-CL', 400, 100, 512), {'bidirectional': True}), 1, 64000, True),
         ('dcunet20', 'DCUNet', (('dcunet20', 512, 128, 512), {}), 1, 64000, True),
         ('dcunet20_large_test_py', 'DCUNet', (('dcunet20-large', 1024, 256, 1024), {}), 1, 32000, True),
         ('dcunet16_normalize', 'DCUNet', (('dcunet16', 512, 128, 512), {'normalize': True}), 2, 64000, True),
         ('frcrn', 'FRCRN', ((320, 160, 640), {}), 1, 64000, True),
         ('carn', 'CARN', ((320, 160, 512), {}), 2, 64000, False),
         ('crn_mag_phase', 'CRN', ((320, 160, 320), {}), 2, 64000, False)]


@pytest.fixture(autouse=True)
def _fp32_math_in_stock_layers():
    """the real convolutions / LSTMs of CARN and CRN are stock cuDNN (out of scope); cuDNN defaults to TF32 for
    convolutions, the parity target is the fp32 reference, so TF32 is switched off for these tests"""
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32 = old


def _si_snr_db(ns, est, clean):
    n = min(est.shape[1], clean.shape[1])
    return -ns.losses.SI_SNR_loss(est[:, :n].double(), clean[:, :n].double()).item()


def _wav_of(out):
    return out[1] if isinstance(out, (tuple, list)) else out


def _compare(ns, out, out_ref, clean, name, istft=None):
    """(spec, wav) of the CUDA drop-in against (spec, wav) of the CPU original.

    DCUNet's bounded-tanh mask divides the atan2 ANGLE by the mask magnitude (_1903_03107_dcunet.py:177), so it jumps
    where the mask crosses the negative real axis: a bin whose mask lies within rounding of that cut lands on either
    side depending on summation order (true of the reference against itself).  For DCUNet (`istft` given) such bins
    are counted (at most 1e-4 of all bins), replaced by the reference's value, and the waveform of the repaired
    spectrum is what is held to the bar; every other model is compared as is."""
    got, want = _wav_of(out).cpu(), _wav_of(out_ref)
    assert got.shape == want.shape, (name, got.shape, want.shape)
    if istft is not None:
        spec_a, spec_b = out[0], out_ref[0].to(out[0].device)
        cut = (spec_a - spec_b).abs() > 2e-3 * spec_b.abs().max()
        assert cut.sum().item() <= 1e-4 * cut.numel(), (name, cut.sum().item())
        if cut.any():
            with torch.no_grad():
                got = torch.clamp(istft(torch.where(cut, spec_b, spec_a).contiguous()), -1, 1).cpu()
    err = rel_l2(got, want)
    delta = abs(_si_snr_db(ns, got, clean) - _si_snr_db(ns, want, clean))
    assert err <= TOL_WAV, (name, err)
    assert delta <= TOL_SISNR_DB, (name, delta)
    return err, delta


@pytest.mark.parametrize('name,cls,ctor,B,L,chained', CASES, ids=[c[0] for c in CASES])
def test_reference_model_classes_cpu_original_vs_cuda_dropin(name, cls, ctor, B, L, chained, built_lib):
    orig, drop = ref_loader.load_original(), ref_loader.load_dropin()
    args, kw = ctor
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        ref = getattr(orig, cls)(*args, **kw).eval()
        sd = seeded_state_dict(ref, seed=21)
        ref.load_state_dict(sd)
        model = getattr(drop, cls)(*args, **kw).eval()
        assert type(model).__module__ == type(ref).__module__           # the same unmodified model file ...
        assert type(model) is not type(ref)                             # ... bound to different hot-path modules
        assert type(model.stft).__module__ == 'speech_enhancement_b200.conv_stft'
        assert type(ref.stft).__module__ == 'conv_stft'
        model.load_state_dict(sd)
        model.to(DEV)
        x, clean = synth_noisy_speech(B, L, seed=22, return_clean=True)
        if name == 'dcunet20_large_test_py':
            x = torch.rand(1, 32000, generator=torch.Generator().manual_seed(3))   # test.py:15 draws U(0, 1)
            clean = x
        n0 = built_lib.se_launch_count()
        with torch.no_grad():
            out_ref = ref(x)
            out_mod = model(x.to(DEV))
        assert built_lib.se_launch_count() > n0                            # the CUDA library did the work
        istft = model.istft if cls == 'DCUNet' else None
        _compare(orig, out_mod, out_ref, clean, name + ' (module by module)', istft)
        if not chained:
            with pytest.raises(RuntimeError):
                dropin.accelerate(model)
            return
        dropin.accelerate(model)
        calls = []
        view = model._se_chained
        inner = view._forward_fused
        view.__dict__['_forward_fused'] = lambda inp: (calls.append(1), inner(inp))[1]
        with torch.no_grad():
            out_ch = model(x.to(DEV))
        assert calls == [1], 'accelerate(): the chained forward did not run'
        _compare(orig, out_ch, out_ref, clean, name + ' (chained)', istft)
        # chained vs module by module (for DCUNet with the same treatment of the branch-cut bins)
        _compare(orig, out_ch, (out_mod[0], out_mod[1].cpu()), clean, name + ' (chained vs modular)', istft)
        if istft is None:
            assert rel_l2(out_ch[1].cpu(), out_mod[1].cpu()) <= TOL_PATHS, name
        # outside eval / no_grad the reference's own forward runs, loudly
        with pytest.warns(RuntimeWarning, match='module-by-module'):
            _wav_of(model(x[:1, :8000].to(DEV)))
        dropin.decelerate(model)
        assert 'forward' not in model.__dict__
