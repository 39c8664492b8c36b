"""Pins the CPU oracle against vectors produced by the unmodified reference
(tests/golden/make_golden.py).  Runs on CPU; no /root/reference needed."""
import numpy as np
import pytest
import torch

import oracle
from oracle.modules_ref import OPS, ComplexConv2dRef, ComplexConvTranspose2dRef
from speech_enhancement_b200 import callers
from speech_enhancement_b200.synth import seeded_state_dict, synth_noisy_speech
from conftest import golden, rand, rel_l2
from cases import STFT_CASES, CONV_CASES

TOL = 2e-6   # oracle and reference call the same ATen kernels; anything above fp32 noise is a bug


@pytest.mark.parametrize('i', range(len(STFT_CASES)))
def test_stft_istft_vs_reference(i):
    g = golden('stft')
    win, hop, fft, L = STFT_CASES[i]
    assert list(g[f'{i}_cfg'][:4]) == [win, hop, fft, L]
    x = rand((2, L), 100 + i, 0.1)
    w, _ = oracle.init_kernels_ref(win, fft)
    winv, window = oracle.init_kernels_ref(win, fft, inverse=True)
    for nm, buf in (('fwd', w), ('inv', winv), ('window', window)):
        flat = buf.reshape(-1).double()
        np.testing.assert_allclose(flat[::97].float().numpy(), g[f'{i}_{nm}_samples'], rtol=0, atol=2e-7)
        sums = np.array([flat.sum().item(), flat.abs().sum().item(), (flat * flat).sum().item()])
        np.testing.assert_allclose(sums, g[f'{i}_{nm}_sums'], rtol=1e-6, atol=1e-6)
    spec = oracle.stft_ref(x, w, hop, fft // 2)
    assert spec.shape == g[f'{i}_spec'].shape
    assert rel_l2(spec, g[f'{i}_spec']) < TOL
    wav = oracle.istft_ref(torch.from_numpy(g[f'{i}_spec']), winv, window, hop, fft // 2)
    assert wav.shape == g[f'{i}_wav'].shape
    assert rel_l2(wav, g[f'{i}_wav']) < TOL
    wav2 = oracle.istft_ref(torch.from_numpy(g[f'{i}_spec']), winv, window, hop, fft // 2, output_length=L - 7)
    assert wav2.shape == g[f'{i}_wav_len'].shape and rel_l2(wav2, g[f'{i}_wav_len']) < TOL
    if f'{i}_mags' in g:
        mags, phase = oracle.stft_magphase_ref(x, w, hop, fft // 2)
        assert rel_l2(mags, g[f'{i}_mags']) < TOL
        # phase compared through the unit vector (atan2 branch cut)
        d = torch.remainder(phase - torch.from_numpy(g[f'{i}_phase']) + np.pi, 2 * np.pi) - np.pi
        assert (d.abs() * (mags > 1e-3)).max() < 1e-3
        wavp = oracle.istft_ref(torch.from_numpy(g[f'{i}_mags']), winv, window, hop, fft // 2,
                                phase=torch.from_numpy(g[f'{i}_phase']))
        assert rel_l2(wavp, g[f'{i}_wav_polar']) < TOL


def test_stft_not_centered_and_1d_input():
    g = golden('stft')
    x = rand((3000,), 77, 0.1)
    w, _ = oracle.init_kernels_ref(400, 512)
    winv, window = oracle.init_kernels_ref(400, 512, inverse=True)
    spec = oracle.stft_ref(x, w, 100, 256, center=False)
    assert spec.shape == g['nc_spec'].shape and rel_l2(spec, g['nc_spec']) < TOL
    wav = oracle.istft_ref(spec, winv, window, 100, 256, center=False)
    assert wav.shape == g['nc_wav'].shape and rel_l2(wav, g['nc_wav']) < TOL


@pytest.mark.parametrize('i', range(len(CONV_CASES)))
def test_cconv_vs_reference(i):
    g = golden('cconv')
    name, transposed, cin, cout, k, kw, H, W, B = CONV_CASES[i]
    kw = dict(kw)
    kw.setdefault('bias', False)
    m = (ComplexConvTranspose2dRef if transposed else ComplexConv2dRef)(cin, cout, k, **kw)
    sd = {key: rand(v.shape, 1000 + 10 * i + j, 0.2) for j, (key, v) in enumerate(sorted(m.state_dict().items()))}
    m.load_state_dict(sd)
    x = rand((B, cin, H, W), 5000 + i)
    with torch.no_grad():
        y = m(x)
    assert y.shape == g[f'{name}_y'].shape
    assert rel_l2(y, g[f'{name}_y']) < TOL
    # packed single-GEMM form == 4-conv form, checked in fp64
    import torch.nn.functional as F
    rc, ic = m.real_conv, m.imag_conv
    xd = x.double()
    if transposed:
        wp = oracle.pack_convT_weight_ref(rc.weight.double(), ic.weight.double())
        yp = F.conv_transpose2d(xd, wp, None, rc.stride, rc.padding, rc.output_padding)
    else:
        wp = oracle.pack_conv_weight_ref(rc.weight.double(), ic.weight.double())
        yp = F.conv2d(xd, wp, None, rc.stride, rc.padding)
    if rc.bias is not None:
        bp = torch.cat([rc.bias - ic.bias, rc.bias + ic.bias]).double()
        yp = yp + bp.view(1, -1, 1, 1)
    assert rel_l2(yp, g[f'{name}_y']) < TOL


def test_masks_vs_reference():
    g = golden('masks')
    B, F, T = 2, 9, 21
    nr, ni, mr, mi = (rand((B, F, T), 900 + j, s) for j, s in enumerate((0.5, 0.5, 1.5, 1.5)))
    for mode in 'RCE':
        assert rel_l2(oracle.mask_dccrn_ref(mode, nr, ni, mr, mi), g[f'dccrn_{mode}']) < TOL
    for method in ('unbounded', 'bounded_sigmoid', 'bounded_tanh'):
        assert rel_l2(oracle.mask_dcunet_ref(method, nr, ni, mr, mi), g[f'dcunet_{method}']) < TOL
    for name, fn in (('carn', oracle.mask_carn_ref), ('frcrn', oracle.mask_frcrn_ref)):
        spec = torch.from_numpy(g[f'{name}_noisy_spec'])
        net = torch.from_numpy(g[f'{name}_net'])
        f = spec.shape[1] // 2
        est = fn(spec[:, :f], spec[:, f:], net[:, 0], net[:, 1])
        assert est.shape == g[f'{name}_est_spec'].shape
        assert rel_l2(est, g[f'{name}_est_spec']) < TOL


MODELS = [('dccrn_cl', callers.DCCRN, ('dccrn-CL', 400, 100, 512)),
          ('dccrn_c', callers.DCCRN, ('dccrn-C', 400, 100, 512)),
          ('dcunet20', callers.DCUNet, ('dcunet20', 512, 128, 512)),
          ('dcunet10', callers.DCUNet, ('dcunet10', 512, 128, 512)),
          ('frcrn', callers.FRCRN, (320, 160, 640))]


@pytest.mark.parametrize('name,cls,args', MODELS)
def test_model_hosts_vs_reference(name, cls, args):
    """caller hosts + oracle ops reproduce the reference model forward; state_dict keys identical"""
    g = golden('models')
    model = cls(*args, ops=OPS)
    assert sorted(model.state_dict().keys()) == list(g[f'{name}_keys'])
    model.load_state_dict(seeded_state_dict(model, seed=7))
    model.eval()
    x = synth_noisy_speech(2, int(g[f'{name}_L'][0]), seed=11)
    with torch.no_grad():
        spec, wav = model(x)
    assert wav.shape == g[f'{name}_wav'].shape
    assert rel_l2(wav, g[f'{name}_wav']) < 2e-5
    assert rel_l2(spec.reshape(-1)[::53], g[f'{name}_spec_samples']) < 2e-5
