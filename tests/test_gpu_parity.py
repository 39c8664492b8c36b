"""Parity of the CUDA path (through the C ABI) against the committed reference vectors and the
CPU oracle.  Tolerances (fp32 path, stated per SURVEY.md 8c):
  module level  rel-L2 <= 1e-5 for the exact-fp32 SIMT engine, <= 2e-4 for the bf16x3 tensor engine
  model level   enhanced-waveform rel-L2 <= 1e-3 (north-star bar), we assert 3e-4
"""
import numpy as np
import pytest
import torch

import oracle
from oracle.modules_ref import OPS
import speech_enhancement_b200 as se
from speech_enhancement_b200 import _abi, callers, functional as SF
from speech_enhancement_b200.synth import seeded_state_dict, synth_noisy_speech
from conftest import golden, rand, rel_l2
from cases import STFT_CASES, CONV_CASES

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
TOL_EXACT = 1e-5
# exact-fp32 SIMT kernels / bf16x3 tensor-core framing / fp32 shared-memory FFT (fft 512 only)
TOL_STFT = {'simt': 1e-5, 'umma': 3e-5, 'fft': 1e-5}
ENGINES = ['simt', 'umma', 'fft']


def _skip_unless_fft_covers(engine, win, hop, fft):
    if engine == 'fft' and _abi.lib().se_stft_fft_supported(win, hop, fft) != 1:
        pytest.skip(f'FFT engine does not cover {win}/{hop}/{fft}')


def _stft_pair(engine, *args, **kw):
    st, ist = se.ConvSTFT(*args, **kw).to(DEV), se.ConviSTFT(*args, **kw).to(DEV)
    st.engine = ist.engine = engine
    return st, ist


@pytest.fixture(autouse=True)
def _lib(built_lib):
    assert built_lib.se_device_ok() == 1, 'these tests need a compute-capability 10.x device'


@pytest.mark.parametrize('engine', ENGINES)
@pytest.mark.parametrize('i', range(len(STFT_CASES)))
def test_stft_istft_golden(i, engine):
    g = golden('stft')
    tol = TOL_STFT[engine]
    win, hop, fft, L = STFT_CASES[i]
    _skip_unless_fft_covers(engine, win, hop, fft)
    x = rand((2, L), 100 + i, 0.1).to(DEV)
    st, ist = _stft_pair(engine, win, hop, fft)
    with torch.no_grad():
        spec = st(x)
        assert spec.shape == g[f'{i}_spec'].shape
        assert rel_l2(spec.cpu(), g[f'{i}_spec']) < tol
        gs = torch.from_numpy(g[f'{i}_spec']).to(DEV)
        wav = ist(gs)
        assert wav.shape == g[f'{i}_wav'].shape and rel_l2(wav.cpu(), g[f'{i}_wav']) < tol
        wav2 = ist(gs, output_length=L - 7)
        assert wav2.shape == g[f'{i}_wav_len'].shape and rel_l2(wav2.cpu(), g[f'{i}_wav_len']) < tol
        # round trip (size-independent property): istft(stft(x)) == x[:L']
        rt = ist(spec)
        assert rel_l2(rt.cpu(), x[:, :rt.shape[1]].cpu()) < tol
        if f'{i}_mags' in g:
            stm = se.ConvSTFT(win, hop, fft, return_mag_phase=True).to(DEV)
            stm.engine = engine
            mags, phase = stm(x)
            assert rel_l2(mags.cpu(), g[f'{i}_mags']) < tol
            gm, gp = torch.from_numpy(g[f'{i}_mags']).to(DEV), torch.from_numpy(g[f'{i}_phase']).to(DEV)
            d = torch.remainder(phase - gp + np.pi, 2 * np.pi) - np.pi
            # phase of a weak bin amplifies the spectrum error by 1/|bin|: looser for the bf16x3 engine
            assert (d.abs() * (gm > 1e-3)).max() < (1e-3 if engine == 'simt' else 5e-3)
            assert rel_l2(ist(gm, gp).cpu(), g[f'{i}_wav_polar']) < tol


@pytest.mark.parametrize('engine', ENGINES)
def test_stft_input_forms_and_not_centered(engine):
    g = golden('stft')
    tol = TOL_STFT[engine]
    x = rand((3000,), 77, 0.1).to(DEV)
    st, ist = _stft_pair(engine, 400, 100, 512, center=False)
    with torch.no_grad():
        spec = st(x)
        assert spec.shape == g['nc_spec'].shape and rel_l2(spec.cpu(), g['nc_spec']) < tol
        assert torch.equal(spec, st(x[None, None, :])) and torch.equal(spec, st(x[None, :]))
        wav = ist(spec)
        # without centring the first/last samples divide by window^2 ~ 0 (+1e-8): ill-conditioned edges
        assert wav.shape == g['nc_wav'].shape and rel_l2(wav.cpu(), g['nc_wav']) < (1e-4 if engine == 'simt' else 2e-3)
        assert rel_l2(wav[:, 100:-100].cpu(), g['nc_wav'][:, 100:-100]) < tol
    with pytest.raises(RuntimeError):
        se.ConvSTFT(400, 100, 512).to(DEV)(torch.zeros(1, 100, device=DEV))     # reflect pad > signal


@pytest.mark.parametrize('win,hop,fft,L,B', [(400, 100, 512, 64000, 8), (512, 128, 512, 64000, 8),
                                             (320, 160, 640, 64000, 4), (400, 100, 512, 160000, 2)])
@pytest.mark.parametrize('engine', ENGINES)
def test_stft_full_size_vs_oracle(win, hop, fft, L, B, engine):
    tol = TOL_STFT[engine]
    _skip_unless_fft_covers(engine, win, hop, fft)
    x = synth_noisy_speech(B, L, seed=3)
    st, ist = _stft_pair(engine, win, hop, fft)
    with torch.no_grad():
        spec = st(x.to(DEV))
        spec_ref = oracle.stft_ref(x, st.weight.cpu(), hop, fft // 2)
        assert spec.shape == spec_ref.shape and rel_l2(spec.cpu(), spec_ref) < tol
        wav = ist(spec)
        wav_ref = oracle.istft_ref(spec_ref, ist.weight.cpu(), ist.window.cpu(), hop, fft // 2)
        assert wav.shape == wav_ref.shape and rel_l2(wav.cpu(), wav_ref) < tol
        assert rel_l2(wav.cpu(), x[:, :wav.shape[1]]) < tol


@pytest.mark.parametrize('win,hop,L,B,center', [(400, 100, 64000, 3, True), (512, 128, 37001, 2, True),
                                                (320, 160, 5000, 5, True), (400, 100, 1311, 1, False),
                                                (512, 64, 9000, 2, True), (37, 11, 2000, 2, True),
                                                (400, 100, 700, 2, True), (512, 256, 8192, 2, False)])
def test_stft_fft_engine_matches_exact_simt(win, hop, L, B, center):
    """the FFT kernels against the exact-fp32 dense projection on ragged sizes: odd frame counts, clips shorter
    than a tile, hop == win, tiny windows, eight overlapping frames, output_length crops"""
    x = rand((B, L), 5, 0.3).to(DEV)
    st, ist = _stft_pair('fft', win, hop, 512, center=center)
    st0, ist0 = _stft_pair('simt', win, hop, 512, center=center)
    with torch.no_grad():
        spec, spec0 = st(x), st0(x)
        assert spec.shape == spec0.shape and rel_l2(spec.cpu(), spec0.cpu()) < 2e-6
        assert (spec[:, 257] == 0).all() and (spec[:, 513] == 0).all()       # Im of DC and Nyquist
        sp = spec0 + 0.05 * torch.randn_like(spec0)                           # not a consistent STFT
        for kw in ({}, {'output_length': L - 5}, {'output_length': 17}):
            wav, wav0 = ist(sp, **kw), ist0(sp, **kw)
            assert wav.shape == wav0.shape
            interior = slice(hop, -hop) if not center and wav.shape[1] > 3 * hop else slice(None)
            assert rel_l2(wav[:, interior].cpu(), wav0[:, interior].cpu()) < (5e-6 if center else 2e-5)


def test_stft_fft_engine_only_for_the_dft_basis():
    """`engine='auto'` takes the FFT kernels only when `weight` is the closed-form basis: a checkpoint with any other
    basis keeps the dense projection (and its result), `engine='fft'` refuses it, fft sizes other than 512 too"""
    x = rand((2, 4000), 9, 0.1).to(DEV)
    st, ist = _stft_pair('auto', 400, 100, 512)
    assert st._pick_engine('ConvSTFT', st.weight[0, 0], False) == 'fft'
    assert ist._pick_engine('ConviSTFT', ist.window, True) == 'fft'
    with torch.no_grad():
        st.weight[40, 0, 7] += 1e-3
        ist.weight[3, 0, 100] += 1e-4
        assert st._pick_engine('ConvSTFT', st.weight[0, 0], False) == 'umma'
        assert ist._pick_engine('ConviSTFT', ist.window, True) == 'umma'
        spec = st(x)
        ref = oracle.stft_ref(x.cpu(), st.weight.cpu(), 100, 256)
        assert rel_l2(spec.cpu(), ref) < 3e-5
        wav = ist(spec)
        assert rel_l2(wav.cpu(), oracle.istft_ref(ref, ist.weight.cpu(), ist.window.cpu(), 100, 256)) < 3e-5
        st.engine = 'fft'
        with pytest.raises(RuntimeError):
            st(x)
    with pytest.raises(RuntimeError):
        _stft_pair('fft', 320, 160, 640)[0](x)


def _conv_case(i, engine):
    name, transposed, cin, cout, k, kw, H, W, B = CONV_CASES[i]
    kw = dict(kw)
    kw.setdefault('bias', False)
    m = (se.ComplexConvTranspose2d if transposed else se.ComplexConv2d)(cin, cout, k, **kw)
    sd = {key: rand(v.shape, 1000 + 10 * i + j, 0.2) for j, (key, v) in enumerate(sorted(m.state_dict().items()))}
    m.load_state_dict(sd)
    m.to(DEV)
    m.engine = engine
    x = rand((B, cin, H, W), 5000 + i).to(DEV)
    with torch.no_grad():
        return name, m, x, m(x)


@pytest.mark.parametrize('i', range(len(CONV_CASES)))
def test_cconv_golden_simt(i):
    g = golden('cconv')
    name, m, x, y = _conv_case(i, _abi.ENGINE_SIMT)
    assert y.shape == g[f'{name}_y'].shape
    assert rel_l2(y.cpu(), g[f'{name}_y']) < TOL_EXACT


@pytest.mark.parametrize('i', range(len(CONV_CASES)))
def test_cconv_golden_auto(i):
    """whatever engine AUTO picks must also meet the (looser, bf16x3) module tolerance"""
    g = golden('cconv')
    name, m, x, y = _conv_case(i, _abi.ENGINE_AUTO)
    assert rel_l2(y.cpu(), g[f'{name}_y']) < 2e-4


@pytest.mark.parametrize('i', range(len(CONV_CASES)))
def test_cconv_golden_umma(i):
    """every golden geometry forced through the tcgen05 bf16x3 engine (module tolerance 2e-4;
    the split keeps ~16 mantissa bits per operand so the observed error is ~1e-5)"""
    g = golden('cconv')
    name, m, x, y = _conv_case(i, _abi.ENGINE_UMMA)
    assert y.shape == g[f'{name}_y'].shape
    err = rel_l2(y.cpu(), g[f'{name}_y'])
    assert err < 2e-4, err


# (transposed, cin, cout, kernel, kwargs, H, W, B): real layer shapes of SURVEY.md 8a
LAYER_SHAPES = [
    (False, 32, 64, (5, 2), dict(stride=(2, 1), padding=(2, 0)), 128, 643, 2),          # DCCRN enc1
    (False, 128, 256, (5, 2), dict(stride=(2, 1), padding=(2, 0)), 32, 643, 1),         # DCCRN enc3
    (False, 256, 256, (5, 2), dict(stride=(2, 1), padding=(2, 0)), 8, 643, 2),          # DCCRN enc5
    (True, 512, 256, (5, 2), dict(stride=(2, 1), padding=(2, 0), output_padding=(1, 0)), 4, 642, 2),   # dec0
    (True, 256, 64, (5, 2), dict(stride=(2, 1), padding=(2, 0), output_padding=(1, 0)), 32, 642, 1),   # dec3
    (True, 64, 2, (5, 2), dict(stride=(2, 1), padding=(2, 0), output_padding=(1, 0)), 128, 642, 1),    # dec5
    (False, 64, 128, (7, 5), dict(stride=(2, 2), padding=(3, 2)), 257, 501, 1),         # DCUNet enc2
    (False, 128, 180, (5, 3), dict(stride=(2, 1), padding=(2, 1)), 3, 32, 3),           # DCUNet enc9 (N=192)
    (True, 360, 128, (5, 3), dict(stride=(2, 1), padding=(2, 1)), 2, 32, 3),            # DCUNet dec0
    (True, 256, 128, (5, 3), dict(stride=(2, 2), padding=(2, 1)), 9, 63, 2),            # DCUNet dec3
    (True, 128, 64, (1, 7), dict(stride=(1, 1), padding=(0, 3)), 65, 126, 1),           # DCUNet dec8-like
    (True, 256, 128, (5, 2), dict(stride=(2, 1), padding=(0, 0)), 37, 403, 1),          # FRCRN dec3
]


@pytest.mark.parametrize('case', range(len(LAYER_SHAPES)))
def test_cconv_layer_shapes_umma_vs_simt_and_oracle(case):
    transposed, cin, cout, k, kw, H, W, B = LAYER_SHAPES[case]
    m = (se.ComplexConvTranspose2d if transposed else se.ComplexConv2d)(cin, cout, k, bias=False, **kw)
    sd = {key: rand(v.shape, 300 + case * 7 + j, (2.0 / (cin * 5)) ** 0.5)
          for j, (key, v) in enumerate(sorted(m.state_dict().items()))}
    m.load_state_dict(sd)
    x_cpu = rand((B, cin, H, W), 7000 + case)
    m.to(DEV)
    x = x_cpu.to(DEV)
    with torch.no_grad():
        m.engine = _abi.ENGINE_SIMT
        y_simt = m(x)
        m.engine = _abi.ENGINE_UMMA
        y_umma = m(x)
        fn = oracle.cconvT2d_ref if transposed else oracle.cconv2d_ref
        args = dict(stride=kw['stride'], padding=kw['padding'])
        if transposed:
            args['output_padding'] = kw.get('output_padding', 0)
        y_ref = fn(x_cpu.double(), sd['real_conv.weight'].double(), sd['imag_conv.weight'].double(), **args)
    assert y_umma.shape == y_ref.shape
    assert rel_l2(y_simt.cpu(), y_ref) < 1e-5
    err = rel_l2(y_umma.cpu(), y_ref)
    assert err < 5e-5, err
    assert torch.isfinite(y_umma).all()


def test_cconv_container_forms_and_repack():
    m = se.ComplexConv2d(8, 8, (3, 2), padding=(1, 0), bias=False).to(DEV)
    x = rand((2, 8, 9, 7), 1).to(DEV)
    with torch.no_grad():
        y = m(x)
        yr, yi = m([x[:, :4], x[:, 4:]])
        assert torch.equal(torch.cat([yr, yi], 1), y)
        yc = m(torch.complex(x[:, :4], x[:, 4:]))
        assert torch.is_complex(yc) and torch.equal(torch.cat([yc.real, yc.imag], 1), y)
        # in-place weight update (optimizer.step / load_state_dict) must invalidate the packed cache
        m.real_conv.weight.mul_(2.0)
        m.imag_conv.weight.mul_(2.0)
        assert rel_l2(m(x).cpu(), (2 * y).cpu()) < 1e-6
    with pytest.raises(RuntimeError):
        se.ComplexConv2d(8, 8, 3, dilation=2).to(DEV)(x)


def test_masks_golden_and_strided_inputs():
    g = golden('masks')
    B, F, T = 2, 9, 21
    nr, ni, mr, mi = (rand((B, F, T), 900 + j, s).to(DEV) for j, s in enumerate((0.5, 0.5, 1.5, 1.5)))
    for mode in 'RCE':
        assert rel_l2(SF.mask_dccrn(mode, nr, ni, mr, mi).cpu(), g[f'dccrn_{mode}']) < TOL_EXACT
    for method in ('unbounded', 'bounded_sigmoid', 'bounded_tanh'):
        assert rel_l2(SF.mask_dcunet(method, nr, ni, mr, mi).cpu(), g[f'dcunet_{method}']) < TOL_EXACT
    for name, fn in (('carn', SF.mask_carn), ('frcrn', SF.mask_frcrn)):
        spec = torch.from_numpy(g[f'{name}_noisy_spec']).to(DEV)
        net = torch.from_numpy(g[f'{name}_net']).to(DEV)
        f = spec.shape[1] // 2
        est = fn(spec[:, :f], spec[:, f:], net[:, 0], net[:, 1])       # strided views, consumed in place
        assert est.shape == g[f'{name}_est_spec'].shape
        assert rel_l2(est.cpu(), g[f'{name}_est_spec']) < TOL_EXACT
    # full-size planes vs the oracle (DCCRN-E, 257 x 642)
    nr, ni, mr, mi = (rand((3, 257, 642), 50 + j, s) for j, s in enumerate((0.3, 0.3, 1.0, 1.0)))
    ref = oracle.mask_dccrn_ref('E', nr, ni, mr, mi)
    out = SF.mask_dccrn('E', nr.to(DEV), ni.to(DEV), mr.to(DEV), mi.to(DEV))
    assert rel_l2(out.cpu(), ref) < TOL_EXACT
    mags, ph = rand((2, 161, 50), 5).abs(), rand((2, 161, 50), 6)
    assert rel_l2(SF.polar_to_rect(mags.to(DEV), ph.to(DEV)).cpu(), oracle.polar_to_rect_ref(mags, ph)) < TOL_EXACT


def test_polar_masks_edge_values_match_the_atan2_route():
    """the polar masks are evaluated in rectangular form (unit vectors instead of atan2 -> sincos); exact zeros,
    signed zeros, denormal-scale and huge bins must still give what the reference's atan2 route gives"""
    vals = torch.tensor([0.0, -0.0, 1e-25, -1e-25, 3e-20, 1e-4, -0.3, 1.0, -2.5, 7.0, 1e18, -1e18])

    def planes(mvals_r, mvals_i):
        g = torch.stack(torch.meshgrid(vals, vals, mvals_r, mvals_i, indexing='ij'), 0).reshape(4, 1, -1, 1)
        return g.expand(4, 1, g.shape[2], 3).contiguous()

    # DCUNet's angle / magnitude quirk amplifies a 1-ulp atan2 difference by 1 / |mask|: keep its masks O(1)
    big = torch.tensor([-0.3, 1.0, -2.5, 7.0])
    for fn, ref, (nr, ni, mr, mi) in (
            (lambda *a: SF.mask_dccrn('E', *a), lambda *a: oracle.mask_dccrn_ref('E', *a), planes(vals[2:10], vals[1:9])),
            (lambda *a: SF.mask_dcunet('bounded_tanh', *a), lambda *a: OPS.mask_dcunet('bounded_tanh', *a), planes(big, big))):
        out = fn(nr.to(DEV), ni.to(DEV), mr.to(DEV), mi.to(DEV)).cpu()
        want = torch.as_tensor(ref(nr, ni, mr, mi))
        assert out.shape == want.shape and torch.isfinite(want).all() and torch.isfinite(out).all()
        # error of each complex bin relative to that bin's magnitude (a component that is ~0 by cancellation has no
        # relative accuracy of its own in either route)
        f = want.shape[1] // 2
        mag = (want[:, :f] ** 2 + want[:, f:] ** 2).sqrt()
        err = ((out[:, :f] - want[:, :f]) ** 2 + (out[:, f:] - want[:, f:]) ** 2).sqrt() / (mag + 1e-9)
        assert err.max() < 2e-5, err.max()


MODELS = [('dccrn_cl', callers.DCCRN, ('dccrn-CL', 400, 100, 512)),
          ('dccrn_c', callers.DCCRN, ('dccrn-C', 400, 100, 512)),
          ('dcunet20', callers.DCUNet, ('dcunet20', 512, 128, 512)),
          ('dcunet10', callers.DCUNet, ('dcunet10', 512, 128, 512)),
          ('frcrn', callers.FRCRN, (320, 160, 640))]


@pytest.mark.parametrize('name,cls,args', MODELS)
def test_models_vs_reference_golden(name, cls, args):
    g = golden('models')
    model = cls(*args)
    model.load_state_dict(seeded_state_dict(model, seed=7))
    model.eval().to(DEV)
    x = synth_noisy_speech(2, int(g[f'{name}_L'][0]), seed=11).to(DEV)
    with torch.no_grad():
        spec, wav = model(x)
    assert wav.shape == g[f'{name}_wav'].shape
    err = rel_l2(wav.cpu(), g[f'{name}_wav'])
    assert err < 3e-4, err


def si_snr_db(est, target, eps=1e-8):
    """losses.py:62-84 restated (per-utterance SI-SNR in dB, mean over the batch)."""
    dot = (est * target).sum(1, keepdim=True)
    s = dot * target / target.pow(2).sum(1, keepdim=True)
    return (10 * torch.log10((s.pow(2).sum(1) + eps) / ((est - s).pow(2).sum(1) + eps))).mean().item()


@pytest.mark.parametrize('name,cls,args,B', [('dccrn_cl', callers.DCCRN, ('dccrn-CL', 400, 100, 512), 2),
                                             ('dcunet20', callers.DCUNet, ('dcunet20', 512, 128, 512), 1),
                                             ('frcrn', callers.FRCRN, (320, 160, 640), 1)])
def test_models_full_length_vs_oracle(name, cls, args, B):
    """4 s utterances (BASELINE configs 1-3 shapes) against the CPU oracle: rel-L2 <= 1e-3 and
    SI-SNR delta <= 0.01 dB measured against the oracle's own fp64 output as the target."""
    ref = cls(*args, ops=OPS).eval()
    sd = seeded_state_dict(ref, seed=0)
    ref.load_state_dict(sd)
    ours = cls(*args).eval()
    ours.load_state_dict(sd)
    ours.to(DEV)
    x = synth_noisy_speech(B, 64000, seed=4)
    with torch.no_grad():
        _, wav_ref = ref(x)
        _, wav64 = ref.double()(x.double())
        _, wav = ours(x.to(DEV))
    wav = wav.cpu()
    assert wav.shape == wav_ref.shape
    err = rel_l2(wav, wav_ref)
    assert err < 1e-3, err
    tgt = wav64.float()
    delta = abs(si_snr_db(wav, tgt) - si_snr_db(wav_ref, tgt))
    # both are ~100+ dB against the fp64 target; the bar is on SI-SNR against a correlated target:
    mix = tgt + 0.1 * tgt.std() * torch.randn(tgt.shape, generator=torch.Generator().manual_seed(1))
    delta = abs(si_snr_db(wav, mix) - si_snr_db(wav_ref, mix))
    assert delta < 0.01, delta


def test_dccrn_fused_chain_matches_modular_path_and_oracle():
    """eval/no_grad DCCRN runs the chained engine (BN + PReLU + causal pad + skip concat folded into the
    conv kernels, CL2 between layers); it must agree with the module-by-module path and the CPU oracle."""
    args = ('dccrn-CL', 400, 100, 512)
    ref = callers.DCCRN(*args, ops=OPS).eval()
    sd = seeded_state_dict(ref, seed=5)
    ref.load_state_dict(sd)
    ours = callers.DCCRN(*args).eval()
    ours.load_state_dict(sd)
    ours.to(DEV)
    x = synth_noisy_speech(3, 24000, seed=9)
    with torch.no_grad():
        _, wav_ref = ref(x)
        assert ours._use_fused(x.to(DEV))
        spec_f, wav_f = ours(x.to(DEV))
        ours.fused = False
        spec_m, wav_m = ours(x.to(DEV))
    assert wav_f.shape == wav_m.shape == wav_ref.shape
    assert rel_l2(wav_f.cpu(), wav_m.cpu()) < 1e-4
    assert rel_l2(spec_f.cpu(), spec_m.cpu()) < 1e-4
    assert rel_l2(wav_f.cpu(), wav_ref) < 3e-4
    assert rel_l2(wav_m.cpu(), wav_ref) < 3e-4


@pytest.mark.parametrize('cls,args,L', [(callers.DCUNet, ('dcunet20', 512, 128, 512), 20000),
                                        (callers.DCUNet, ('dcunet10', 512, 128, 512), 16384),
                                        (callers.DCUNet, ('dcunet16', 400, 100, 512), 16384),
                                        (callers.FRCRN, (320, 160, 640), 16000)])
def test_dcunet_frcrn_fused_chain_matches_modular_path_and_oracle(cls, args, L):
    """the chained engine for the other two complex families: plain torch.cat skips with the F.pad to the
    skip's size (DCUNet) and CCBAM-attended complex_concat skips (FRCRN)."""
    ref = cls(*args, ops=OPS).eval()
    sd = seeded_state_dict(ref, seed=6)
    ref.load_state_dict(sd)
    ours = cls(*args).eval()
    ours.load_state_dict(sd)
    ours.to(DEV)
    x = synth_noisy_speech(2, L, seed=10)
    with torch.no_grad():
        spec_ref, wav_ref = ref(x)
        calls, chained = [], ours._forward_fused
        ours._forward_fused = lambda inp: (calls.append(1), chained(inp))[1]
        spec_f, wav_f = ours(x.to(DEV))
        ours.fused = False
        spec_m, wav_m = ours(x.to(DEV))
        assert len(calls) == 1                      # the first call took the chained path, the second did not
        assert wav_f.shape == wav_m.shape == wav_ref.shape

        # DCUNet's bounded-tanh mask divides the atan2 ANGLE by the mask magnitude (_1903_03107_dcunet.py:177), so it
        # jumps where the mask crosses the negative real axis: a bin whose mask lies within rounding of that cut may
        # land on either side (true of the reference against itself at a different summation order).  Such bins are
        # counted (at most 1e-4 of all bins), taken out, and everything else is held to the tight tolerances.
        def agree(spec_a, wav_a, spec_b, wav_b, tol):
            spec_b = torch.as_tensor(spec_b).to(spec_a.device)
            wav_b = torch.as_tensor(wav_b).to(spec_a.device)
            cut = (spec_a - spec_b).abs() > 2e-3 * spec_b.abs().max()
            assert cut.sum().item() <= 1e-4 * cut.numel()
            spec_fix = torch.where(cut, spec_b, spec_a)
            assert rel_l2(spec_fix.cpu(), spec_b.cpu()) < tol
            assert rel_l2(torch.clamp(ours.istft(spec_a.contiguous()), -1, 1).cpu(), wav_a.cpu()) < 1e-6
            assert rel_l2(torch.clamp(ours.istft(spec_fix.contiguous()), -1, 1).cpu(), wav_b.cpu()) < tol

        agree(spec_f, wav_f, spec_m, wav_m, 2e-4)
        agree(spec_f, wav_f, spec_ref, wav_ref, 3e-4)
        agree(spec_m, wav_m, spec_ref, wav_ref, 3e-4)


@pytest.mark.parametrize('B,C,H,W', [(2, 128, 37, 53), (3, 64, 9, 70), (1, 32, 5, 3)])
def test_ccbam_on_cl2_matches_the_module_path(B, C, H, W):
    """csrc/ccbam.cu (statistics, channel attention MLP, spatial pooling, 7x7 attention conv with folded BN, apply)
    against the module's own NCHW forward (ccbam.py:28-106) and the CPU oracle modules"""
    from speech_enhancement_b200 import chain
    m = callers.CCBAM(C, 16).eval()
    sd = seeded_state_dict(m, seed=12)
    m.load_state_dict(sd)
    ref = callers.CCBAM(C, 16, ops=OPS).eval()
    ref.load_state_dict(sd)
    x = rand((B, C, H, W), 77, 1.0)
    with torch.no_grad():
        want = ref(x)
        m.to(DEV)
        y_mod = m(x.to(DEV))
        m.forward_cl2(chain.cl2_from_nchw(x.to(DEV)))           # builds the cached folded operands
        n0 = _abi.lib().se_launch_count()
        y = chain.cl2_to_nchw(m.forward_cl2(chain.cl2_from_nchw(x.to(DEV))))
        assert _abi.lib().se_launch_count() - n0 == 7            # 2 layout conversions + the 5 attention kernels
    assert rel_l2(y_mod.cpu(), want) < TOL_EXACT
    assert rel_l2(y.cpu(), want) < 2e-5                          # CL2 carries ~16 mantissa bits


@pytest.mark.parametrize('B,C,F_,T', [(2, 256, 4, 37), (3, 128, 2, 50), (1, 32, 3, 5), (2, 80, 4, 9), (1, 16, 2, 300), (2, 24, 4, 7)])
def test_cl2_from_lstm_rows_is_the_reference_permute(B, C, F_, T):
    """[B, T, C*F] rows -> CL2 == permute(0, 2, 1).reshape(B, C, F, T) (dccrn.py:164-165) followed by the layout conversion"""
    from speech_enhancement_b200 import chain
    rows = rand((B, T, C * F_), 31, 1.0).to(DEV)
    want = chain.cl2_from_nchw(rows.permute(0, 2, 1).reshape(B, C, F_, T).contiguous())
    got = chain.cl2_from_lstm_rows(rows, B, C, F_, T)
    assert got.data.shape == want.data.shape and torch.equal(got.data, want.data)


def test_host_pipeline_overlapped_copies_give_the_same_waveforms():
    """HostPipeline (H2D / forward / D2H of neighbouring batches on three streams) == model(x)[1], batch by batch"""
    from speech_enhancement_b200.pipeline import HostPipeline
    m = callers.DCCRN('dccrn-CL', 400, 100, 512).eval()
    m.load_state_dict(seeded_state_dict(m, seed=3))
    m.to(DEV)
    xs = [synth_noisy_speech(3, 12000, seed=40 + i).pin_memory() for i in range(5)]
    pipe = HostPipeline(m)
    n = pipe.output_length(xs[0])
    outs = [torch.empty((3, n)).pin_memory() for _ in xs]
    for x, o in zip(xs, outs):
        pipe.submit(x, o)
    pipe.synchronize()
    with torch.no_grad():
        for x, o in zip(xs, outs):
            assert torch.equal(o, m(x.to(DEV))[1].cpu())
    with pytest.raises(RuntimeError):
        pipe.submit(torch.zeros(3, 12000), outs[0])               # not pinned


def test_cl2_roundtrip_and_two_source_block():
    from speech_enhancement_b200 import chain
    x = rand((2, 40, 9, 37), 3).to(DEV)
    act = chain.cl2_from_nchw(x)
    assert act.data.shape == (2, 2, 9, 37, 64)
    assert rel_l2(chain.cl2_to_nchw(act).cpu(), x.cpu()) < 1e-5       # hi + lo keeps ~16 mantissa bits
    # decoder-style block: complex_concat([x, skip]) -> convT, against the module on the concatenated input
    m = se.ComplexConvTranspose2d(80, 24, (5, 2), stride=(2, 1), padding=(2, 0), output_padding=(1, 0), bias=True)
    sd = {k: rand(v.shape, 40 + j, 0.1) for j, (k, v) in enumerate(sorted(m.state_dict().items()))}
    m.load_state_dict(sd)
    m.to(DEV)
    skip = rand((2, 40, 9, 37), 4).to(DEV)
    with torch.no_grad():
        y_ref = m(se.complex_concat([x, skip], dim=1))
        blk = chain.FusedBlock(m, None, torch.nn.LeakyReLU(0.2), src_channels=[40, 40], concat='complex')
        act_y, y = blk([act, chain.cl2_from_nchw(skip)], out_cl2=True, out_nchw=True)
    y_ref = torch.nn.functional.leaky_relu(y_ref, 0.2)
    assert y.shape == y_ref.shape and rel_l2(y.cpu(), y_ref.cpu()) < 5e-5
    assert rel_l2(chain.cl2_to_nchw(act_y).cpu(), y_ref.cpu()) < 5e-5


@pytest.mark.parametrize('proj', ['torch', 'umma'])
@pytest.mark.parametrize('N,T,cin,hid,layers,batch_first', [(6, 37, 64, 256, 1, True), (3, 20, 128, 256, 2, True),
                                                            (5, 11, 32, 128, 1, False), (300, 9, 64, 256, 1, True),
                                                            (51, 7, 64, 256, 1, True), (31, 7, 64, 256, 2, True)])
def test_complex_lstm_recurrent_kernel_vs_torch(N, T, cin, hid, layers, batch_first, proj):
    """csrc/lstm_rec.cu (+ the input projections on the GEMM engine for proj='umma') against torch's own
    fp32 LSTM on the CPU (the reference's arithmetic) and against cuDNN."""
    m = se.ComplexLSTM(cin, hid, num_layers=layers, batch_first=batch_first)
    sd = {k: rand(v.shape, 4000 + j, 0.3) for j, (k, v) in enumerate(sorted(m.state_dict().items()))}
    m.load_state_dict(sd)
    shape = (N, T, cin) if batch_first else (T, N, cin)
    x = rand(shape, 4100)
    tol = TOL_EXACT if proj == 'torch' else 3e-5
    with torch.no_grad():
        ref = m(x)                                   # CPU: stock nn.LSTM
        m.to(DEV)
        m.proj_engine = proj
        n0 = _abi.lib().se_launch_count()
        y = m(x.to(DEV))
        assert _abi.lib().se_launch_count() - n0 >= layers         # the kernels really ran
        m.rec_kernel = False
        y_cudnn = m(x.to(DEV))
    assert y.shape == ref.shape
    assert rel_l2(y.cpu(), ref) < tol
    # cuDNN's own LSTM sits ~1.4e-5 from the CPU result on the 2-layer case (fast-math gates)
    assert rel_l2(y.cpu(), y_cudnn.cpu()) < 5e-5


@pytest.mark.parametrize('M,cin,cout', [(7, 256, 1024), (1284, 64, 96), (40, 520, 36)])
def test_complex_linear_tensor_engine(M, cin, cout):
    m = se.ComplexLinear(cin, cout)
    sd = {k: rand(v.shape, 4200 + j, 0.2) for j, (k, v) in enumerate(sorted(m.state_dict().items()))}
    m.load_state_dict(sd)
    x = rand((3, M, cin), 4300)
    with torch.no_grad():
        ref = m(x)
        m.to(DEV)
        y = m(x.to(DEV))
        m.engine = 'torch'
        y_t = m(x.to(DEV))
    assert y.shape == ref.shape and rel_l2(y.cpu(), ref) < 3e-5 and rel_l2(y_t.cpu(), ref) < TOL_EXACT


# ----------------------------------------------------------------------------------- autograd
def _grads(fn, inputs, seed):
    """d/d inputs of sum(fn(*inputs) * fixed random weights)"""
    inputs = [t.clone().requires_grad_(True) for t in inputs]
    out = fn(*inputs)
    outs = out if isinstance(out, (tuple, list)) else [out]
    loss = 0
    for i, o in enumerate(outs):
        w = rand(tuple(o.shape), seed + i).to(o.device)
        loss = loss + (o * w).sum()
    loss.backward()
    return [t.grad.detach().cpu() for t in inputs], [o.detach().cpu() for o in outs]


@pytest.mark.parametrize('win,hop,fft,L,center', [(400, 100, 512, 4000, True), (512, 128, 512, 4096, True),
                                                  (320, 160, 640, 3200, True), (320, 160, 512, 3360, True),
                                                  (400, 100, 512, 3000, False)])
def test_stft_istft_backward_vs_oracle(win, hop, fft, L, center):
    """the autograd nodes of ConvSTFT / ConviSTFT (adjoint GEMMs on the conv engine) against torch autograd of
    the CPU oracle (= the reference's F.conv1d / conv_transpose1d graph)."""
    x = rand((3, L), 50, 0.1)
    st, ist = se.ConvSTFT(win, hop, fft, center=center).to(DEV), se.ConviSTFT(win, hop, fft, center=center).to(DEV)
    st_o, ist_o = OPS.ConvSTFT(win, hop, fft, center=center), OPS.ConviSTFT(win, hop, fft, center=center)
    (gx,), (spec,) = _grads(st, [x.to(DEV)], 60)
    (gx_o,), (spec_o,) = _grads(st_o, [x], 60)
    assert rel_l2(spec, spec_o) < 3e-5 and rel_l2(gx, gx_o) < 3e-5
    (gs,), (wav,) = _grads(ist, [spec_o.to(DEV)], 70)
    (gs_o,), (wav_o,) = _grads(ist_o, [spec_o], 70)
    assert wav.shape == wav_o.shape and rel_l2(gs, gs_o) < 3e-5
    (gs2,), _ = _grads(lambda s_: ist(s_, output_length=L - 7), [spec_o.to(DEV)], 71)
    (gs2_o,), _ = _grads(lambda s_: ist_o(s_, output_length=L - 7), [spec_o], 71)
    assert rel_l2(gs2, gs2_o) < 3e-5
    # polar branches (CRN): mags/phase out of the STFT and into the iSTFT
    stp, stp_o = se.ConvSTFT(win, hop, fft, center=center, return_mag_phase=True).to(DEV), \
        OPS.ConvSTFT(win, hop, fft, center=center, return_mag_phase=True)
    (gp,), _ = _grads(lambda x_: stp(x_)[0], [x.to(DEV)], 80)
    (gp_o,), _ = _grads(lambda x_: stp_o(x_)[0], [x], 80)
    assert rel_l2(gp, gp_o) < 1e-4
    mags, ph = rand((2, fft // 2 + 1, 9), 81).abs() + 0.1, rand((2, fft // 2 + 1, 9), 82)
    g_mp, _ = _grads(lambda m_, p_: ist(m_, p_), [mags.to(DEV), ph.to(DEV)], 83)
    g_mp_o, _ = _grads(lambda m_, p_: ist_o(m_, p_), [mags, ph], 83)
    assert rel_l2(g_mp[0], g_mp_o[0]) < 3e-5 and rel_l2(g_mp[1], g_mp_o[1]) < 3e-5


def test_mask_backward_all_kinds_vs_oracle():
    """se_mask_bwd against torch autograd of the oracle's pointwise graph, strided inputs included."""
    B, F_, T = 2, 33, 41
    nr, ni, mr, mi = (rand((B, F_, T), 90 + i) for i in range(4))
    cases = [(lambda *a, m=m: SF.mask_dccrn(m, *a), lambda *a, m=m: OPS.mask_dccrn(m, *a)) for m in 'RCE']
    cases += [(lambda *a, m=m: SF.mask_dcunet(m, *a), lambda *a, m=m: OPS.mask_dcunet(m, *a))
              for m in ('unbounded', 'bounded_sigmoid', 'bounded_tanh')]
    cases += [(SF.mask_carn, OPS.mask_carn)]
    for ours, ref in cases:
        g, (o,) = _grads(ours, [t.to(DEV) for t in (nr, ni, mr, mi)], 95)
        g_o, (o_o,) = _grads(ref, [nr, ni, mr, mi], 95)
        assert rel_l2(o, o_o) < TOL_EXACT
        for a, b in zip(g, g_o):
            assert rel_l2(a, b) < 2e-5
    g, _ = _grads(SF.mask_frcrn, [nr.to(DEV), ni.to(DEV), mr[:, 2:].to(DEV), mi[:, 2:].to(DEV)], 96)
    g_o, _ = _grads(OPS.mask_frcrn, [nr, ni, mr[:, 2:].contiguous(), mi[:, 2:].contiguous()], 96)
    for a, b in zip(g, g_o):
        assert rel_l2(a, b) < 2e-5
    # planes that are slices of a bigger tensor (as in the models): grads flow back to the parent
    spec = rand((B, 2 * F_, T), 97).to(DEV).requires_grad_(True)
    net = rand((B, 2, F_, T), 98).to(DEV).requires_grad_(True)
    SF.mask_carn(spec[:, :F_], spec[:, F_:], net[:, 0], net[:, 1]).square().sum().backward()
    spec_c, net_c = spec.detach().cpu().requires_grad_(True), net.detach().cpu().requires_grad_(True)
    OPS.mask_carn(spec_c[:, :F_], spec_c[:, F_:], net_c[:, 0], net_c[:, 1]).square().sum().backward()
    assert rel_l2(spec.grad.cpu(), spec_c.grad) < 2e-5 and rel_l2(net.grad.cpu(), net_c.grad) < 2e-5


def test_carn_train_step_matches_cpu_oracle():
    """BASELINE config 4 at small size: two AdamW steps of CARN (STFT / mask / iSTFT autograd nodes of this
    package around stock real convolutions) against the same steps on the CPU oracle."""
    from speech_enhancement_b200 import training
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        ref = callers.CARN(ops=OPS).train()
        sd = seeded_state_dict(ref, seed=8)
        ref.load_state_dict(sd)
        ours = callers.CARN().train()
        ours.load_state_dict(sd)
        ours.to(DEV)
        mix, clean = synth_noisy_speech(2, 8000, seed=30), synth_noisy_speech(2, 8000, seed=31)
        o_ref, o_ours = training.make_optimizer(ref), training.make_optimizer(ours)
        f_ref, f_ours = training.FlatGradients(ref), training.FlatGradients(ours)
        for step in range(2):
            l_ref = training.train_step(ref, o_ref, f_ref, mix, clean)
            l_ours = training.train_step(ours, o_ours, f_ours, mix.to(DEV), clean.to(DEV))
            assert abs(l_ref.item() - l_ours.item()) < 2e-3 * max(1.0, abs(l_ref.item()))
            assert rel_l2(f_ours.buffer.cpu(), f_ref.buffer) < 2e-3
        for (k, a), (_, b) in zip(ours.state_dict().items(), ref.state_dict().items()):
            if a.dtype.is_floating_point:
                assert rel_l2(a.cpu(), b) < 2e-3, k
    finally:
        torch.backends.cudnn.allow_tf32 = tf32


@pytest.mark.parametrize('i', range(len(CONV_CASES)))
def test_cconv_backward_vs_oracle(i):
    """autograd of ComplexConv2d / ComplexConvTranspose2d (dgrad on the engine with swapped weight axes, wgrad
    on the pixel-reduction GEMM se_cconv_wgrad) against torch autograd of the oracle's four real convolutions."""
    name, transposed, cin, cout, k, kw, H, W, B = CONV_CASES[i]
    kw = dict(kw)
    kw.setdefault('bias', i % 2 == 0)
    mk = (lambda ns: ns.ComplexConvTranspose2d if transposed else ns.ComplexConv2d)
    ours, ref = mk(se)(cin, cout, k, **kw), mk(OPS)(cin, cout, k, **kw)
    sd = {key: rand(v.shape, 6000 + 10 * i + j, 0.2) for j, (key, v) in enumerate(sorted(ref.state_dict().items()))}
    ours.load_state_dict(sd)
    ref.load_state_dict(sd)
    ours.to(DEV)
    x = rand((B, cin, H, W), 7000 + i)
    xo, xr = x.clone().to(DEV).requires_grad_(True), x.clone().requires_grad_(True)
    yo, yr = ours(xo), ref(xr)
    w = rand(tuple(yr.shape), 7100 + i)
    (yo * w.to(DEV)).sum().backward()
    (yr * w).sum().backward()
    assert rel_l2(yo.detach().cpu(), yr.detach()) < 2e-4
    assert rel_l2(xo.grad.cpu(), xr.grad) < 2e-4, name
    for (n1, p1), (n2, p2) in zip(sorted(ours.named_parameters()), sorted(ref.named_parameters())):
        assert n1 == n2 and p1.grad is not None
        assert rel_l2(p1.grad.cpu(), p2.grad) < 2e-4, (name, n1)


def test_dccrn_train_mode_step_vs_oracle():
    """a full DCCRN-CL training-mode forward / backward through the module path (complex BN in batch-statistics
    mode and the LSTM on stock PyTorch, every complex conv / STFT / mask / iSTFT through this package's nodes)."""
    from speech_enhancement_b200 import training
    args = ('dccrn-CL', 400, 100, 512)
    ref = callers.DCCRN(*args, ops=OPS).train()
    sd = seeded_state_dict(ref, seed=13)
    ref.load_state_dict(sd)
    ours = callers.DCCRN(*args).train()
    ours.load_state_dict(sd)
    ours.to(DEV)
    mix, clean = synth_noisy_speech(2, 6000, seed=50), synth_noisy_speech(2, 6000, seed=51)
    f_ref, f_ours = training.FlatGradients(ref), training.FlatGradients(ours)
    with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        losses = []
        for model, m, c in ((ref, mix, clean), (ours, mix.to(DEV), clean.to(DEV))):
            _, wav = model(m)
            loss = training.si_snr_loss(training.pad_or_truncate(wav, c), c)
            loss.backward()
            losses.append(loss.item())
    assert abs(losses[0] - losses[1]) < 1e-3 * max(1.0, abs(losses[0]))
    assert rel_l2(f_ours.buffer.cpu(), f_ref.buffer) < 5e-3


@pytest.mark.parametrize('cls,args,B,L', [(callers.DCCRN, ('dccrn-CL', 400, 100, 512), 2, 8000),
                                          (callers.DCUNet, ('dcunet20', 512, 128, 512), 1, 16384)])
def test_cuda_graph_replay_matches_eager(cls, args, B, L):
    """the chained forward is capturable (no host sync, host-side tensor maps, cached operands): a CUDA-graph
    replay returns exactly what the eager call returns, also after the input changes."""
    from speech_enhancement_b200.graphs import GraphedInference
    model = cls(*args).eval()
    model.load_state_dict(seeded_state_dict(model, seed=14))
    model.to(DEV)
    g = GraphedInference(model)
    for seed in (60, 61):
        x = synth_noisy_speech(B, L, seed=seed).to(DEV)
        with torch.no_grad():
            spec_e, wav_e = model(x)
        spec_g, wav_g = g(x)
        assert torch.equal(wav_g, wav_e) and torch.equal(spec_g, spec_e)
    assert len(g._graphs) == 1


@pytest.mark.parametrize('B,H,W,CO,kh,kw,sh,sw,ph,pw,oph,opw', [
    (2, 16, 37, 2, 5, 2, 2, 1, 2, 0, 1, 0),      # DCCRN last decoder layer (_2008_00264_dccrn.py:104-111)
    (1, 33, 70, 2, 7, 1, 1, 1, 3, 0, 0, 0),      # DCUNet last decoder layer (_1903_03107_dcunet.py:79-81)
    (3, 9, 11, 4, 3, 3, 2, 2, 1, 1, 1, 1),
    (1, 70, 130, 2, 5, 3, 2, 2, 2, 1, 0, 0),     # several tiles in both directions
    (2, 5, 3, 4, 1, 2, 1, 1, 0, 0, 0, 0),
])
def test_col2im_gathers_every_tap_product(B, H, W, CO, kh, kw, sh, sw, ph, pw, oph, opw):
    """se_cconvT_col2im: y[b, co, hi*sh - ph + dh, wi*sw - pw + dw] += P[b, hi, wi, (dh*kw + dw)*CO + co], + bias, LeakyReLU"""
    from speech_enhancement_b200 import _abi
    Ho, Wo = (H - 1) * sh - 2 * ph + kh + oph, (W - 1) * sw - 2 * pw + kw + opw
    ncol = kh * kw * CO
    Np = (ncol + 3) // 4 * 4
    P = rand((B, H, W, Np), 77, 1.0).to(DEV)
    bias = rand((CO,), 78, 1.0).to(DEV)
    slope = 0.25
    y = torch.full((B, CO, Ho, Wo), float('nan'), device=DEV)
    _abi.check(_abi.lib().se_cconvT_col2im(_abi.ptr(P), B, H, W, Np, CO, kh, kw, sh, sw, ph, pw, Ho, Wo, _abi.ptr(bias),
                                           slope, _abi.ptr(y), _abi.stream_ptr()))
    want = torch.zeros((B, CO, Ho + 2 * ph + kh, Wo + 2 * pw + kw), device=DEV, dtype=torch.float64)
    for dh in range(kh):
        for dw in range(kw):
            t = (dh * kw + dw) * CO
            want[:, :, dh:dh + (H - 1) * sh + 1:sh, dw:dw + (W - 1) * sw + 1:sw] += P[..., t:t + CO].permute(0, 3, 1, 2).double()
    want = want[:, :, ph:ph + Ho, pw:pw + Wo] + bias.double().view(1, CO, 1, 1)
    want = torch.where(want > 0, want, want * slope).float()
    assert torch.isfinite(y).all()
    assert (y - want).abs().max().item() <= 2e-6 * max(1.0, want.abs().max().item())
