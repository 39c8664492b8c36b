"""Generate tests/golden/*.npz from the UNMODIFIED reference at /root/reference (CPU, fp32).

Run in the authoring container only (`python tests/golden/make_golden.py`); the GPU box has no
/root/reference, so the vectors are committed.  Everything is seeded: inputs come from
torch.Generator seeds recorded in each file, model weights from
speech_enhancement_b200.synth.seeded_state_dict, so fixtures hold only the reference OUTPUTS
(plus small inputs where regenerating them would need the reference).
"""
import importlib
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = '/root/reference'
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(REF, 'models'))
sys.path.insert(0, REF)
warnings.filterwarnings('ignore')

from speech_enhancement_b200.synth import seeded_state_dict, synth_noisy_speech  # noqa: E402

ref_stft = importlib.import_module('conv_stft')
ref_cnn = importlib.import_module('modules.complex_nn')

# (win, hop, fft, L) -- every triple the reference's smoke blocks use
STFT_CASES = [(400, 100, 512, 4000), (512, 128, 512, 4096), (1024, 256, 1024, 6000), (320, 160, 640, 3200),
              (320, 160, 512, 3200), (512, 256, 512, 4000), (320, 160, 320, 3200), (256, 64, 256, 448)]

# (name, transposed, cin, cout, kernel, kwargs, H, W, B) -- every distinct geometry of SURVEY.md 8a
CONV_CASES = [
    ('dccrn_enc', False, 8, 16, (5, 2), dict(stride=(2, 1), padding=(2, 0)), 16, 12, 2),
    ('dccrn_enc0', False, 2, 32, (5, 2), dict(stride=(2, 1), padding=(2, 0)), 32, 9, 1),
    ('dccrn_dec', True, 16, 8, (5, 2), dict(stride=(2, 1), padding=(2, 0), output_padding=(1, 0)), 8, 11, 2),
    ('dccrn_dec_last', True, 16, 2, (5, 2), dict(stride=(2, 1), padding=(2, 0), output_padding=(1, 0)), 8, 11, 1),
    ('frcrn_enc', False, 8, 8, (5, 2), dict(stride=(2, 1), padding=(0, 0)), 17, 10, 2),
    ('frcrn_dec', True, 8, 8, (5, 2), dict(stride=(2, 1), padding=(0, 0)), 7, 9, 2),
    ('dcunet_first', False, 2, 16, 3, dict(padding=1), 17, 13, 1),
    ('dcunet_7x1', False, 8, 8, (7, 1), dict(stride=(1, 1), padding=(3, 0)), 17, 13, 1),
    ('dcunet_1x7', False, 8, 8, (1, 7), dict(stride=(1, 1), padding=(0, 3)), 17, 13, 1),
    ('dcunet_7x5_s22', False, 8, 16, (7, 5), dict(stride=(2, 2), padding=(3, 2)), 17, 13, 2),
    ('dcunet_7x5_s21', False, 8, 16, (7, 5), dict(stride=(2, 1), padding=(3, 2)), 17, 13, 1),
    ('dcunet_5x3_s22', False, 8, 12, (5, 3), dict(stride=(2, 2), padding=(2, 1)), 9, 14, 2),
    ('dcunet_5x3_s21', False, 8, 12, (5, 3), dict(stride=(2, 1), padding=(2, 1)), 9, 14, 1),
    ('dcunet_T5x3_s21', True, 12, 8, (5, 3), dict(stride=(2, 1), padding=(2, 1)), 5, 7, 2),
    ('dcunet_T5x3_s22', True, 12, 8, (5, 3), dict(stride=(2, 2), padding=(2, 1)), 5, 7, 1),
    ('dcunet_T7x5_s21', True, 8, 8, (7, 5), dict(stride=(2, 1), padding=(3, 2)), 6, 7, 1),
    ('dcunet_T7x5_s22', True, 8, 4, (7, 5), dict(stride=(2, 2), padding=(3, 2)), 6, 7, 2),
    ('dcunet_T1x7', True, 8, 4, (1, 7), dict(stride=(1, 1), padding=(0, 3)), 9, 11, 1),
    ('dcunet_T7x1_last', True, 8, 2, (7, 1), dict(stride=(1, 1), padding=(3, 0)), 9, 11, 1),
    ('ccbam_7x7', False, 4, 2, 7, dict(padding=3), 9, 12, 2),
    ('bias_conv', False, 4, 6, (3, 2), dict(stride=(1, 1), padding=(1, 0), bias=True), 7, 6, 2),
    ('bias_convT', True, 6, 4, (3, 2), dict(stride=(2, 1), padding=(1, 0), output_padding=(1, 0), bias=True), 4, 6, 2),
]


def rand(shape, seed, scale=1.0):
    return torch.randn(shape, generator=torch.Generator().manual_seed(seed)) * scale


def save(name, **arrays):
    path = os.path.join(HERE, name + '.npz')
    np.savez_compressed(path, **{k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v))
                                 for k, v in arrays.items()})
    print(f'{name}.npz: {os.path.getsize(path) / 1024:.1f} KiB')


@torch.no_grad()
def gen_stft():
    out = {}
    for i, (win, hop, fft, L) in enumerate(STFT_CASES):
        x = rand((2, L), 100 + i, 0.1)
        st, ist = ref_stft.ConvSTFT(win, hop, fft), ref_stft.ConviSTFT(win, hop, fft)
        spec = st(x)
        out[f'{i}_cfg'] = np.array([win, hop, fft, L, 100 + i])
        out[f'{i}_spec'] = spec
        out[f'{i}_wav'] = ist(spec)
        out[f'{i}_wav_len'] = ist(spec, output_length=L - 7)
        if i in (0, 6):            # mag/phase branch: DCCRN triple and the CRN triple (its only user)
            mags, phase = ref_stft.ConvSTFT(win, hop, fft, return_mag_phase=True)(x)
            out[f'{i}_mags'], out[f'{i}_phase'] = mags, phase
            out[f'{i}_wav_polar'] = ist(mags, phase)
        # basis fingerprints: strided samples + fp64 sums of the reference buffers
        for nm, buf in (('fwd', st.weight), ('inv', ist.weight), ('window', ist.window)):
            flat = buf.reshape(-1).double()
            out[f'{i}_{nm}_samples'] = flat[::97].float()
            out[f'{i}_{nm}_sums'] = np.array([flat.sum().item(), flat.abs().sum().item(), (flat * flat).sum().item()])
    # non-centred and 1-D / 3-D input forms
    x = rand((3000,), 77, 0.1)
    st = ref_stft.ConvSTFT(400, 100, 512, center=False)
    ist = ref_stft.ConviSTFT(400, 100, 512, center=False)
    out['nc_spec'] = st(x)
    out['nc_wav'] = ist(st(x))
    save('stft', **out)


@torch.no_grad()
def gen_conv():
    out = {}
    for i, (name, transposed, cin, cout, k, kw, H, W, B) in enumerate(CONV_CASES):
        cls = ref_cnn.ComplexConvTranspose2d if transposed else ref_cnn.ComplexConv2d
        kw = dict(kw)
        kw.setdefault('bias', False)
        m = cls(cin, cout, k, **kw)
        sd = {key: rand(v.shape, 1000 + 10 * i + j, 0.2) for j, (key, v) in enumerate(sorted(m.state_dict().items()))}
        m.load_state_dict(sd)
        x = rand((B, cin, H, W), 5000 + i)
        out[f'{name}_y'] = m(x)
    save('cconv', **out)


@torch.no_grad()
def gen_masks():
    dccrn = importlib.import_module('_2008_00264_dccrn')
    dcunet = importlib.import_module('_1903_03107_dcunet')
    out = {}
    B, F, T = 2, 9, 21
    nr, ni, mr, mi = (rand((B, F, T), 900 + j, s) for j, s in enumerate((0.5, 0.5, 1.5, 1.5)))
    for mode in 'RCE':
        self = type('S', (), {'masking': mode, '_return_mag_phase': dccrn.DCCRN._return_mag_phase})()
        re, im = dccrn.DCCRN._mask_processing(self, nr, ni, mr, mi)
        out[f'dccrn_{mode}'] = torch.cat([re, im], 1)
    self = type('S', (), {'_return_mag_phase': dcunet.DCUNet._return_mag_phase})()
    for method in ('unbounded', 'bounded_sigmoid', 'bounded_tanh'):
        y = dcunet.DCUNet._mask_processing(self, torch.stack([mr, mi], 1), torch.stack([nr, ni], 1), method)
        out[f'dcunet_{method}'] = y.reshape(B, 2 * F, T)
    # CARN (_2104_05267_carn.py:161-169) and FRCRN (_2206_07293_frcrn.py:141-152) masks are inline
    # in forward(): run the reference models and capture the mask inputs with forward hooks.
    for name, mod, cls, args, tail in (('carn', '_2104_05267_carn', 'CARN', (320, 160, 512), 'linear'),
                                       ('frcrn', '_2206_07293_frcrn', 'FRCRN', (320, 160, 640), 'final_conv')):
        model = getattr(importlib.import_module(mod), cls)(*args)
        model.load_state_dict(seeded_state_dict(model, seed=3))
        model.eval()
        cap = {}
        model.stft.register_forward_hook(lambda m, i, o: cap.__setitem__('spec', o))
        getattr(model, tail).register_forward_hook(lambda m, i, o: cap.__setitem__('net', o))
        est_spec, _ = model(synth_noisy_speech(2, 2400, seed=5))
        net = cap['net']
        if name == 'carn':                      # [B, T, fft+2] -> [B, 2, F, T]  (:156-158)
            b, t, _ = net.shape
            net = net.permute(0, 2, 1).reshape(b, 2, -1, t)
        out[f'{name}_noisy_spec'] = cap['spec']
        out[f'{name}_net'] = net
        out[f'{name}_est_spec'] = est_spec
    save('masks', **out)


@torch.no_grad()
def gen_models():
    out = {}
    specs = [('dccrn_cl', '_2008_00264_dccrn', 'DCCRN', ('dccrn-CL', 400, 100, 512), 4000),
             ('dccrn_c', '_2008_00264_dccrn', 'DCCRN', ('dccrn-C', 400, 100, 512), 3000),
             ('dcunet20', '_1903_03107_dcunet', 'DCUNet', ('dcunet20', 512, 128, 512), 8192),
             ('dcunet10', '_1903_03107_dcunet', 'DCUNet', ('dcunet10', 512, 128, 512), 4096),
             ('frcrn', '_2206_07293_frcrn', 'FRCRN', (320, 160, 640), 3200)]
    for name, mod, cls, args, L in specs:
        model = getattr(importlib.import_module(mod), cls)(*args)
        model.load_state_dict(seeded_state_dict(model, seed=7))
        model.eval()
        x = synth_noisy_speech(2, L, seed=11)
        spec, wav = model(x)
        out[f'{name}_L'] = np.array([L])
        out[f'{name}_wav'] = wav
        out[f'{name}_spec_sums'] = np.array([spec.double().sum().item(), spec.double().abs().sum().item()])
        out[f'{name}_spec_samples'] = spec.reshape(-1)[::53]
        out[f'{name}_keys'] = np.array(sorted(model.state_dict().keys()))
    save('models', **out)


if __name__ == '__main__':
    torch.set_num_threads(8)
    gen_stft()
    gen_conv()
    gen_masks()
    gen_models()
