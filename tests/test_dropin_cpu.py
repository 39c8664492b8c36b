"""dropin.install(): the reference's UNMODIFIED model files bind to this package's classes
(SURVEY.md 8b "Module names").  Needs /root/reference, so it runs in the authoring container only
and is skipped on the GPU box; runs in a subprocess because it edits sys.modules."""
import os
import subprocess
import sys

import pytest

REF = '/root/reference'
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CODE = r'''
import sys, warnings
warnings.filterwarnings('ignore')
sys.path.insert(0, %r)
import numpy as np
import speech_enhancement_b200 as se
from speech_enhancement_b200 import dropin
names = dropin.install(%r)
assert set(names) == {'conv_stft', 'models.conv_stft', 'modules.complex_nn', 'models.modules.complex_nn'}
import _2008_00264_dccrn as dccrn, _1903_03107_dcunet as dcunet, _2206_07293_frcrn as frcrn
import _2104_05267_carn as carn, _1809_01405_crn as crn
g = np.load(%r)
for name, model in (('dccrn_cl', dccrn.DCCRN('dccrn-CL', 400, 100, 512)),
                    ('dcunet20', dcunet.DCUNet('dcunet20', 512, 128, 512)),
                    ('frcrn', frcrn.FRCRN(320, 160, 640))):
    assert type(model.stft) is se.ConvSTFT and type(model.istft) is se.ConviSTFT, name
    convs = [m for m in model.modules() if isinstance(m, (se.ComplexConv2d, se.ComplexConvTranspose2d))]
    assert len(convs) >= 12, (name, len(convs))
    assert not any(type(m).__module__.startswith('modules.') and 'ComplexConv' in type(m).__name__
                   for m in model.modules())
    assert sorted(model.state_dict().keys()) == list(g[name + '_keys']), name
assert type(carn.CARN(320, 160, 512).stft) is se.ConvSTFT
assert type(crn.CRN(320, 160, 320).stft) is se.ConvSTFT
import models                                   # the package-relative names resolve to the same classes
assert models.ConvSTFT is se.ConvSTFT and models.ComplexConv2d is se.ComplexConv2d
dropin.uninstall()
assert 'conv_stft' not in sys.modules
print('dropin ok')
'''


@pytest.mark.skipif(not os.path.isdir(REF), reason='reference checkout not present on this machine')
def test_reference_models_bind_to_this_package():
    code = CODE % (ROOT, REF, os.path.join(ROOT, 'tests', 'golden', 'models.npz'))
    res = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and 'dropin ok' in res.stdout, res.stdout[-2000:] + res.stderr[-4000:]


def test_accelerate_recognises_reference_instances_and_refuses_others():
    """dropin.accelerate() on instances of the staged reference classes (oracle/_ref): structure check on the CPU,
    and the loud fallback to the reference's own forward for a CPU input"""
    import warnings
    import torch
    from oracle import ref_loader
    from speech_enhancement_b200 import dropin
    if not ref_loader.available():
        pytest.skip(ref_loader.why_unavailable())
    orig, drop = ref_loader.load_original(), ref_loader.load_dropin()
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        for cls, args in (('DCCRN', ('dccrn-CL', 400, 100, 512)), ('DCUNet', ('dcunet10', 512, 128, 512)),
                          ('FRCRN', (320, 160, 640))):
            m = getattr(drop, cls)(*args).eval()
            assert dropin.accelerate(m) is m and m._se_chained is not None
            assert dropin.accelerate(m) is m                            # idempotent
            dropin.decelerate(m)
            assert 'forward' not in m.__dict__
        with pytest.raises(RuntimeError, match='no chained forward'):
            dropin.accelerate(drop.CARN(320, 160, 512))
        with pytest.raises(RuntimeError, match='not this package'):     # the reference bound to its OWN modules
            dropin.accelerate(orig.DCCRN('dccrn-CL', 400, 100, 512))
        with pytest.raises(RuntimeError, match='is_complex=False'):
            dropin.accelerate(drop.DCCRN('dccrn-CL', 400, 100, 512, is_complex=False))
    # the original binding still runs on the CPU next to the drop-in binding (one process, two bindings)
    with torch.no_grad(), warnings.catch_warnings():
        warnings.simplefilter('ignore')
        wav = orig.DCCRN('dccrn-CL', 400, 100, 512).eval()(torch.zeros(1, 4000))[1]
    assert wav.shape == (1, 3988)


def test_phase_folded_transposed_conv_prototype_matches_torch():
    """scripts/phase_fold_prototype.py (the planned N-folding of the stride-2 decoder layers) reproduces
    F.conv_transpose2d for the DCCRN and DCUNet decoder geometries"""
    import importlib.util
    import os
    import torch
    import torch.nn.functional as F
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'scripts', 'phase_fold_prototype.py')
    spec = importlib.util.spec_from_file_location('phase_fold_prototype', path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    g = torch.Generator().manual_seed(5)
    for k, st, pd, op, groups in [((5, 2), (2, 1), (2, 0), (1, 0), 6), ((7, 5), (2, 2), (3, 2), (0, 0), 12),
                                  ((5, 3), (2, 1), (2, 1), (0, 0), 9)]:
        x = torch.randn(2, 6, 5, 7, dtype=torch.float64, generator=g)
        w = torch.randn(6, 4, *k, dtype=torch.float64, generator=g)
        want = F.conv_transpose2d(x, w, stride=st, padding=pd, output_padding=op)
        got, n = mod.conv_transpose_folded(x, w, st, pd, op)
        assert n == groups and got.shape == want.shape
        assert (got - want).abs().max().item() < 1e-12
