"""Make the reference's UNMODIFIED model files bind to the B200 path.

The reference has no plugin registry: its model files reach the hot path through module
names -- the flat ones (`from conv_stft import ConvSTFT, ConviSTFT`, _2008_00264_dccrn.py:8;
`from modules.complex_nn import ...`, :9; `from .complex_nn import ...`, modules/ccbam.py:5),
which resolve because train.py:2 / test.py:2 put `models/` on sys.path, and the package ones
(`models/__init__.py:9-10`).  `install()` seeds all four names in `sys.modules` BEFORE the model
files are imported, so `DCCRN(...)`, `DCUNet(...)`, `FRCRN(...)`, `CARN(...)`, `CRN(...)` build
their STFT and complex-conv layers from this package ("Module names").

    import speech_enhancement_b200.dropin as dropin
    dropin.install('/path/to/Speech-Enhancement')      # adds <root>/models and <root> to sys.path
    from _2008_00264_dccrn import DCCRN                 # the reference file, untouched
    model = DCCRN('dccrn-CL', 400, 100, 512).cuda().eval()
"""
import sys
import types

_NAMES_STFT = ('conv_stft', 'models.conv_stft')
_NAMES_CNN = ('modules.complex_nn', 'models.modules.complex_nn')
_saved = {}


def _complex_nn_module():
    """One module object carrying every public symbol of the reference's complex_nn.py."""
    from .modules import complex_nn, complex_aux
    mod = types.ModuleType('complex_nn')
    mod.__doc__ = 'speech_enhancement_b200 stand-in for models/modules/complex_nn.py'
    for src in (complex_nn, complex_aux):
        for k in ('complex_concat', 'split_complex', 'merge_real_imag', 'ComplexConv', 'ComplexConv2d',
                  'ComplexConvTranspose2d', 'ComplexLinear', 'ComplexLSTM', 'ComplexBatchNorm2d',
                  'ComplexPReLU', 'ComplexReLU', 'ComplexLeakyReLU'):
            if hasattr(src, k):
                setattr(mod, k, getattr(src, k))
    return mod


def install(reference_root=None):
    """Route the reference's imports of the hot-path modules to this package.

    Must run before any `models/_*.py` is imported (already-imported model files keep the
    classes they bound).  Returns the list of sys.modules names that were seeded."""
    import os
    from . import conv_stft
    already = [m for m in sys.modules if m.startswith('_') and m[1:5].isdigit() or m.startswith('models._')]
    if already:
        raise RuntimeError(f'dropin.install(): reference model files already imported ({already[:3]}); '
                           'call install() first')
    cnn = _complex_nn_module()
    for name in _NAMES_STFT:
        _saved.setdefault(name, sys.modules.get(name))
        sys.modules[name] = conv_stft
    for name in _NAMES_CNN:
        _saved.setdefault(name, sys.modules.get(name))
        sys.modules[name] = cnn
    if reference_root is not None:
        for p in (os.path.join(reference_root, 'models'), reference_root):
            if p not in sys.path:
                sys.path.insert(0, p)
    return list(_NAMES_STFT + _NAMES_CNN)


def uninstall():
    """Undo install() (does not un-import model files that already bound to this package)."""
    for name, old in list(_saved.items()):
        if old is None:
            sys.modules.pop(name, None)
        else:
            sys.modules[name] = old
        del _saved[name]
