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
    dropin.accelerate(model)                             # optional: chained engine behind model.forward

`install()` alone gives the module-by-module drop-in: every ConvSTFT / ComplexConv2d / ... call of the reference's
forward is one call into libse_b200.so, with NCHW fp32 tensors between the layers exactly as the reference has
them.  `accelerate(model)` additionally routes eval-mode / no-grad CUDA forwards of a reference DCCRN / DCUNet /
FRCRN INSTANCE through the chained engine (chain.py): complex BN, activation, causal pad and skip concat are
folded into the conv kernels and activations stay in the engine's CL2 layout between layers.  Training, autograd
and CPU calls keep the reference's own forward.
"""
import sys
import types
import warnings

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


# ---------------------------------------------------------------------------------- accelerate()
class _Chained:
    """Adapter that lets the chained forwards written for callers.{DCCRN,DCUNet,FRCRN} run on an INSTANCE of the
    reference's own class: attribute lookups fall through to the wrapped model (same attribute names: `stft`,
    `istft`, `encoder.layers[i].conv/.norm/.act`, `decoder.layers[i].conv_transposed/...`, `lstm`, `fft_size`,
    `masking`, `first_conv`, `final_conv`: _2008_00264_dccrn.py:88-143, _1903_03107_dcunet.py:45-121,
    _2206_07293_frcrn.py:62-117), the folded-operand caches live here."""

    def __init__(self, model, host_cls):
        from . import callers
        d = self.__dict__
        d['_m'], d['_ops'], d['_fb'], d['_final_fb'], d['fused'] = model, callers.cuda_ops(), None, None, True
        d['_host'] = host_cls

    def __getattr__(self, name):
        host = self.__dict__['_host']
        if name in ('_fused_blocks', '_forward_fused', '_final_block'):
            return types.MethodType(getattr(host, name), self)
        return getattr(self.__dict__['_m'], name)


def _structure_problem(model, kind):
    """None when `model` has the layer structure the chained forward walks, else a description"""
    from . import conv_stft
    from .modules import complex_nn, complex_aux
    import torch.nn as nn
    if not isinstance(getattr(model, 'stft', None), conv_stft.ConvSTFT) or \
            not isinstance(getattr(model, 'istft', None), conv_stft.ConviSTFT):
        return 'stft / istft are not this package\'s modules (was dropin.install() called before the model was built?)'
    blocks = []
    if kind == 'DCUNet':
        blocks.append(('first_conv', getattr(model, 'first_conv', None), 'conv'))
    try:
        blocks += [(f'encoder.layers[{i}]', l, 'conv') for i, l in enumerate(model.encoder.layers)]
        blocks += [(f'decoder.layers[{i}]', l, 'conv_transposed') for i, l in enumerate(model.decoder.layers)]
    except AttributeError as e:
        return f'no encoder.layers / decoder.layers ({e})'
    for name, blk, attr in blocks:
        conv = getattr(blk, attr, None)
        want = complex_nn.ComplexConv2d if attr == 'conv' else complex_nn.ComplexConvTranspose2d
        if not isinstance(conv, want):
            return f'{name}.{attr} is {type(conv).__name__}, not this package\'s {want.__name__} (is_complex=False?)'
        norm = getattr(blk, 'norm', None)
        if not isinstance(norm, (complex_aux.ComplexBatchNorm2d, nn.Identity)):
            return f'{name}.norm is {type(norm).__name__}'
    if kind == 'FRCRN':
        if not isinstance(getattr(model, 'final_conv', None), nn.Conv2d):
            return 'final_conv is not an nn.Conv2d'
        if len(getattr(model.decoder, 'skip_connection_attention_layers', ())) != len(model.decoder.layers):
            return 'decoder.skip_connection_attention_layers does not pair with decoder.layers'
    if kind in ('DCCRN', 'FRCRN') and not hasattr(model, 'lstm'):
        return 'no lstm block'
    return None


def accelerate(model):
    """Put the chained engine behind `model.forward` of a reference DCCRN / DCUNet / FRCRN instance (built after
    `install()`; the class is the reference's own, unmodified).  Returns `model`.

    The chained forward serves CUDA fp32 (or int16 PCM) inputs in eval mode under torch.no_grad().  Everything else (training,
    autograd, CPU tensors) runs the reference's own forward, i.e. the module-by-module drop-in; the first such
    call per reason emits a RuntimeWarning so the slower path is never taken silently.  A model whose structure
    is not recognised raises here instead of being left half-accelerated."""
    import torch
    from . import callers
    kind = type(model).__name__
    hosts = {'DCCRN': callers.DCCRN, 'DCUNet': callers.DCUNet, 'FRCRN': callers.FRCRN}
    if kind not in hosts:
        raise RuntimeError(f'dropin.accelerate(): no chained forward for {kind}; supported: {sorted(hosts)} '
                           '(CARN / CRN use the accelerated STFT / mask / iSTFT modules through install() alone)')
    if getattr(model, '_se_chained', None) is not None:
        return model
    problem = _structure_problem(model, kind)
    if problem is not None:
        raise RuntimeError(f'dropin.accelerate({kind}): {problem}')
    if kind == 'FRCRN':
        for att in model.decoder.skip_connection_attention_layers:
            att.forward_cl2 = types.MethodType(callers.CCBAM.forward_cl2, att)
            att._attn_operands = types.MethodType(callers.CCBAM._attn_operands, att)
    view = _Chained(model, hosts[kind])
    reference_forward = model.forward
    warned = set()

    def forward(x):
        if model.training:
            why = 'the model is in training mode (batch-statistics complex BN is not folded)'
        elif torch.is_grad_enabled():
            why = 'autograd is enabled (call under torch.no_grad() for the chained engine)'
        elif not (torch.is_tensor(x) and x.is_cuda and x.dtype in (torch.float32, torch.int16)):
            why = 'the input is not a CUDA float32 (or int16 PCM) tensor'
        else:
            return view._forward_fused(x)
        if getattr(model.istft, 'pcm16_out', False):
            # the reference's own forward ends in clamp_(wav, -1, 1), which would wreck int16 samples
            raise RuntimeError(f'{kind}: istft.pcm16_out needs the chained forward, but {why}')
        if why not in warned:
            warned.add(why)
            warnings.warn(f'{kind}: running the reference\'s module-by-module forward because {why}',
                          RuntimeWarning, stacklevel=2)
        return reference_forward(x)

    model.__dict__['_se_chained'] = view
    model.__dict__['_se_reference_forward'] = reference_forward
    model.forward = forward
    return model


def decelerate(model):
    """undo accelerate(): `model.forward` is the reference's own method again"""
    if model.__dict__.pop('_se_chained', None) is not None:
        model.__dict__.pop('forward', None)
        model.__dict__.pop('_se_reference_forward', None)
    return model
