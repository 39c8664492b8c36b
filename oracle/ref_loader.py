"""Loads the staged, unmodified reference (oracle/_ref, see stage_ref.py) -- test infrastructure.

Two bindings of the SAME reference model files (models/_2008_00264_dccrn.py, _1903_03107_dcunet.py,
_2206_07293_frcrn.py, _2104_05267_carn.py, _1809_01405_crn.py):

  load_original()  the reference exactly as it ships: its own conv_stft.py and modules/complex_nn.py
                   (the fp32 PyTorch CPU path: the parity oracle and the timed `kind: "reference"` baseline)
  load_dropin()    the same model files after speech_enhancement_b200.dropin.install(): ConvSTFT / ConviSTFT /
                   ComplexConv2d / ComplexConvTranspose2d resolve to the CUDA package (the drop-in under test)

The reference's files import each other through FLAT module names (`from conv_stft import ...`,
`from modules.complex_nn import ...`, `from architectures import ...`; they resolve because train.py:2 /
test.py:2 put models/ on sys.path), so both bindings want the same sys.modules keys.  Each loader therefore
imports inside a scope that restores sys.path and removes the flat names from sys.modules afterwards; the class
objects keep their own module globals, so the two bindings coexist in one process.
"""
import contextlib
import importlib
import os
import sys
import types

from . import stage_ref

REF_DIR = stage_ref.DST
MODEL_FILES = {'DCCRN': '_2008_00264_dccrn', 'DCUNet': '_1903_03107_dcunet', 'FRCRN': '_2206_07293_frcrn',
               'CARN': '_2104_05267_carn', 'GCARN': '_2104_05267_carn', 'CRN': '_1809_01405_crn'}
_cache = {}


def available():
    """True when oracle/_ref holds the staged reference and every file matches its manifest digest"""
    return stage_ref.verify(REF_DIR)


def why_unavailable():
    return (f'{REF_DIR} is absent or modified: run `python oracle/stage_ref.py` (or __graft_entry__.build()) on a '
            'machine that has the reference checkout')


def _is_ref_module(name, mod):
    f = getattr(mod, '__file__', None) or ''
    return f.startswith(REF_DIR + os.sep)


@contextlib.contextmanager
def _flat_import_scope():
    names0, path0 = set(sys.modules), list(sys.path)
    sys.path.insert(0, REF_DIR)
    sys.path.insert(0, os.path.join(REF_DIR, 'models'))
    try:
        yield
    finally:
        sys.path[:] = path0
        for k in list(sys.modules):
            if k not in names0 and (_is_ref_module(k, sys.modules[k]) or k in ('conv_stft', 'modules.complex_nn',
                                                                            'models.conv_stft',
                                                                            'models.modules.complex_nn')):
                del sys.modules[k]


def _collect():
    ns = types.SimpleNamespace()
    for cls, file in MODEL_FILES.items():
        setattr(ns, cls, getattr(importlib.import_module(file), cls))
    ns.complex_nn = importlib.import_module('modules.complex_nn')
    ns.conv_stft = importlib.import_module('conv_stft')
    ns.CCBAM = importlib.import_module('modules').CCBAM
    ns.ConvSTFT, ns.ConviSTFT = ns.conv_stft.ConvSTFT, ns.conv_stft.ConviSTFT
    return ns


def _require():
    if not available():
        raise RuntimeError(why_unavailable())


def load_original():
    """namespace of the reference's own classes, bound to the reference's own hot-path modules"""
    _require()
    if 'original' not in _cache:
        clash = [k for k in ('conv_stft', 'modules', 'modules.complex_nn', 'architectures') if k in sys.modules]
        if clash:
            raise RuntimeError(f'ref_loader.load_original(): {clash} already in sys.modules (dropin installed?)')
        with _flat_import_scope():
            ns = _collect()
            ns.losses = importlib.import_module('losses')
        _cache['original'] = ns
    return _cache['original']


def load_dropin():
    """namespace of the reference's own model classes with the hot-path modules routed to the CUDA package"""
    _require()
    if 'dropin' not in _cache:
        from speech_enhancement_b200 import dropin
        with _flat_import_scope():
            dropin.install(REF_DIR)
            try:
                ns = _collect()
            finally:
                dropin.uninstall()
        _cache['dropin'] = ns
    return _cache['dropin']
