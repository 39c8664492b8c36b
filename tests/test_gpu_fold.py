"""Phase-folded transposed convolution (the sh*sw output phases of a ComplexConvTranspose2d as column groups of one
stride-1 convolution, chain.FusedBlock.fold) against the per-tap form of the same engine and the fp64 CPU oracle."""
import pytest
import torch

import oracle
import speech_enhancement_b200 as se
from speech_enhancement_b200 import chain
from conftest import rand, rel_l2

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'

#        cin2, cout2, kernel, stride, padding, output_padding, B, H,  W,  two sources, bias
CASES = [(128, 32, (5, 2), (2, 1), (2, 0), (1, 0), 2, 64, 70, True, False),     # DCCRN dec4
         (256, 64, (5, 2), (2, 1), (2, 0), (1, 0), 3, 32, 41, True, False),     # DCCRN dec3
         (128, 64, (7, 5), (2, 2), (3, 2), (0, 0), 2, 33, 29, True, False),     # DCUNet dec7
         (64, 24, (5, 3), (2, 2), (2, 1), (0, 0), 2, 9, 17, False, True),       # 24 channels: padded per-phase group
         (64, 48, (3, 3), (2, 2), (0, 0), (1, 1), 1, 7, 11, False, True),       # no padding: offsets start below zero
         (32, 90, (5, 3), (2, 1), (2, 1), (0, 0), 2, 6, 13, False, False),      # 90 channels (DCUNet-large pads to 192)
         (64, 16, (4, 4), (2, 2), (1, 1), (0, 0), 2, 8, 8, False, True),        # even kernel
         (64, 32, (2, 3), (3, 1), (0, 1), (2, 0), 2, 5, 9, False, False),       # kernel smaller than the stride
         (64, 128, (5, 3), (2, 2), (2, 1), (0, 0), 2, 9, 12, False, True),      # 128 channels: only ONE axis fits 256 columns
         (128, 64, (7, 5), (2, 2), (3, 2), (1, 1), 1, 12, 10, True, False)]     # output_padding on both axes


@pytest.mark.parametrize('cin2,cout2,k,s,p,op,B,H,W,two,bias', CASES)
def test_folded_block_matches_per_tap_form_and_oracle(cin2, cout2, k, s, p, op, B, H, W, two, bias, monkeypatch):
    m = se.ComplexConvTranspose2d(cin2, cout2, k, stride=s, padding=p, output_padding=op, bias=bias)
    sd = {kk: rand(v.shape, 60 + j, 0.1) for j, (kk, v) in enumerate(sorted(m.state_dict().items()))}
    m.load_state_dict(sd)
    m.to(DEV)
    srcs_c = [cin2 // 2, cin2 // 2] if two else [cin2]
    xs = [rand((B, c, H, W), 70 + i).to(DEV) for i, c in enumerate(srcs_c)]
    full = se.complex_concat(xs, dim=1) if two else xs[0]
    rc, ic = m.real_conv, m.imag_conv
    want = oracle.cconvT2d_ref(full.cpu().double(), rc.weight.cpu().double(), ic.weight.cpu().double(),
                               rc.bias.cpu().double() if bias else None, ic.bias.cpu().double() if bias else None,
                               s, p, op, (1, 1))
    want = torch.nn.functional.leaky_relu(want, 0.2)
    outs = {}
    for mode in ('0', '1'):
        monkeypatch.setattr(chain, 'FOLD_MODE', mode)
        blk = chain.FusedBlock(m, None, torch.nn.LeakyReLU(0.2), src_channels=srcs_c, concat='complex')
        one_axis_cols = min(v for v in (s[0], s[1], s[0] * s[1]) if v > 1) * ((cout2 + 31) // 32 * 32)
        assert (blk.fold is not None) == (mode == '1' and one_axis_cols <= 256)
        with torch.no_grad():
            act, _ = blk([chain.cl2_from_nchw(x) for x in xs], out_cl2=True)
            # poison check for the pad channels: consumers may read them (they must be exact zeros)
            assert (act.data[..., cout2:] == 0).all()
            outs[mode] = chain.cl2_to_nchw(act).cpu()
    assert outs['1'].shape == want.shape
    assert rel_l2(outs['0'], want) < 2e-4
    assert rel_l2(outs['1'], want) < 2e-4
    assert rel_l2(outs['1'], outs['0']) < 2e-5


#            kind  cin2 cout2 kernel  stride  padding  B  H   W   two    bias   pad_left
PIXEL_CASES = [('C', 64, 64, (7, 1), (1, 1), (3, 0), 2, 37, 29, False, False, 0),     # DCUNet enc0: pairs along H
               ('C', 64, 64, (1, 7), (1, 1), (0, 3), 2, 19, 41, False, True, 0),      # DCUNet enc1: pairs along W
               ('T', 128, 64, (1, 7), (1, 1), (0, 3), 2, 21, 33, True, False, 0),     # DCUNet dec8: stride-1 transposed
               ('C', 32, 64, (5, 2), (2, 1), (2, 0), 3, 40, 27, False, False, 1),     # DCCRN enc1: stride 2 + causal pad
               ('C', 64, 24, (3, 3), (1, 1), (1, 1), 2, 9, 10, False, True, 0),       # padded column groups, even / odd extents
               ('T', 64, 32, (3, 5), (1, 1), (1, 2), 1, 8, 7, False, True, 0)]


@pytest.mark.parametrize('kind,cin2,cout2,k,s,p,B,H,W,two,bias,pad_left', PIXEL_CASES)
def test_pixel_pair_folded_block_matches_per_tap_form_and_oracle(kind, cin2, cout2, k, s, p, B, H, W, two, bias, pad_left,
                                                                 monkeypatch):
    tr = kind == 'T'
    mk = se.ComplexConvTranspose2d if tr else se.ComplexConv2d
    m = mk(cin2, cout2, k, stride=s, padding=p, bias=bias)
    sd = {kk: rand(v.shape, 80 + j, 0.1) for j, (kk, v) in enumerate(sorted(m.state_dict().items()))}
    m.load_state_dict(sd)
    m.to(DEV)
    srcs_c = [cin2 // 2, cin2 // 2] if two else [cin2]
    xs = [rand((B, c, H, W), 90 + i).to(DEV) for i, c in enumerate(srcs_c)]
    full = (se.complex_concat(xs, dim=1) if two else xs[0]).cpu().double()
    if pad_left:
        full = torch.nn.functional.pad(full, (pad_left, 0, 0, 0))
    rc, ic = m.real_conv, m.imag_conv
    br, bi = (rc.bias.cpu().double(), ic.bias.cpu().double()) if bias else (None, None)
    if tr:
        want = oracle.cconvT2d_ref(full, rc.weight.cpu().double(), ic.weight.cpu().double(), br, bi, s, p, (0, 0), (1, 1))
    else:
        want = oracle.cconv2d_ref(full, rc.weight.cpu().double(), ic.weight.cpu().double(), br, bi, s, p, (1, 1))
    want = torch.nn.functional.leaky_relu(want, 0.01)
    outs = {}
    for mode in ('0', '1'):
        monkeypatch.setattr(chain, 'FOLD_MODE', mode)
        blk = chain.FusedBlock(m, None, torch.nn.LeakyReLU(0.01), src_channels=srcs_c, concat='complex')
        assert (blk.fold is not None) == (mode == '1') and (blk.fold is None or blk.fold['kind'] == 'pixel')
        with torch.no_grad():
            act, _ = blk([chain.cl2_from_nchw(x) for x in xs], pad_left=pad_left, out_cl2=True)
            assert (act.data[..., cout2:] == 0).all()
            outs[mode] = chain.cl2_to_nchw(act).cpu()
    assert outs['1'].shape == want.shape
    assert rel_l2(outs['0'], want) < 2e-4
    assert rel_l2(outs['1'], want) < 2e-4
    assert rel_l2(outs['1'], outs['0']) < 2e-5


def test_auto_mode_folds_the_layers_the_cost_model_names(monkeypatch):
    monkeypatch.setattr(chain, 'FOLD_MODE', 'auto')
    mk = lambda cin, cout, k, s, p: se.ComplexConvTranspose2d(cin, cout, k, stride=s, padding=p)
    assert chain.FusedBlock(mk(128, 32, (5, 2), (2, 1), (2, 0))).fold is not None      # DCCRN dec4
    assert chain.FusedBlock(mk(256, 64, (5, 2), (2, 1), (2, 0))).fold is not None      # DCCRN dec3
    assert chain.FusedBlock(mk(512, 128, (5, 2), (2, 1), (2, 0))).fold is None         # N = 256 already
    f = chain.FusedBlock(mk(256, 64, (7, 5), (2, 2), (3, 2))).fold                      # DCUNet dec7: all 4 phases = 256
    assert f is not None and (f['fh'], f['fw']) == (2, 1) and len(f['launches']) == 2   # columns lose; H only, 2 launches
    assert chain.FusedBlock(mk(256, 128, (5, 2), (2, 1), (0, 0))).fold is None         # FRCRN decoder
    # pixel-pair folding measured no gain on the short-K layers it applies to: built only under SE_CONV_FOLD=1
    assert chain.FusedBlock(mk(128, 64, (1, 7), (1, 1), (0, 3))).fold is None           # DCUNet dec8: stride 1, no phases
    assert chain.FusedBlock(se.ComplexConv2d(64, 64, (7, 1), padding=(3, 0))).fold is None
    monkeypatch.setattr(chain, 'FOLD_MODE', '1')
    f = chain.FusedBlock(se.ComplexConv2d(64, 64, (7, 1), padding=(3, 0))).fold
    assert f is not None and f['kind'] == 'pixel' and f['axis'] == 0
    monkeypatch.setattr(chain, 'FOLD_MODE', 'auto')
    assert chain.FusedBlock(se.ComplexConv2d(128, 256, (5, 2), stride=(2, 1), padding=(2, 0))).fold is None
    assert chain.FusedBlock(mk(128, 32, (5, 2), (2, 1), (2, 0))).fold['kind'] == 'phase' 


#             kind cin2 cout2 kernel stride padding  B  H   W   two   bias  pad_left
ROW_CASES = [('C', 64, 64, (1, 7), (1, 1), (0, 3), 2, 5, 501, False, False, 0),      # DCUNet enc1
             ('T', 128, 64, (1, 7), (1, 1), (0, 3), 2, 3, 300, True, False, 0),      # DCUNet dec8 (two sources)
             ('C', 64, 128, (1, 2), (1, 1), (0, 0), 1, 4, 131, False, True, 1),      # causal pad: shifts 0 and 1
             ('C', 128, 256, (1, 5), (1, 1), (0, 0), 2, 2, 77, False, False, 0),     # N = 256 (split operand form), W < 128
             ('T', 64, 32, (1, 3), (1, 1), (0, 1), 3, 7, 129, False, True, 0),
             ('C', 64, 48, (1, 9), (2, 1), (0, 4), 2, 9, 260, False, False, 0)]      # row stride 2, nine column shifts (> 8)


@pytest.mark.parametrize('kind,cin2,cout2,k,s,p,B,H,W,two,bias,pad_left', ROW_CASES)
def test_single_row_kernels_share_one_halo_box(kind, cin2, cout2, k, s, p, B, H, W, two, bias, pad_left, built_lib):
    """1 x k kernels: the k kernel columns are column-shifted views of ONE staged A box (UMMA descriptors with a matrix
    base offset) instead of k separate loads; against the fp64 oracle and the exact-fp32 SIMT engine"""
    tr = kind == 'T'
    mk = se.ComplexConvTranspose2d if tr else se.ComplexConv2d
    m = mk(cin2, cout2, k, stride=s, padding=p, bias=bias)
    sd = {kk: rand(v.shape, 180 + j, 0.1) for j, (kk, v) in enumerate(sorted(m.state_dict().items()))}
    m.load_state_dict(sd)
    m.to(DEV)
    srcs_c = [cin2 // 2, cin2 // 2] if two else [cin2]
    xs = [rand((B, c, H, W), 190 + i).to(DEV) for i, c in enumerate(srcs_c)]
    full = se.complex_concat(xs, dim=1) if two else xs[0]
    if pad_left:
        full = torch.nn.functional.pad(full, (pad_left, 0, 0, 0))
    rc, ic = m.real_conv, m.imag_conv
    br, bi = (rc.bias.cpu().double(), ic.bias.cpu().double()) if bias else (None, None)
    fd = full.cpu().double()
    if tr:
        want = oracle.cconvT2d_ref(fd, rc.weight.cpu().double(), ic.weight.cpu().double(), br, bi, s, p, (0, 0), (1, 1))
    else:
        want = oracle.cconv2d_ref(fd, rc.weight.cpu().double(), ic.weight.cpu().double(), br, bi, s, p, (1, 1))
    with torch.no_grad():
        blk = chain.FusedBlock(m, None, None, src_channels=srcs_c, concat='complex')
        act, _ = blk([chain.cl2_from_nchw(x) for x in xs], pad_left=pad_left, out_cl2=True)
        got = chain.cl2_to_nchw(act).cpu()
        m.engine = 2                                           # module path on the tensor engine (NCHW in / out)
        got_mod = m(full.contiguous()).cpu()
        m.engine = 1                                           # exact fp32 SIMT
        got_simt = m(full.contiguous()).cpu()
    assert got.shape == want.shape
    assert rel_l2(got_simt, want) < 1e-5
    assert rel_l2(got, want) < 2e-4 and rel_l2(got_mod, want) < 2e-4
    assert rel_l2(got, got_simt) < 5e-5
