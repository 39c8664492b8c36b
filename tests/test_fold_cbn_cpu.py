"""Host-side algebra of two GPU paths, checked on the CPU against stock PyTorch:
  * chain.fold_transposed_weight: the phase-folded operand of a transposed convolution reproduces F.conv_transpose2d
    when evaluated the way the engine evaluates it (stride-1 convolution over the input grid, phases as column groups);
  * complex_aux._cbn_closed_form: statistics -> whitening coefficients of the fused train-mode complex BN, forward and
    backward, against the module's plain-PyTorch path (complex_nn.py:244-322)."""
import pytest
import torch
import torch.nn.functional as F

from speech_enhancement_b200 import chain
from speech_enhancement_b200.modules.complex_aux import ComplexBatchNorm2d, _cbn_closed_form


@pytest.mark.parametrize('k,s,p,op', [((5, 2), (2, 1), (2, 0), (1, 0)), ((7, 5), (2, 2), (3, 2), (0, 0)),
                                      ((3, 3), (2, 2), (0, 0), (1, 1)), ((4, 4), (2, 2), (1, 1), (0, 0)),
                                      ((2, 3), (3, 1), (0, 1), (2, 0)), ((5, 3), (2, 1), (2, 1), (0, 0))])
def test_folded_operand_reproduces_conv_transpose(k, s, p, op):
    g = torch.Generator().manual_seed(3)
    cin, cout, B, H, W = 6, 5, 2, 7, 9
    x = torch.randn(B, cin, H, W, dtype=torch.float64, generator=g)
    w = torch.randn(cin, cout, *k, dtype=torch.float64, generator=g)          # ConvTranspose2d layout
    want = F.conv_transpose2d(x, w, stride=s, padding=p, output_padding=op)
    Ho, Wo = want.shape[2:]
    oh, ow = chain._phase_offsets(k[0], s[0], p[0]), chain._phase_offsets(k[1], s[1], p[1])
    es = sorted({e for v in oh.values() for _, e in v})
    fs = sorted({f for v in ow.values() for _, f in v})
    es, fs = list(range(es[0], es[-1] + 1)), list(range(fs[0], fs[-1] + 1))
    cg, nph = 8, s[0] * s[1]
    fold = dict(oh=oh, ow=ow, es=es, fs=fs, cg=cg, nph=nph)
    wf = chain.fold_transposed_weight(w.permute(1, 0, 2, 3).contiguous(), s[0], s[1], fold)   # packed is [Cout, Cin, kh, kw]
    # the engine's evaluation: forward convolution, stride 1, padding (-es[0], -fs[0]) on the top / left and zero fill
    # at the bottom / right, over the quotient grid; column group ph of quotient pixel (q, p) is output (q*sh + r, p*sw + c)
    Hq, Wq = -(-Ho // s[0]), -(-Wo // s[1])
    pt, pl = -es[0], -fs[0]
    pb, pr = max(0, Hq - 1 + es[-1] - (H - 1)), max(0, Wq - 1 + fs[-1] - (W - 1))
    xp = F.pad(x, (max(pl, 0), pr, max(pt, 0), pb))
    xp = xp[:, :, max(-pt, 0):, max(-pl, 0):]
    acc = F.conv2d(xp, wf)[:, :, :Hq, :Wq]                                     # [B, nph*cg, Hq, Wq]
    y = acc.view(B, s[0], s[1], cg, Hq, Wq)[:, :, :, :cout].permute(0, 3, 4, 1, 5, 2).reshape(B, cout, Hq * s[0], Wq * s[1])
    assert (y[:, :, :Ho, :Wo] - want).abs().max().item() < 1e-12
    assert (acc.view(B, s[0], s[1], cg, Hq, Wq)[:, :, :, cout:] == 0).all()   # pad columns of every phase stay zero


@pytest.mark.parametrize('affine', [True, False])
def test_cbn_closed_form_forward_backward_matches_module(affine):
    torch.manual_seed(0)
    B, C, H, W = 3, 4, 5, 7
    m = ComplexBatchNorm2d(2 * C, affine=affine).train()
    m.fused_train = False
    if affine:
        with torch.no_grad():
            m.Wrr.uniform_(0.8, 1.2); m.Wii.uniform_(0.8, 1.2); m.Br.normal_(); m.Bi.normal_()
    x = (torch.randn(B, 2 * C, H, W) * 2 + 1.5).requires_grad_()
    go = torch.randn(B, 2 * C, H, W)
    y = m(x)
    y.backward(go)
    xd = x.detach()
    xr, xi = xd[:, :C].double(), xd[:, C:].double()
    red, n = (0, 2, 3), B * H * W
    S = torch.stack([xr.sum(red), xi.sum(red), (xr * xr).sum(red), (xi * xi).sum(red), (xr * xi).sum(red)], 1)
    Wd = tuple(t.detach().double() for t in (m.Wrr, m.Wri, m.Wii)) if affine else None
    Bd = tuple(t.detach().double() for t in (m.Br, m.Bi)) if affine else None
    coef, M, V = _cbn_closed_form(S, n, m.eps, Wd, Bd)
    v = lambda t: t.view(1, C, 1, 1)
    yr = v(coef[0]) * xr + v(coef[1]) * xi + v(coef[4])
    yi = v(coef[2]) * xr + v(coef[3]) * xi + v(coef[5])
    assert (torch.cat([yr, yi], 1).float() - y.detach()).abs().max().item() < 5e-6
    # running statistics as the module's lerp_ left them after one step from (0, 1): momentum 0.1
    assert torch.allclose(m.RMr, 0.1 * M[0].float(), atol=1e-6) and torch.allclose(m.RVri, 0.1 * V[1].float(), atol=1e-6)
    gr, gi = go[:, :C].double(), go[:, C:].double()
    G = torch.stack([gr.sum(red), gi.sum(red), (gr * xr).sum(red), (gr * xi).sum(red), (gi * xr).sum(red),
                     (gi * xi).sum(red)], 1)
    S_ = S.clone().requires_grad_()
    leaves = [S_]
    W_ = B_ = None
    if affine:
        W_ = tuple(t.clone().requires_grad_() for t in Wd)
        B_ = tuple(t.clone().requires_grad_() for t in Bd)
        leaves += list(W_) + list(B_)
    coef2, _, _ = _cbn_closed_form(S_, n, m.eps, W_, B_)
    grads = torch.autograd.grad(list(coef2), leaves, [G[:, 2], G[:, 3], G[:, 4], G[:, 5], G[:, 0], G[:, 1]])
    dS = grads[0]
    dxr = v(coef[0]) * gr + v(coef[2]) * gi + v(dS[:, 0]) + 2 * v(dS[:, 2]) * xr + v(dS[:, 4]) * xi
    dxi = v(coef[1]) * gr + v(coef[3]) * gi + v(dS[:, 1]) + v(dS[:, 4]) * xr + 2 * v(dS[:, 3]) * xi
    assert (torch.cat([dxr, dxi], 1).float() - x.grad).abs().max().item() < 5e-6
    if affine:
        for got, p in zip(grads[1:], (m.Wrr, m.Wri, m.Wii, m.Br, m.Bi)):
            assert (got.float() - p.grad).abs().max().item() < 1e-5 * max(p.grad.abs().max().item(), 1.0)
