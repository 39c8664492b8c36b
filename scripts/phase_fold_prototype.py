"""CPU prototype of the phase-folded transposed convolution planned for the small-N decoder layers.

A ConvTranspose2d with stride (sh, sw) computes sh*sw output phases (ho % sh, wo % sw) from the SAME input tile with
different taps.  Folding the phases into the GEMM's N dimension turns kh*kw MMAs of N = Cout columns into
n_e * n_f MMAs of N = sh*sw*Cout columns (n_e, n_f = distinct input-row / -column offsets over all phases), with zero
weights where a phase has no tap at an offset.  This script builds the folded weights, evaluates the folded form with
plain matmuls and checks it against F.conv_transpose2d; it also prints the MMA counts of both forms.

  python scripts/phase_fold_prototype.py
"""
import itertools

import torch
import torch.nn.functional as F


def offsets(k, s, p):
    """{phase r: [(tap d, input offset e)]} with ho = s*q + r reading input row q + e through tap d"""
    out = {}
    for r in range(s):
        out[r] = [(d, (r + p - d) // s) for d in range(k) if (r + p - d) % s == 0]
    return out


def fold_weights(w, stride, padding):
    """w [Ci, Co, kh, kw] -> ({(e, f): W [Ci, sh*sw*Co]}, sorted e list, sorted f list); column (r*sw + s)*Co + co"""
    Ci, Co, kh, kw = w.shape
    (sh, sw), (ph, pw) = stride, padding
    oh, ow = offsets(kh, sh, ph), offsets(kw, sw, pw)
    es = sorted({e for v in oh.values() for _, e in v})
    fs = sorted({f for v in ow.values() for _, f in v})
    folded = {(e, f): torch.zeros(Ci, sh * sw * Co, dtype=w.dtype) for e in es for f in fs}
    for r, s in itertools.product(range(sh), range(sw)):
        for (dh, e), (dw, f) in itertools.product(oh[r], ow[s]):
            c0 = (r * sw + s) * Co
            folded[(e, f)][:, c0:c0 + Co] = w[:, :, dh, dw]
    return folded, es, fs


def conv_transpose_folded(x, w, stride, padding, output_padding):
    B, Ci, H, W = x.shape
    Co, kh, kw = w.shape[1:]
    (sh, sw), (ph, pw), (oph, opw) = stride, padding, output_padding
    Ho, Wo = (H - 1) * sh - 2 * ph + kh + oph, (W - 1) * sw - 2 * pw + kw + opw
    folded, es, fs = fold_weights(w, stride, padding)
    Hq, Wq = -(-Ho // sh), -(-Wo // sw)                       # quotient grid the GEMM's M dimension walks
    pad_t, pad_l = max(0, -es[0]), max(0, -fs[0])
    pad_b, pad_r = max(0, Hq - 1 + es[-1] - (H - 1)), max(0, Wq - 1 + fs[-1] - (W - 1))
    xp = F.pad(x, (pad_l, pad_r, pad_t, pad_b))               # the engine gets these zeros from TMA out-of-bounds fill
    acc = torch.zeros(B, Hq, Wq, sh * sw * Co, dtype=x.dtype)
    for e, f in itertools.product(es, fs):                    # one MMA group per (row offset, column offset)
        tile = xp[:, :, pad_t + e:pad_t + e + Hq, pad_l + f:pad_l + f + Wq]          # [B, Ci, Hq, Wq]
        acc += torch.einsum('bchw,cn->bhwn', tile, folded[(e, f)])
    # epilogue: column range (r, s) of quotient pixel (q, p) is output pixel (sh*q + r, sw*p + s)
    y = acc.view(B, Hq, Wq, sh, sw, Co).permute(0, 5, 1, 3, 2, 4).reshape(B, Co, Hq * sh, Wq * sw)
    return y[:, :, :Ho, :Wo], len(es) * len(fs)


if __name__ == '__main__':
    torch.manual_seed(0)
    cases = [  # (name, Ci, Co, k, stride, padding, output_padding, H, W)
        ('DCCRN dec4', 128, 32, (5, 2), (2, 1), (2, 0), (1, 0), 16, 21),
        ('DCCRN dec3', 256, 64, (5, 2), (2, 1), (2, 0), (1, 0), 8, 21),
        ('DCUNet dec7', 256, 64, (7, 5), (2, 2), (3, 2), (0, 0), 17, 13),
        ('DCUNet dec5', 256, 128, (5, 3), (2, 2), (2, 1), (0, 0), 9, 11),
        ('FRCRN dec', 256, 128, (5, 2), (2, 1), (0, 0), (0, 0), 7, 12),
    ]
    for name, Ci, Co, k, st, pd, op, H, W in cases:
        x = torch.randn(2, Ci, H, W, dtype=torch.float64)
        w = torch.randn(Ci, Co, *k, dtype=torch.float64) / (Ci * k[0] * k[1]) ** 0.5
        want = F.conv_transpose2d(x, w, stride=st, padding=pd, output_padding=op)
        got, groups = conv_transpose_folded(x, w, st, pd, op)
        err = (got - want).abs().max().item()
        taps = k[0] * k[1]
        print(f'{name:12s} k={k} s={st}: max |err| {err:.2e}; per 16-channel k-step: {taps} MMAs of N={Co} today -> '
              f'{groups} MMAs of N={st[0] * st[1] * Co} folded ({taps}/{groups * st[0] * st[1]} weight slots used)')
        assert got.shape == want.shape and err < 1e-12
