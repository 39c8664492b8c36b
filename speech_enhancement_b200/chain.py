"""Layer chaining on the tcgen05 conv engine: activations stay in CL2 between layers.

The reference runs every encoder / decoder block as  F.pad -> ComplexConv -> ComplexBatchNorm2d ->
activation  with a `complex_concat` / `torch.cat` of the skip in front of every decoder block
(_2008_00264_dccrn.py:27-33, :51-57, :113-121; _1903_03107_dcunet.py:84-95; _2206_07293_frcrn.py:90-102).
In eval mode all of that is linear glue around the convolution, so a `FusedBlock` folds it into the
conv kernel's operands and epilogue:

  * complex BN with running statistics is a per-channel 2x2 affine  y = Z (x - M) + B
    (complex_nn.py:252-322): Z is folded into the packed weight rows, Z(b - M) + B into the bias;
  * PReLU / LeakyReLU / ReLU are the epilogue's  v > 0 ? v : slope * v;
  * the causal frame pad is a coordinate shift with TMA zero fill (`pad_left`);
  * the skip concat is a second K range read from the skip's own CL2 tensor (`chan_map` permutes the
    weight columns at pack time), the T-trim before it is the tensor map's extent.

Everything here is host-side orchestration; the arithmetic is in libse_b200.so (se_cconv_gemm,
se_cconv_small_cin, se_cl2_from_nchw, se_cl2_to_nchw).
"""
import ctypes as C

import os
import torch
import torch.nn as nn

from . import _abi


def _cp(c):
    return (c + 63) // 64 * 64


class Act:
    """A CL2 activation: bf16 [B, 2, Hp, Wp, Cp] (value = hi + lo) with logical extent C x H x W."""
    __slots__ = ('data', 'C', 'H', 'W')

    def __init__(self, data, C_, H, W):
        self.data, self.C, self.H, self.W = data, C_, H, W

    @property
    def B(self):
        return self.data.shape[0]

    @property
    def Cp(self):
        return self.data.shape[4]

    def trimmed(self, H=None, W=None):
        """same storage, smaller logical extent (the reference's x[..., :-1] before a decoder block)"""
        return Act(self.data, self.C, self.H if H is None else H, self.W if W is None else W)


def cl2_empty(B, C_, H, W, device):
    return Act(torch.empty((B, 2, H, W, _cp(C_)), device=device, dtype=torch.bfloat16), C_, H, W)


def cl2_from_nchw(x):
    x = _abi.require_cuda_f32(x, 'cl2_from_nchw(x)').contiguous()
    B, C_, H, W = x.shape
    out = cl2_empty(B, C_, H, W, x.device)
    with torch.cuda.device(x.device), _abi.stage('cl2'):
        _abi.check(_abi.lib().se_cl2_from_nchw(_abi.ptr(x), B, C_, H, W, _abi.ptr(out.data), out.Cp,
                                               _abi.stream_ptr()))
    return out


def cl2_from_lstm_rows(rows, B, C_, F_, T):
    """[B, T, C*F] fp32 (the LSTM block's output, column part*(C/2)*F + c*F + f) -> CL2 Act [B, C, F, T]: the reference's
    permute(0, 2, 1).reshape(B, C, F, T) (_2008_00264_dccrn.py:164-165) fused with the layout conversion"""
    rows = _abi.require_cuda_f32(rows, 'cl2_from_lstm_rows(rows)')
    if rows.dim() != 3 or rows.shape[0] != B or rows.shape[1] != T or rows.shape[2] != C_ * F_ or rows.stride(2) != 1 \
            or rows.stride(0) != T * rows.stride(1):
        raise RuntimeError(f'cl2_from_lstm_rows: expected [B={B}, T={T}, {C_ * F_}] rows, got {tuple(rows.shape)}')
    out = cl2_empty(B, C_, F_, T, rows.device)
    with torch.cuda.device(rows.device), _abi.stage('cl2'):
        _abi.check(_abi.lib().se_cl2_from_lstm_rows(_abi.ptr(rows), rows.stride(1), B, C_, F_, T, _abi.ptr(out.data),
                                                    out.Cp, _abi.stream_ptr()))
    return out


def cl2_to_nchw(act):
    d = act.data
    if d.shape[2] != act.H or d.shape[3] != act.W:
        raise RuntimeError('cl2_to_nchw: trimmed views are not supported')
    y = torch.empty((act.B, act.C, act.H, act.W), device=d.device, dtype=torch.float32)
    with torch.cuda.device(d.device):
        _abi.check(_abi.lib().se_cl2_to_nchw(_abi.ptr(d), act.B, act.C, act.H, act.W, act.Cp, _abi.ptr(y),
                                             _abi.stream_ptr()))
    return y


def _pair(v):
    return (v, v) if isinstance(v, int) else tuple(v)


def _slope_of(act):
    if act is None or isinstance(act, nn.Identity):
        return 1.0, ()
    if isinstance(act, nn.PReLU):
        if act.weight.numel() != 1:
            raise RuntimeError('FusedBlock: per-channel PReLU is not supported')
        return None, (act.weight,)
    if isinstance(act, nn.LeakyReLU):
        return float(act.negative_slope), ()
    if isinstance(act, nn.ReLU):
        return 0.0, ()
    raise RuntimeError(f'FusedBlock: unsupported activation {type(act).__name__}')


def concat_channel_map(src_channels, mode):
    """module input channel behind every GEMM column, sources laid out back to back (each padded to 64).

    mode 'complex': the module sees complex_concat(sources) = [all real halves | all imaginary halves]
    (complex_nn.py:4-16); mode 'plain': torch.cat(sources, 1) (_1903_03107_dcunet.py:93)."""
    maps = []
    if mode == 'complex':
        half_total = sum(c // 2 for c in src_channels)
        off = 0
        for c in src_channels:
            h = c // 2
            m = [off + j for j in range(h)] + [half_total + off + j for j in range(h)]
            maps.append(m + [-1] * (_cp(c) - c))
            off += h
    elif mode == 'plain':
        off = 0
        for c in src_channels:
            maps.append(list(range(off, off + c)) + [-1] * (_cp(c) - c))
            off += c
    else:
        raise ValueError(mode)
    return [v for m in maps for v in m]


def _phase_offsets(k, s, p):
    """{phase r: [(tap d, input offset e)]}: output index s*q + r reads input index q + e through kernel tap d"""
    return {r: [(d, (r + p - d) // s) for d in range(k) if (r + p - d) % s == 0] for r in range(s)}


def _mma_cost(n):
    """cycles of the split-operand MMAs of one 16-channel k-step at N accumulator columns (scripts/mma_rate.cu:
    a 128 x N x 16 MMA costs max(N / 2, 84); concat form a_hi x [b_hi | b_lo] + a_lo x b_hi up to N = 128, three
    N-wide MMAs beyond)"""
    return max(n, 84) + max(n / 2, 84) if n <= 128 else 3 * max(n / 2, 84)


def fold_transposed_weight(p4, sh, sw, fold):
    """packed transposed-conv operand [Cout2, Cin2, kh, kw] (tap (dh, dw) takes input pixel (hi, wi) to output pixel
    (hi*sh - ph + dh, wi*sw - pw + dw)) -> the operand of the equivalent stride-1 FORWARD convolution over the input
    grid, [phases * cg, Cin2, n_e, n_f]: row (r*sw + s)*cg + co holds, at input offset (e, f), the tap through which
    output phase (r, s) reads input pixel (q + e, p + f); zero where the phase has no such tap."""
    es, fs, cg, nph = fold['es'], fold['fs'], fold['cg'], fold['nph']
    cout2, cin2 = p4.shape[0], p4.shape[1]
    wf = torch.zeros((nph * cg, cin2, len(es), len(fs)), device=p4.device, dtype=p4.dtype)
    for r in range(sh):
        for sph in range(sw):
            c0 = (r * sw + sph) * cg
            for dh, e in fold['oh'][r]:
                for dw, fo in fold['ow'][sph]:
                    wf[c0:c0 + cout2, :, e - es[0], fo - fs[0]] = p4[:, :, dh, dw]
    return wf


def fold_pixel_pairs_weight(p4, axis, stride, cg):
    """forward-convolution operand [Cout2, Cin2, kh, kw] (out[q] = sum_d x[q*s - p + d] w[d]) -> the operand of the
    convolution that computes TWO adjacent output pixels along `axis` (0 = rows, 1 = columns) per GEMM row:
    [2 * cg, Cin2, kh', kw'] with the folded axis extended to k + s taps and walked at stride 2 s.  Output 2 q + ph
    reads input q * 2s - p + (ph * s + d), so column group ph carries w[d] at tap ph * s + d and zeros elsewhere."""
    cout2, cin2, kh, kw = p4.shape
    ext = (kh + stride, kw) if axis == 0 else (kh, kw + stride)
    wf = torch.zeros((2 * cg, cin2) + ext, device=p4.device, dtype=p4.dtype)
    for ph in range(2):
        o = ph * stride
        if axis == 0:
            wf[ph * cg:ph * cg + cout2, :, o:o + kh, :] = p4
        else:
            wf[ph * cg:ph * cg + cout2, :, :, o:o + kw] = p4
    return wf


# SE_CONV_FOLD: 'auto' (fold where the MMA cost model says it pays), '0' never, '1' wherever the shape allows
FOLD_MODE = os.environ.get('SE_CONV_FOLD', 'auto')


class FusedBlock:
    fold = None                 # phase-folding plan of a transposed layer (set in __init__ where it applies)

    """conv / convT (+ eval-mode complex BN) (+ leaky activation) of one reference block, prepared for
    the chained engine.  `conv` is a ComplexConv2d / ComplexConvTranspose2d of this package."""

    def __init__(self, conv, norm=None, act=None, src_channels=None, concat='complex'):
        self.conv, self.norm, self.act = conv, norm, act
        rc = conv.real_conv
        self.transposed = bool(conv._transposed)
        if self.transposed:
            ci, co = rc.weight.shape[0], rc.weight.shape[1]
        else:
            co, ci = rc.weight.shape[0], rc.weight.shape[1]
        self.cin2, self.cout2 = 2 * ci, 2 * co
        self.kh, self.kw = rc.weight.shape[2], rc.weight.shape[3]
        self.sh, self.sw = _pair(rc.stride)
        self.ph, self.pw = _pair(rc.padding)
        self.oph, self.opw = _pair(rc.output_padding) if self.transposed else (0, 0)
        self.src_channels = list(src_channels) if src_channels is not None else [self.cin2]
        if sum(self.src_channels) != self.cin2:
            raise RuntimeError(f'FusedBlock: sources {self.src_channels} do not add up to {self.cin2} channels')
        self.concat = concat
        # transposed layers with 2 or 4 output channels run in scatter form (1x1 GEMM over the taps + gather):
        # one pass over the input instead of one >= 84-cycle MMA per tap for a handful of useful columns
        self.scatter_form = self.transposed and self.cout2 in (2, 4) and self.kh * self.kw * self.cout2 <= 128
        # transposed layers with few output channels and stride > 1 run PHASE-FOLDED: the sh*sw output phases read the
        # same input tile through different taps, so they become column groups of ONE stride-1 convolution over the
        # input grid (n_e x n_f input offsets instead of kh x kw taps, N = phases * channels instead of channels);
        # every MMA pays the 128-row A-operand read (>= 84 cycles) however narrow N is, so fewer, wider MMAs win
        self.fold = None
        if self.transposed and not self.scatter_form and self.sh * self.sw > 1 and FOLD_MODE != '0':
            oh, ow = _phase_offsets(self.kh, self.sh, self.ph), _phase_offsets(self.kw, self.sw, self.pw)
            cg = (self.cout2 + 31) // 32 * 32
            unfolded = self.kh * self.kw * _mma_cost((self.cout2 + 15) // 16 * 16)
            best = None
            # fold both axes, or only one (the other axis' phases then get one launch each).  'auto' keeps to the concat
            # operand form (<= 128 folded columns): measured on DCUNet dec7 (4 phases x 64 = 256 columns, split form) the
            # fully folded launch LOSES, 6.97 -> 9.98 ms (weight-slab fill bound), although the MMA model favours it
            for fh, fw in ((self.sh, self.sw), (self.sh, 1), (1, self.sw)):
                if fh * fw <= 1 or fh * fw * cg > 256:
                    continue
                launches, cost = [], 0.0
                for rh in ([None] if fh > 1 or self.sh == 1 else range(self.sh)):
                    for rw in ([None] if fw > 1 or self.sw == 1 else range(self.sw)):
                        poh = oh if rh is None else {0: oh[rh]}
                        pow_ = ow if rw is None else {0: ow[rw]}
                        es = sorted({e for v in poh.values() for _, e in v})
                        fs = sorted({f for v in pow_.values() for _, f in v})
                        if not es or not fs:
                            launches = None            # a phase no tap reaches (kernel smaller than the stride): not folded
                            break
                        es, fs = list(range(es[0], es[-1] + 1)), list(range(fs[0], fs[-1] + 1))
                        launches.append(dict(rh=rh, rw=rw, oh=poh, ow=pow_, es=es, fs=fs))
                        cost += len(es) * len(fs) * _mma_cost(fh * fw * cg)
                    if launches is None:
                        break
                if not launches or any(len(l['es']) * len(l['fs']) > 49 for l in launches):
                    continue
                gain = unfolded / cost                # both per pixel of the full quotient grid (sh * sw output pixels)
                if FOLD_MODE != '1' and (gain <= 1.1 or fh * fw * cg > 128):
                    continue
                if best is None or gain > best['gain']:
                    best = dict(kind='phase', fh=fh, fw=fw, cg=cg, nph=fh * fw, launches=launches, gain=gain)
            self.fold = best
        # narrow layers WITHOUT output phases (forward convolutions, stride-1 transposed ones): two adjacent output
        # pixels along one axis share a GEMM row -- that axis is walked at twice the stride with k + s taps, the two
        # pixels are two column groups (N = 2 x channels).  k taps of N columns per pixel become (k + s) / 2 of 2 N.
        # MEASURED (profiles/r2_pixel_fold_layers.txt, 64 x 4 s): no gain -- DCUNet 7x1 1.61 -> 1.64 ms, 1x7 1.96 -> 1.94,
        # transposed 1x7 3.61 -> 3.43, DCCRN enc1 (stride 2, element stride 4 in the TMA box) 0.47 -> 0.62 ms: these
        # short-K layers are bound by per-tile fill + epilogue, not by the MMA count the model looks at.  So this form
        # is built only on request (SE_CONV_FOLD=1; tests/test_gpu_fold.py keeps it correct), never in 'auto'.
        if (self.fold is None and not self.scatter_form and FOLD_MODE != '0'
                and (not self.transposed or (self.sh, self.sw) == (1, 1))):
            cg = (self.cout2 + 31) // 32 * 32
            npad = (self.cout2 + 15) // 16 * 16
            best = None
            for axis, (k, st) in enumerate(((self.kh, self.sh), (self.kw, self.sw))):
                gain = 2 * k * _mma_cost(npad) / ((k + st) * _mma_cost(2 * cg))
                if best is None or gain > best[1]:
                    best = (axis, gain)
            if 2 * cg <= 256 and (self.kh + self.sh) * (self.kw + self.sw) <= 49 and \
                    FOLD_MODE == '1':
                self.fold = dict(kind='pixel', axis=best[0], cg=cg, nph=2, gain=best[1])
        self._key = None
        self._ops = None

    # ---- folded operands, rebuilt when any parameter / buffer changes
    def _operands(self):
        conv, norm = self.conv, self.norm
        slope_const, slope_t = _slope_of(self.act)
        ts = self._param_tensors()
        has_bn = norm is not None and not isinstance(norm, nn.Identity)
        if has_bn:
            if norm.training or not norm.track_running_stats:
                raise RuntimeError('FusedBlock: complex BN can only be folded in eval mode with running stats')
            ts += [norm.RMr, norm.RMi, norm.RVrr, norm.RVri, norm.RVii]
            if norm.affine:
                ts += [norm.Wrr, norm.Wri, norm.Wii, norm.Br, norm.Bi]
        ts += list(slope_t)
        key = _abi.cache_key(ts)
        if key == self._key:
            return self._ops
        with torch.no_grad():
            packed, bias, co, ci, kh, kw = self._packed_source()    # fp32 [2co, 2ci*kh*kw], [2co] | None
            dev = packed.device
            if bias is None:
                bias = torch.zeros(2 * co, device=dev, dtype=torch.float32)
            if has_bn:
                Zrr, Zri, Zir, Zii = norm._affine_2x2(norm.RVrr, norm.RVri, norm.RVii)
                Wr, Wi = packed[:co], packed[co:]
                br, bi = bias[:co] - norm.RMr, bias[co:] - norm.RMi
                packed = torch.cat([Zrr[:, None] * Wr + Zri[:, None] * Wi, Zir[:, None] * Wr + Zii[:, None] * Wi], 0)
                nbr, nbi = Zrr * br + Zri * bi, Zir * br + Zii * bi
                if norm.affine:
                    nbr, nbi = nbr + norm.Br, nbi + norm.Bi
                bias = torch.cat([nbr, nbi])
            packed = packed.contiguous().float()
            bias = bias.contiguous().float()
            slope = slope_const if slope_const is not None else float(slope_t[0].detach().float().item())
            cmap = concat_channel_map(self.src_channels, self.concat)
            cpt = len(cmap)
            cmap_t = torch.tensor(cmap, device=dev, dtype=torch.int32)
            lib = _abi.lib()
            taps = kh * kw
            wq = torch.empty((lib.se_cconv_wq_elems(2 * co, cpt, taps),), device=dev, dtype=torch.bfloat16)
            with torch.cuda.device(dev):
                _abi.check(lib.se_cconv_prep_weight(_abi.ptr(packed), 2 * co, 2 * ci, taps, _abi.ptr(cmap_t), cpt,
                                                    _abi.ptr(wq), _abi.stream_ptr()))
            wq_sc = None
            if self.scatter_form:
                # taps folded into the output columns of a 1x1 convolution: row tap*2co + o <- packed[o, :, tap]
                p3 = packed.view(2 * co, 2 * ci, taps).permute(2, 0, 1).reshape(taps * 2 * co, 2 * ci).contiguous()
                wq_sc = torch.empty((lib.se_cconv_wq_elems(taps * 2 * co, cpt, 1),), device=dev, dtype=torch.bfloat16)
                with torch.cuda.device(dev):
                    _abi.check(lib.se_cconv_prep_weight(_abi.ptr(p3), taps * 2 * co, 2 * ci, 1, _abi.ptr(cmap_t), cpt,
                                                        _abi.ptr(wq_sc), _abi.stream_ptr()))
            wq_fold = bias_fold = None
            if self.fold is not None:
                f = self.fold
                cg, nph = f['cg'], f['nph']
                p4 = packed.view(2 * co, 2 * ci, kh, kw)
                if f['kind'] == 'phase':
                    # one operand per launch: the phases of a folded axis side by side, one phase of an unfolded axis
                    wfs = [fold_transposed_weight(p4, f['fh'], f['fw'], dict(l, cg=cg, nph=nph)) for l in f['launches']]
                else:
                    if self.transposed:       # stride-1 transposed convolution = forward convolution with flipped taps
                        p4 = p4.flip(2, 3)
                    wfs = [fold_pixel_pairs_weight(p4, f['axis'], self.sh if f['axis'] == 0 else self.sw, cg)]
                bias_fold = torch.zeros((nph, cg), device=dev, dtype=torch.float32)
                bias_fold[:, :2 * co] = bias
                bias_fold = bias_fold.reshape(-1).contiguous()
                wq_fold = []
                for wf in wfs:
                    ntap = wf.shape[2] * wf.shape[3]
                    wf = wf.reshape(nph * cg, 2 * ci * ntap).contiguous()
                    wq_f = torch.empty((lib.se_cconv_wq_elems(nph * cg, cpt, ntap),), device=dev, dtype=torch.bfloat16)
                    with torch.cuda.device(dev):
                        _abi.check(lib.se_cconv_prep_weight(_abi.ptr(wf), nph * cg, 2 * ci, ntap, _abi.ptr(cmap_t), cpt,
                                                            _abi.ptr(wq_f), _abi.stream_ptr()))
                    wq_fold.append(wq_f)
        self._ops = {'packed': packed, 'bias': bias, 'slope': slope, 'wq': wq, 'cpt': cpt, 'wq_sc': wq_sc,
                     'wq_fold': wq_fold, 'bias_fold': bias_fold}
        self._key = key
        return self._ops

    def _param_tensors(self):
        conv = self.conv
        ts = [conv.real_conv.weight, conv.imag_conv.weight]
        if conv.real_conv.bias is not None:
            ts += [conv.real_conv.bias, conv.imag_conv.bias]
        return ts

    def _packed_source(self):
        return self.conv._packed()

    def out_size(self, H, W, pad_top=0, pad_left=0):
        if self.transposed:
            return ((H - 1) * self.sh - 2 * self.ph + self.kh + self.oph,
                    (W - 1) * self.sw - 2 * self.pw + self.kw + self.opw)
        return ((H + pad_top + 2 * self.ph - self.kh) // self.sh + 1,
                (W + pad_left + 2 * self.pw - self.kw) // self.sw + 1)

    def __call__(self, srcs, pad_top=0, pad_left=0, out_cl2=True, out_nchw=False, y_out=None, y_strides=None):
        """srcs: list of Act with the same logical H, W.  Returns (Act | None, NCHW fp32 tensor | None).
        `y_out` (fp32 CUDA tensor) + `y_strides` = (batch, channel, row, column) strides in elements: the fp32 copy of
        the output is written there instead of a fresh NCHW tensor -- e.g. straight into the [B, T, C*F] rows the LSTM
        block reads (_2008_00264_dccrn.py:161-163), with no permute / reshape copy in between."""
        ops = self._operands()
        if len(srcs) != len(self.src_channels) or any(s.C != c for s, c in zip(srcs, self.src_channels)):
            raise RuntimeError(f'FusedBlock: expected sources with channels {self.src_channels}')
        # sources may differ in extent: the smaller one is zero-extended at its bottom / right edge
        H, W, B = max(s.H for s in srcs), max(s.W for s in srcs), srcs[0].B
        if any(s.B != B for s in srcs):
            raise RuntimeError('FusedBlock: sources disagree on the batch size')
        Ho, Wo = self.out_size(H, W, pad_top, pad_left)
        dev = srcs[0].data.device
        y_act = cl2_empty(B, self.cout2, Ho, Wo, dev) if out_cl2 else None
        y = torch.empty((B, self.cout2, Ho, Wo), device=dev, dtype=torch.float32) if out_nchw else None
        if y_out is not None:
            if out_nchw or self.scatter_form or y_out.dtype != torch.float32 or not y_out.is_cuda or y_strides is None:
                raise RuntimeError('FusedBlock: y_out needs y_strides, fp32 CUDA storage and out_nchw=False')
            y = y_out
        a = _abi.ConvArgs()
        a.transposed, a.B, a.nsrc = int(self.transposed), B, len(srcs)
        for i, s in enumerate(srcs):
            a.src[i] = s.data.data_ptr()
            a.src_Cp[i], a.src_Hpitch[i], a.src_Wpitch[i] = s.Cp, s.data.shape[2], s.data.shape[3]
            a.src_H[i], a.src_W[i] = s.H, s.W
            a.src_C[i] = s.C
        a.H, a.W, a.pad_top, a.pad_left = H, W, pad_top, pad_left
        a.Cout2, a.kh, a.kw, a.sh, a.sw, a.ph, a.pw, a.Ho, a.Wo = (self.cout2, self.kh, self.kw, self.sh,
                                                                    self.sw, self.ph, self.pw, Ho, Wo)
        a.wq, a.bias, a.slope = ops['wq'].data_ptr(), ops['bias'].data_ptr(), ops['slope']
        a.y_nchw = y.data_ptr() if y is not None else None
        a.y_cl2 = y_act.data.data_ptr() if y_act is not None else None
        a.y_Cp = y_act.Cp if y_act is not None else 0
        if y_out is not None:
            a.y_bs, a.y_cs, a.y_hs, a.y_ws = (int(v) for v in y_strides)
        if self.scatter_form and out_nchw and not out_cl2:
            ncol = self.kh * self.kw * self.cout2
            Np = (ncol + 3) // 4 * 4
            P = torch.empty((B, H, W, Np), device=dev, dtype=torch.float32)
            a.transposed, a.Cout2, a.kh, a.kw, a.sh, a.sw, a.ph, a.pw, a.Ho, a.Wo = 0, ncol, 1, 1, 1, 1, 0, 0, H, W
            a.wq, a.bias, a.slope = ops['wq_sc'].data_ptr(), None, 1.0
            a.y_nchw = P.data_ptr()
            a.y_cs, a.y_ws, a.y_hs, a.y_bs = 1, Np, W * Np, H * W * Np
            with torch.cuda.device(dev), _abi.stage('cconv'):
                lib = _abi.lib()
                _abi.check(lib.se_cconv_gemm(C.byref(a), _abi.stream_ptr()))
                _abi.check(lib.se_cconvT_col2im(_abi.ptr(P), B, H, W, Np, self.cout2, self.kh, self.kw, self.sh, self.sw,
                                                self.ph, self.pw, Ho, Wo, _abi.ptr(ops['bias']), ops['slope'],
                                                _abi.ptr(y), _abi.stream_ptr()))
            return None, y
        if self.fold is not None and y is None and y_act is not None:
            f = self.fold
            a.transposed = 0
            a.Cout2 = f['nph'] * f['cg']
            a.fold_cg, a.fold_cout2 = f['cg'], self.cout2
            a.bias = ops['bias_fold'].data_ptr()
            if f['kind'] == 'phase':
                # phase-folded form: stride-1 forward convolution(s) over the input grid, the phases of the folded axes as
                # column groups; an unfolded axis with stride > 1 gets one launch per phase, each covering every s-th row /
                # column of the output
                a.sh, a.sw, a.fold_h, a.fold_w = 1, 1, f['fh'], f['fw']
                with torch.cuda.device(dev), _abi.stage('cconv'):
                    for l, wq_f in zip(f['launches'], ops['wq_fold']):
                        a.kh, a.kw = len(l['es']), len(l['fs'])
                        a.ph, a.pw = -l['es'][0], -l['fs'][0]
                        a.out_h_mul, a.out_h_add = (self.sh, l['rh']) if l['rh'] is not None else (0, 0)
                        a.out_w_mul, a.out_w_add = (self.sw, l['rw']) if l['rw'] is not None else (0, 0)
                        a.wq = wq_f.data_ptr()
                        _abi.check(_abi.lib().se_cconv_gemm(C.byref(a), _abi.stream_ptr()))
                return y_act, y
            # pixel-pair form: the folded axis is walked at twice the stride with k + s taps
            kh, kw, sh, sw, ph, pw = self.kh, self.kw, self.sh, self.sw, self.ph, self.pw
            if self.transposed:           # forward equivalent of a stride-1 transposed convolution
                ph, pw = kh - 1 - ph, kw - 1 - pw
            if f['axis'] == 0:
                a.kh, a.kw, a.sh, a.sw, a.fold_h, a.fold_w = kh + sh, kw, 2 * sh, sw, 2, 1
            else:
                a.kh, a.kw, a.sh, a.sw, a.fold_h, a.fold_w = kh, kw + sw, sh, 2 * sw, 1, 2
            a.ph, a.pw = ph, pw
            a.wq = ops['wq_fold'][0].data_ptr()
        with torch.cuda.device(dev), _abi.stage('cconv'):
            _abi.check(_abi.lib().se_cconv_gemm(C.byref(a), _abi.stream_ptr()))
        return y_act, y

    def first_layer(self, x, xs_b, xs_c, xs_h, B, H, W, pad_top=0, pad_left=0, out_cl2=True, out_nchw=False):
        """the 2-input-channel first encoder layer, reading fp32 planes in place through strides
        (`x` is any tensor whose data_ptr is the first element of channel 0)."""
        if self.transposed:
            raise RuntimeError('FusedBlock.first_layer: forward convolution only')
        ops = self._operands()
        Ho, Wo = self.out_size(H, W, pad_top, pad_left)
        dev = x.device
        y_act = cl2_empty(B, self.cout2, Ho, Wo, dev) if out_cl2 else None
        y = torch.empty((B, self.cout2, Ho, Wo), device=dev, dtype=torch.float32) if out_nchw else None
        with torch.cuda.device(dev), _abi.stage('cconv'):
            _abi.check(_abi.lib().se_cconv_small_cin(
                _abi.ptr(x), xs_b, xs_c, xs_h, B, self.cin2, H, W, pad_top, pad_left, _abi.ptr(ops['packed']),
                _abi.ptr(ops['bias']), ops['slope'], self.cout2, self.kh, self.kw, self.sh, self.sw, self.ph,
                self.pw, _abi.ptr(y), _abi.ptr(y_act.data) if y_act is not None else None,
                y_act.Cp if y_act is not None else 0, Ho, Wo, _abi.stream_ptr()))
        return y_act, y


class RealConvBlock(FusedBlock):
    """a plain (real) nn.Conv2d with even channel counts on the chained engine: its weight is already the
    packed operand [Cout, Cin*kh*kw] (FRCRN's final_conv, /root/reference/models/frcrn.py:final_conv)."""

    def __init__(self, conv, act=None):
        w = conv.weight
        if w.shape[0] % 2 or w.shape[1] % 2 or _pair(conv.dilation) != (1, 1) or conv.groups != 1:
            raise RuntimeError('RealConvBlock: needs even channel counts, no dilation, groups == 1')
        self.conv, self.norm, self.act = conv, None, act
        self.transposed = False
        self.cout2, self.cin2 = w.shape[0], w.shape[1]
        self.kh, self.kw = w.shape[2], w.shape[3]
        self.sh, self.sw = _pair(conv.stride)
        self.ph, self.pw = _pair(conv.padding)
        self.oph, self.opw = 0, 0
        self.src_channels = [self.cin2]
        self.concat = 'plain'
        self.scatter_form = False
        self._key = None
        self._ops = None

    def _param_tensors(self):
        return [self.conv.weight] + ([self.conv.bias] if self.conv.bias is not None else [])

    def _packed_source(self):
        w = self.conv.weight.detach().float()
        b = self.conv.bias.detach().float() if self.conv.bias is not None else None
        return w.reshape(w.shape[0], -1), b, w.shape[0] // 2, w.shape[1] // 2, w.shape[2], w.shape[3]


# side streams for independent linear-projection launches (0: all on the current stream; SE_LINEAR_STREAMS overrides)
FANOUT_STREAMS = int(os.environ.get('SE_LINEAR_STREAMS', '4'))
# output columns per GEMM launch of a linear layer (<= 256: one accumulator's TMEM columns)
LINEAR_CHUNK = int(os.environ.get('SE_LINEAR_CHUNK', '256'))
_SIDE = {}


def _side_streams(dev, n):
    pool = _SIDE.setdefault(torch.device(dev).index if torch.device(dev).index is not None else torch.cuda.current_device(), [])
    while len(pool) < n:
        pool.append(torch.cuda.Stream(device=dev))
    return pool[:n]


class LinearOperand:
    """y = x W^T + b on the bf16x3 GEMM engine for one or several stacked nn.Linear-style weights
    (the LSTM input projections and ComplexLinear of /root/reference/models/modules/complex_nn.py:93-145).

    `parts` is a callable returning [(weight [N_i, K], bias [N_i] | None | tuple of biases to sum), ...];
    the operand (weights split into bf16 hi|lo, <= 256 output columns per launch) is rebuilt whenever one
    of the tensors changes.  x may be a column slice of a wider row-major matrix (row stride `ld`).
    """

    def __init__(self, parts):
        self._parts = parts
        self._key = None
        self._ops = None

    def _operands(self):
        parts = self._parts()
        flat = []
        for w, b in parts:
            flat.append(w)
            flat.extend(b if isinstance(b, (tuple, list)) else [b])
        key = _abi.cache_key(flat)
        if key != self._key:
            with torch.no_grad():
                ws, bs = [], []
                for w, b in parts:
                    ws.append(w.detach().float())
                    if isinstance(b, (tuple, list)):
                        b = sum(t.detach().float() for t in b if t is not None) if any(t is not None for t in b) else None
                    bs.append(b.detach().float() if b is not None else torch.zeros(w.shape[0], device=w.device))
                W = torch.cat(ws, 0).contiguous()
                bias = torch.cat(bs, 0).contiguous()
                N, K = W.shape
                Cp = _cp(K)
                chunks = []
                lib = _abi.lib()
                with torch.cuda.device(W.device):
                    for c0 in range(0, N, LINEAR_CHUNK):
                        n = min(LINEAR_CHUNK, N - c0)
                        wq = torch.empty(lib.se_cconv_wq_elems(n, Cp, 1), device=W.device, dtype=torch.bfloat16)
                        _abi.check(lib.se_cconv_prep_weight(_abi.ptr(W[c0:c0 + n]), n, K, 1, None, Cp, _abi.ptr(wq),
                                                            _abi.stream_ptr()))
                        chunks.append((c0, n, wq))
            self._ops = dict(N=N, K=K, Cp=Cp, chunks=chunks, bias=bias, W=W)
            self._key = key
        return self._ops

    def __call__(self, x, out=None, out_col=0):
        """x: [M, K] fp32 CUDA with unit column stride (row stride free).  Writes columns
        [out_col, out_col + N) of `out` ([M, >= out_col + N] fp32 contiguous; allocated when None)."""
        return self.many([(x, out, out_col)])[0]

    def many(self, jobs):
        """several independent products [(x, out | None, out_col), ...] with this operand (see `linear_many`)"""
        return linear_many([(self, x, out, col) for x, out, col in jobs])


def linear_many(jobs):
    """Several independent products [(operand, x, out | None, out_col), ...].  Every (job, <= 256-column
    chunk) pair is its own persistent GEMM launch of one CTA per SM whose last wave is mostly empty (161 tiles on
    148 SMs for the DCCRN LSTM projection), so the launches are dealt round-robin to side streams and the block
    scheduler fills the idle SMs of one launch's tail with the next launch's CTAs; joined before returning."""
    lib = _abi.lib()
    outs, launches = [], []
    dev = jobs[0][1].device
    with torch.cuda.device(dev), _abi.stage('linear'):
        for op, x, out, out_col in jobs:
            ops = op._operands()
            N, Cp = ops['N'], ops['Cp']
            M, K = x.shape
            if K != ops['K'] or x.stride(1) != 1 or x.dtype != torch.float32 or not x.is_cuda:
                raise RuntimeError(f'LinearOperand: expected fp32 CUDA [M, {ops["K"]}] with unit column stride')
            if out is None:
                out = torch.empty((M, N), device=x.device, dtype=torch.float32)
            outs.append(out)
            if M == 0:
                continue
            ldo = out.stride(0)
            H, W = (2, M // 2) if M % 2 == 0 else (1, M)   # two image rows let two accumulators share a weight unit
            img = torch.empty((1, 2, H, W, Cp), device=x.device, dtype=torch.bfloat16)
            _abi.check(lib.se_cl2_from_rows(_abi.ptr(x), M, K, x.stride(0), _abi.ptr(img), Cp, _abi.stream_ptr()))
            for c0, n, wq in ops['chunks']:
                a = _abi.ConvArgs()
                a.transposed, a.B, a.nsrc = 0, 1, 1
                a.src[0] = img.data_ptr()
                a.src_Cp[0], a.src_Hpitch[0], a.src_Wpitch[0] = Cp, H, W
                a.H, a.W, a.Cout2, a.kh, a.kw, a.sh, a.sw, a.Ho, a.Wo = H, W, n, 1, 1, 1, 1, H, W
                a.wq, a.bias, a.slope = wq.data_ptr(), ops['bias'].data_ptr() + 4 * c0, 1.0
                a.y_nchw = out.data_ptr() + 4 * (out_col + c0)
                a.y_cs, a.y_ws, a.y_hs = 1, ldo, W * ldo
                launches.append((a, img))
        if len(launches) <= 1 or not FANOUT_STREAMS:
            for a, _ in launches:
                _abi.check(lib.se_cconv_gemm(C.byref(a), _abi.stream_ptr()))
        else:
            cur = torch.cuda.current_stream(dev)
            side = _side_streams(dev, min(FANOUT_STREAMS, len(launches)))
            fork = torch.cuda.Event()
            fork.record(cur)
            for i, (a, _) in enumerate(launches):
                s = side[i % len(side)]
                if i < len(side):
                    s.wait_event(fork)
                _abi.check(lib.se_cconv_gemm(C.byref(a), C.c_void_p(s.cuda_stream)))
            for s in side:
                cur.wait_stream(s)                          # `img` / `out` are next touched on `cur`, after the join
    return outs
