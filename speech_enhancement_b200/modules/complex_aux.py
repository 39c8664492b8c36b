"""Out-of-scope neighbours of the hot path, on stock PyTorch.

Same class names, constructor signatures, parameter/buffer names and numerics as
/root/reference/models/modules/complex_nn.py:93-360 so reference checkpoints load and the
caller models behave identically; they are the "next" rows of SURVEY.md 8f.  ComplexLSTM inference on CUDA runs its time loop
in csrc/lstm_rec.cu (8f rank 3); everything else here is stock PyTorch on whatever device the
input lives on.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import _abi
from .complex_nn import split_complex, merge_real_imag, _halve


def _lstm_rec_ok(lstm, x):
    """the persistent recurrent kernel (csrc/lstm_rec.cu) serves inference on a unidirectional fp32 LSTM"""
    return (x.is_cuda and x.dtype == torch.float32 and not lstm.bidirectional and lstm.proj_size == 0
            and not (torch.is_grad_enabled() and (x.requires_grad or lstm.weight_hh_l0.requires_grad))
            and not (lstm.training and lstm.dropout > 0 and lstm.num_layers > 1)
            and _abi.lib().se_lstm_rec_supported(lstm.hidden_size) == 1)


def _lstm_layer_params(lstm, layer):
    g = lambda n: getattr(lstm, f'{n}_l{layer}')
    bias = (g('bias_ih'), g('bias_hh')) if lstm.bias else None
    return g('weight_ih'), g('weight_hh'), bias


def _lstm_rec(pre, col0, w_hh, N, T, H):
    """time loop of one LSTM layer over pre-gates pre[N*T, ld] (columns col0 .. col0+4H)."""
    h = torch.empty((N, T, H), device=pre.device, dtype=torch.float32)
    with _abi.stage('lstm'):
        _abi.check(_abi.lib().se_lstm_rec_fwd(pre.data_ptr() + 4 * col0, pre.stride(0), _abi.ptr(w_hh.contiguous()),
                                              N, T, H, _abi.ptr(h), _abi.stream_ptr()))
    return h


def lstm_pair_forward(owner, lstms, x, engine):
    """Several nn.LSTMs (same shape, zero initial state, output sequence only) over the SAME input
    x [N, T, in] (batch-first, contiguous): the layer-0 input projections of all of them are one
    GEMM, each time loop runs in se_lstm_rec_fwd.  engine 'umma' puts the projections on the bf16x3
    tensor-core GEMM, 'torch' leaves them to fp32 F.linear."""
    from ..chain import LinearOperand
    # `x` may be a list of row blocks [n_i, T, in] that are stacked on the batch axis WITHOUT a copy: views with a
    # free row stride (the real / imaginary halves of a [B, T, 2*in] tensor) go through the projection one by one
    parts = list(x) if isinstance(x, (list, tuple)) else [x]
    T = parts[0].shape[1]
    N = sum(p.shape[0] for p in parts)
    dev = parts[0].device
    H, L = lstms[0].hidden_size, lstms[0].num_layers
    cache = owner.__dict__.setdefault('_proj_ops', {})
    outs = []
    with torch.cuda.device(dev):
        if engine == 'umma':
            op = cache.get('l0')
            if op is None:
                op = cache['l0'] = LinearOperand(lambda: [(_lstm_layer_params(m, 0)[0], _lstm_layer_params(m, 0)[2])
                                                          for m in lstms])
            if len(parts) == 1:
                pre = op(parts[0].reshape(N * T, -1))                  # [N*T, len(lstms) * 4H]
            else:
                pre = torch.empty((N * T, len(lstms) * 4 * H), device=dev, dtype=torch.float32)
                r0, jobs = 0, []
                for p in parts:
                    rows = p.shape[0] * T
                    jobs.append((p.reshape(rows, -1), pre[r0:r0 + rows], 0))
                    r0 += rows
                op.many(jobs)
        else:
            x2 = torch.cat(parts, 0).reshape(N * T, -1) if len(parts) > 1 else parts[0].reshape(N * T, -1)
            pre = torch.cat([F.linear(x2, _lstm_layer_params(m, 0)[0],
                                      sum(_lstm_layer_params(m, 0)[2]) if m.bias else None) for m in lstms], 1)
        first = None
        if len(lstms) == 2:
            # both time loops in one launch (grid.y = LSTM)
            first = [torch.empty((N, T, H), device=dev, dtype=torch.float32) for _ in lstms]
            with _abi.stage('lstm'):
                _abi.check(_abi.lib().se_lstm_rec_pair_fwd(
                    _abi.ptr(pre), pre.stride(0), _abi.ptr(_lstm_layer_params(lstms[0], 0)[1].contiguous()),
                    _abi.ptr(_lstm_layer_params(lstms[1], 0)[1].contiguous()), N, T, H, _abi.ptr(first[0]),
                    _abi.ptr(first[1]), _abi.stream_ptr()))
        hs = [first[i] if first is not None else _lstm_rec(pre, i * 4 * H, _lstm_layer_params(m, 0)[1], N, T, H)
              for i, m in enumerate(lstms)]
        for layer in range(1, L):
            if len(lstms) == 2 and engine == 'umma':
                # deeper layers of the two LSTMs (FRCRN: ComplexLSTM(num_layers=2), _2206_07293_frcrn.py:113) read different
                # inputs -- each its own previous layer -- so their projections fill the two column halves of ONE pre-gate
                # matrix and both time loops again share a launch (twice the CTAs per launch at half-empty batches)
                from ..chain import linear_many
                pl = torch.empty((N * T, 2 * 4 * H), device=dev, dtype=torch.float32)
                jobs = []
                for i, m in enumerate(lstms):
                    op = cache.get((i, layer))
                    if op is None:
                        op = cache[(i, layer)] = LinearOperand(
                            lambda m=m, layer=layer: [(_lstm_layer_params(m, layer)[0], _lstm_layer_params(m, layer)[2])])
                    jobs.append((op, hs[i].reshape(N * T, H), pl, i * 4 * H))
                linear_many(jobs)
                nxt = [torch.empty((N, T, H), device=dev, dtype=torch.float32) for _ in lstms]
                with _abi.stage('lstm'):
                    _abi.check(_abi.lib().se_lstm_rec_pair_fwd(
                        _abi.ptr(pl), pl.stride(0), _abi.ptr(_lstm_layer_params(lstms[0], layer)[1].contiguous()),
                        _abi.ptr(_lstm_layer_params(lstms[1], layer)[1].contiguous()), N, T, H, _abi.ptr(nxt[0]),
                        _abi.ptr(nxt[1]), _abi.stream_ptr()))
                hs = nxt
                continue
            for i, m in enumerate(lstms):
                w_ih, w_hh, bias = _lstm_layer_params(m, layer)
                if engine == 'umma':
                    op = cache.get((i, layer))
                    if op is None:
                        op = cache[(i, layer)] = LinearOperand(
                            lambda m=m, layer=layer: [(_lstm_layer_params(m, layer)[0], _lstm_layer_params(m, layer)[2])])
                    pl = op(hs[i].reshape(N * T, H))
                else:
                    pl = F.linear(hs[i].reshape(N * T, H), w_ih, sum(bias) if bias else None)
                hs[i] = _lstm_rec(pl, 0, w_hh, N, T, H)
        outs = hs
    return outs


class ComplexLinear(_abi.InvalidatesCaches, nn.Module):
    """independent real / imaginary linears on the last-dim halves (complex_nn.py:93-113)."""

    def __init__(self, in_channels, out_channels, **kwargs):
        super().__init__()
        ci, co = _halve(in_channels, out_channels)
        self.real_linear = nn.Linear(ci, co, **kwargs)
        self.imag_linear = nn.Linear(ci, co, **kwargs)

    engine = 'umma'         # 'umma': bf16x3 tensor-core GEMM for CUDA fp32 inference; 'torch': nn.Linear

    def forward(self, x):
        if (self.engine == 'umma' and torch.is_tensor(x) and x.is_cuda and x.dtype == torch.float32 and x.dim() >= 2
                and not (torch.is_grad_enabled() and (x.requires_grad or self.real_linear.weight.requires_grad))):
            from ..chain import LinearOperand
            ops = self.__dict__.setdefault('_tc_ops', None)
            if ops is None:
                ops = self.__dict__['_tc_ops'] = tuple(
                    LinearOperand(lambda lin=lin: [(lin.weight, lin.bias)]) for lin in (self.real_linear, self.imag_linear))
            K, No = self.real_linear.in_features, self.real_linear.out_features
            x2 = x.reshape(-1, 2 * K)
            if x2.stride(1) != 1:
                x2 = x2.contiguous()
            out = torch.empty((x2.shape[0], 2 * No), device=x.device, dtype=torch.float32)
            from ..chain import linear_many
            linear_many([(ops[0], x2[:, :K], out, 0), (ops[1], x2[:, K:], out, No)])
            return out.reshape(*x.shape[:-1], 2 * No)
        real, imag = split_complex(x, dim=-1)
        return merge_real_imag(x, self.real_linear(real), self.imag_linear(imag), dim=-1)


class ComplexLSTM(_abi.InvalidatesCaches, nn.Module):
    """complex LSTM as four real LSTM passes (complex_nn.py:115-145)."""

    def __init__(self, in_channels, hidden_channels, **kwargs):
        super().__init__()
        assert in_channels % 2 == 0, f'in_channels must be a factor of 2, current channels: {in_channels}'
        assert hidden_channels % 2 == 0, f'hidden_channels must be a factor of 2, current channels: {hidden_channels}'
        self.real_lstm = nn.LSTM(in_channels // 2, hidden_channels // 2, **kwargs)
        self.imag_lstm = nn.LSTM(in_channels // 2, hidden_channels // 2, **kwargs)

    def forward(self, x):
        real, imag = split_complex(x, dim=-1)
        # both inputs share each LSTM's weights: run them stacked on the batch axis (2 calls, not 4)
        bdim = 0 if self.real_lstm.batch_first else 1
        if (self.rec_kernel and self.proj_engine == 'umma' and self.real_lstm.batch_first and x.dim() == 3
                and x.is_contiguous() and _lstm_rec_ok(self.real_lstm, x)):
            # contiguous batch-first input: the two halves are projected in place (row stride 2*in), no stacked copy
            by_real, by_imag = lstm_pair_forward(self, [self.real_lstm, self.imag_lstm], [real, imag], self.proj_engine)
            r2r, i2r = torch.chunk(by_real, 2, dim=0)
            r2i, i2i = torch.chunk(by_imag, 2, dim=0)
            # [real | imag] written straight into the halves of the output (complex_nn.py:142-145 without the cat copy)
            hid = r2r.shape[-1]
            out = torch.empty(r2r.shape[:-1] + (2 * hid,), device=x.device, dtype=r2r.dtype)
            torch.sub(r2r, i2i, out=out[..., :hid])
            torch.add(r2i, i2r, out=out[..., hid:])
            return out
        both = torch.cat([real, imag], dim=bdim)
        if self.rec_kernel and _lstm_rec_ok(self.real_lstm, both):
            bf = self.real_lstm.batch_first
            xb = (both if bf else both.transpose(0, 1)).contiguous()
            by_real, by_imag = lstm_pair_forward(self, [self.real_lstm, self.imag_lstm], xb, self.proj_engine)
            if not bf:
                by_real, by_imag = by_real.transpose(0, 1), by_imag.transpose(0, 1)
        else:
            # cuDNN RNNs default to TF32 math on CUDA; the parity target is the fp32 reference
            with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
                by_real, _ = self.real_lstm(both)
                by_imag, _ = self.imag_lstm(both)
        r2r, i2r = torch.chunk(by_real, 2, dim=bdim)
        r2i, i2i = torch.chunk(by_imag, 2, dim=bdim)
        return merge_real_imag(x, r2r - i2i, r2i + i2r, dim=-1)

    rec_kernel = True       # class switch: False forces the stock cuDNN path
    proj_engine = 'umma'    # input projections: 'umma' (bf16x3 tensor-core GEMM) or 'torch' (fp32 F.linear)

    def flatten_parameters(self):
        self.real_lstm.flatten_parameters()
        self.imag_lstm.flatten_parameters()


def _cbn_closed_form(S, n, eps, W, Bv):
    """[C, 5] sums -> (Zrr, Zri, Zir, Zii, cr, ci, M, V): the batch statistics and whitening of complex_nn.py:266-322
    on C-sized fp64 tensors (differentiable: the backward of the fused train-mode kernels runs autograd through it)."""
    Mr, Mi = S[:, 0] / n, S[:, 1] / n
    Vrr, Vii, Vri = S[:, 2] / n - Mr * Mr, S[:, 3] / n - Mi * Mi, S[:, 4] / n - Mr * Mi
    vrr, vii = Vrr + eps, Vii + eps
    s_ = (vrr * vii - Vri * Vri).sqrt()
    t_ = (vrr + vii + 2 * s_).sqrt()
    rst = (s_ * t_).reciprocal()
    Urr, Uii, Uri = (s_ + vii) * rst, (s_ + vrr) * rst, -Vri * rst
    if W is None:
        Zrr, Zri, Zir, Zii = Urr, Uri, Uri, Uii
    else:
        Wrr, Wri, Wii = W
        Zrr, Zri, Zir, Zii = Wrr * Urr + Wri * Uri, Wrr * Uri + Wri * Uii, Wri * Urr + Wii * Uri, Wri * Uri + Wii * Uii
    cr, ci = -(Zrr * Mr + Zri * Mi), -(Zir * Mr + Zii * Mi)
    if Bv is not None:
        cr, ci = cr + Bv[0], ci + Bv[1]
    return (Zrr, Zri, Zir, Zii, cr, ci), (Mr, Mi), (Vrr, Vri, Vii)


class _CBNTrainFn(torch.autograd.Function):
    """train-mode complex BN as four HBM passes of csrc/cbn_train.cu around C-sized host algebra (one autograd node
    instead of the ~40 full-tensor ops of complex_nn.py:244-322 and their backward)"""

    @staticmethod
    def forward(ctx, x, eps, Wrr, Wri, Wii, Br, Bi):
        B, C2, H, W = x.shape
        C = C2 // 2
        lib = _abi.lib()
        n = B * H * W
        with torch.cuda.device(x.device), _abi.stage('cbn'):
            splits = lib.se_cbn_splits(B, C, H, W)
            partial = torch.empty((C, splits, 5), device=x.device, dtype=torch.float64)
            _abi.check(lib.se_cbn_stats(_abi.ptr(x), B, C, H, W, _abi.ptr(partial), _abi.stream_ptr()))
            S = partial.sum(1)
            Wd = None if Wrr is None else (Wrr.detach().double(), Wri.detach().double(), Wii.detach().double())
            Bd = None if Br is None else (Br.detach().double(), Bi.detach().double())
            coef, M, V = _cbn_closed_form(S, n, eps, Wd, Bd)
            coef_t = torch.stack(coef).float().contiguous()
            y = torch.empty_like(x)
            _abi.check(lib.se_cbn_apply(_abi.ptr(x), _abi.ptr(coef_t), B, C, H, W, _abi.ptr(y), _abi.stream_ptr()))
        ctx.save_for_backward(x, S, Wrr, Wri, Wii, Br, Bi)
        ctx.eps = eps
        stats = torch.stack([M[0], M[1], V[0], V[1], V[2]]).float()
        ctx.mark_non_differentiable(stats)
        return y, stats

    @staticmethod
    def backward(ctx, g, _g_stats):
        x, S, Wrr, Wri, Wii, Br, Bi = ctx.saved_tensors
        B, C2, H, W = x.shape
        C = C2 // 2
        lib = _abi.lib()
        g = g.contiguous()
        with torch.cuda.device(x.device), _abi.stage('cbn_bwd'):
            splits = lib.se_cbn_splits(B, C, H, W)
            partial = torch.empty((C, splits, 6), device=x.device, dtype=torch.float64)
            _abi.check(lib.se_cbn_bwd_stats(_abi.ptr(x), _abi.ptr(g), B, C, H, W, _abi.ptr(partial), _abi.stream_ptr()))
            G = partial.sum(1)
            with torch.enable_grad():
                S_ = S.detach().requires_grad_(True)
                leaves = [S_]
                Wd = Bd = None
                if Wrr is not None:
                    Wd = tuple(t.detach().double().requires_grad_(True) for t in (Wrr, Wri, Wii))
                    leaves += list(Wd)
                if Br is not None:
                    Bd = tuple(t.detach().double().requires_grad_(True) for t in (Br, Bi))
                    leaves += list(Bd)
                coef, _, _ = _cbn_closed_form(S_, B * H * W, ctx.eps, Wd, Bd)
                # dL/dZrr = sum gr xr, dL/dZri = sum gr xi, dL/dZir = sum gi xr, dL/dZii = sum gi xi, dL/dc = sum g
                grads = torch.autograd.grad(list(coef), leaves, [G[:, 2], G[:, 3], G[:, 4], G[:, 5], G[:, 0], G[:, 1]],
                                            allow_unused=True)
            dS = grads[0]
            k = torch.stack([coef[0].detach(), coef[1].detach(), coef[2].detach(), coef[3].detach(),
                             dS[:, 0], 2 * dS[:, 2], dS[:, 4], dS[:, 1], dS[:, 4], 2 * dS[:, 3]]).float().contiguous()
            dx = torch.empty_like(x)
            _abi.check(lib.se_cbn_bwd_apply(_abi.ptr(x), _abi.ptr(g), _abi.ptr(k), B, C, H, W, _abi.ptr(dx),
                                            _abi.stream_ptr()))
        out = [dx, None]
        i = 1
        if Wrr is not None:
            out += [grads[i].to(Wrr.dtype), grads[i + 1].to(Wri.dtype), grads[i + 2].to(Wii.dtype)]
            i += 3
        else:
            out += [None, None, None]
        if Br is not None:
            out += [grads[i].to(Br.dtype), grads[i + 1].to(Bi.dtype)]
        else:
            out += [None, None]
        return tuple(out)


class ComplexBatchNorm2d(_abi.InvalidatesCaches, nn.Module):
    """2x2 whitening batch norm of Trabelsi et al. (complex_nn.py:148-329)."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, track_running_stats=True,
                 complex_axis=1):
        super().__init__()
        self.num_features = num_features // 2
        self.eps, self.momentum, self.affine = eps, momentum, affine
        self.track_running_stats, self.complex_axis = track_running_stats, complex_axis
        n = self.num_features
        if affine:
            for name in ('Wrr', 'Wri', 'Wii', 'Br', 'Bi'):
                setattr(self, name, nn.Parameter(torch.empty(n)))
        else:
            for name in ('Wrr', 'Wri', 'Wii', 'Br', 'Bi'):
                self.register_parameter(name, None)
        if track_running_stats:
            self.register_buffer('RMr', torch.zeros(n))
            self.register_buffer('RMi', torch.zeros(n))
            self.register_buffer('RVrr', torch.ones(n))
            self.register_buffer('RVri', torch.zeros(n))
            self.register_buffer('RVii', torch.ones(n))
            self.register_buffer('num_batches_tracked', torch.tensor(0, dtype=torch.long))
        else:
            for name in ('RMr', 'RMi', 'RVrr', 'RVri', 'RVii', 'num_batches_tracked'):
                self.register_parameter(name, None)
        self.reset_parameters()

    def reset_running_stats(self):
        if self.track_running_stats:
            self.RMr.zero_(); self.RMi.zero_()
            self.RVrr.fill_(1); self.RVri.zero_(); self.RVii.fill_(1)
            self.num_batches_tracked.zero_()

    def reset_parameters(self):
        self.reset_running_stats()
        if self.affine:
            with torch.no_grad():
                self.Br.zero_(); self.Bi.zero_()
                self.Wrr.fill_(1); self.Wii.fill_(1)
                self.Wri.uniform_(-.9, +.9)

    def _affine_2x2(self, Vrr, Vri, Vii):
        """Z = W * V^{-1/2} per channel (complex_nn.py:283-317)."""
        Vrr, Vii = Vrr + self.eps, Vii + self.eps
        s = (Vrr * Vii - Vri * Vri).sqrt()
        t = (Vrr + Vii + 2 * s).sqrt()
        rst = (s * t).reciprocal()
        Urr, Uii, Uri = (s + Vii) * rst, (s + Vrr) * rst, -Vri * rst
        if not self.affine:
            return Urr, Uri, Uri, Uii
        Wrr, Wri, Wii = self.Wrr, self.Wri, self.Wii
        if Vrr.dim() > 1:
            shape = [1] * Vrr.dim(); shape[1] = -1
            Wrr, Wri, Wii = Wrr.view(shape), Wri.view(shape), Wii.view(shape)
        return (Wrr * Urr + Wri * Uri, Wrr * Uri + Wri * Uii,
                Wri * Urr + Wii * Uri, Wri * Uri + Wii * Uii)

    fused_train = True      # class switch: False keeps the train-mode forward on plain PyTorch ops

    def _forward_train_fused(self, x):
        """batch statistics + whitening as csrc/cbn_train.cu passes (one autograd node), running stats by lerp_"""
        f = 0.0
        if self.training and self.track_running_stats:
            self.num_batches_tracked += 1
            f = 1.0 / self.num_batches_tracked.item() if self.momentum is None else self.momentum
        a = (self.Wrr, self.Wri, self.Wii, self.Br, self.Bi) if self.affine else (None,) * 5
        y, stats = _CBNTrainFn.apply(x.contiguous(), float(self.eps), *a)
        if self.track_running_stats:
            with torch.no_grad():
                self.RMr.lerp_(stats[0], f); self.RMi.lerp_(stats[1], f)
                self.RVrr.lerp_(stats[2], f); self.RVri.lerp_(stats[3], f); self.RVii.lerp_(stats[4], f)
        return y

    def forward(self, inputs):
        training = self.training or not self.track_running_stats
        if (training and self.fused_train and inputs.is_cuda and inputs.dtype == torch.float32 and inputs.dim() == 4
                and self.complex_axis == 1 and inputs.shape[0] * inputs.shape[2] * inputs.shape[3] > 0):
            return self._forward_train_fused(inputs)
        xr, xi = torch.chunk(inputs, 2, dim=self.complex_axis)
        vdim = [1] * xr.dim(); vdim[1] = xr.size(1)
        if training:
            f = 0.0
            if self.training and self.track_running_stats:
                self.num_batches_tracked += 1
                f = 1.0 / self.num_batches_tracked.item() if self.momentum is None else self.momentum
            red = [d for d in range(xr.dim()) if d != 1]
            Mr, Mi = xr.mean(red, keepdim=True), xi.mean(red, keepdim=True)
            xr, xi = xr - Mr, xi - Mi
            Vrr = (xr * xr).mean(red, keepdim=True)
            Vri = (xr * xi).mean(red, keepdim=True)
            Vii = (xi * xi).mean(red, keepdim=True)
            if self.track_running_stats:
                with torch.no_grad():
                    self.RMr.lerp_(Mr.reshape(-1), f); self.RMi.lerp_(Mi.reshape(-1), f)
                    self.RVrr.lerp_(Vrr.reshape(-1), f); self.RVri.lerp_(Vri.reshape(-1), f)
                    self.RVii.lerp_(Vii.reshape(-1), f)
        else:
            xr, xi = xr - self.RMr.view(vdim), xi - self.RMi.view(vdim)
            Vrr, Vri, Vii = self.RVrr.view(vdim), self.RVri.view(vdim), self.RVii.view(vdim)
        Zrr, Zri, Zir, Zii = self._affine_2x2(Vrr, Vri, Vii)
        yr, yi = Zrr * xr + Zri * xi, Zir * xr + Zii * xi
        if self.affine:
            yr, yi = yr + self.Br.view(vdim), yi + self.Bi.view(vdim)
        return torch.cat([yr, yi], self.complex_axis)

    def extra_repr(self):
        return (f'{self.num_features}, eps={self.eps}, momentum={self.momentum}, affine={self.affine}, '
                f'track_running_stats={self.track_running_stats}')


class ComplexPReLU(nn.Module):
    """PReLU applied as a 'complex product' of two real PReLUs (complex_nn.py:337-357)."""

    def __init__(self, **kwargs):
        super().__init__()
        self.real_prelu = nn.PReLU(**kwargs)
        self.imag_prelu = nn.PReLU(**kwargs)

    def forward(self, x):
        real, imag = split_complex(x, dim=1)
        out_r = self.real_prelu(real) - self.imag_prelu(imag)
        out_i = self.imag_prelu(real) + self.real_prelu(imag)
        return merge_real_imag(x, out_r, out_i, dim=1)


# the reference aliases these to the plain real activations (complex_nn.py:359-360)
ComplexReLU = nn.ReLU
ComplexLeakyReLU = nn.LeakyReLU
