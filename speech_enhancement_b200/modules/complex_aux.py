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


def lstm_forward(lstm, x):
    """nn.LSTM forward (zero initial state, output sequence only) with the input projection as one GEMM
    and the time loop in se_lstm_rec_fwd.  x: [N, T, in] if batch_first else [T, N, in]."""
    if not lstm.batch_first:
        x = x.transpose(0, 1)
    x = x.contiguous()
    N, T, _ = x.shape
    H = lstm.hidden_size
    with torch.cuda.device(x.device):
        for layer in range(lstm.num_layers):
            w_ih, w_hh = getattr(lstm, f'weight_ih_l{layer}'), getattr(lstm, f'weight_hh_l{layer}')
            bias = getattr(lstm, f'bias_ih_l{layer}') + getattr(lstm, f'bias_hh_l{layer}') if lstm.bias else None
            pre = F.linear(x, w_ih, bias)                       # [N, T, 4H], fp32 (no TF32: torch default)
            h = torch.empty((N, T, H), device=x.device, dtype=torch.float32)
            with _abi.stage('lstm'):
                _abi.check(_abi.lib().se_lstm_rec_fwd(_abi.ptr(pre), _abi.ptr(w_hh.contiguous()), N, T, H,
                                                      _abi.ptr(h), _abi.stream_ptr()))
            x = h
    return x if lstm.batch_first else x.transpose(0, 1)


class ComplexLinear(nn.Module):
    """independent real / imaginary linears on the last-dim halves (complex_nn.py:93-113)."""

    def __init__(self, in_channels, out_channels, **kwargs):
        super().__init__()
        ci, co = _halve(in_channels, out_channels)
        self.real_linear = nn.Linear(ci, co, **kwargs)
        self.imag_linear = nn.Linear(ci, co, **kwargs)

    def forward(self, x):
        real, imag = split_complex(x, dim=-1)
        return merge_real_imag(x, self.real_linear(real), self.imag_linear(imag), dim=-1)


class ComplexLSTM(nn.Module):
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
        both = torch.cat([real, imag], dim=bdim)
        if self.rec_kernel and _lstm_rec_ok(self.real_lstm, both):
            by_real, by_imag = lstm_forward(self.real_lstm, both), lstm_forward(self.imag_lstm, both)
        else:
            # cuDNN RNNs default to TF32 math on CUDA; the parity target is the fp32 reference
            with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
                by_real, _ = self.real_lstm(both)
                by_imag, _ = self.imag_lstm(both)
        r2r, i2r = torch.chunk(by_real, 2, dim=bdim)
        r2i, i2i = torch.chunk(by_imag, 2, dim=bdim)
        return merge_real_imag(x, r2r - i2i, r2i + i2r, dim=-1)

    rec_kernel = True       # class switch: False forces the stock cuDNN path

    def flatten_parameters(self):
        self.real_lstm.flatten_parameters()
        self.imag_lstm.flatten_parameters()


class ComplexBatchNorm2d(nn.Module):
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

    def forward(self, inputs):
        xr, xi = torch.chunk(inputs, 2, dim=self.complex_axis)
        vdim = [1] * xr.dim(); vdim[1] = xr.size(1)
        training = self.training or not self.track_running_stats
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
