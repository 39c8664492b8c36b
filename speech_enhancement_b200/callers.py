"""Caller models of the hot path: DCCRN, DCUNet, FRCRN hosts written for this package.

The reference's model files (/root/reference/models/_2008_00264_dccrn.py, _1903_03107_dcunet.py,
_2206_07293_frcrn.py) are the callers of the hot path; they can be used unchanged through
`dropin.install()`.  These hosts exist so that bench / smoke / tests run on a box without the
reference checkout.  They keep the reference's constructor arguments, forward contract
`model(wav[B,L]) -> (spec[B,2F,T], wav[B,L'])` and state_dict keys (so reference checkpoints
load), and differ in one respect: the inline mask math + torch.cat is one fused kernel call.

`ops` selects the implementation of the hot-path pieces (ConvSTFT, ConviSTFT, ComplexConv2d,
ComplexConvTranspose2d, mask functions).  The default is the CUDA implementation of this
package; the CPU oracle passes its own namespace to obtain the timed CPU baseline.
"""
import types

import torch
import torch.nn as nn
import torch.nn.functional as F

from .modules.complex_nn import complex_concat, split_complex, merge_real_imag
from .modules.complex_aux import ComplexBatchNorm2d, ComplexLSTM, ComplexLinear


def cuda_ops():
    from . import conv_stft, functional
    from .modules import complex_nn
    return types.SimpleNamespace(
        ConvSTFT=conv_stft.ConvSTFT, ConviSTFT=conv_stft.ConviSTFT,
        ComplexConv2d=complex_nn.ComplexConv2d, ComplexConvTranspose2d=complex_nn.ComplexConvTranspose2d,
        mask_dccrn=functional.mask_dccrn, mask_dcunet=functional.mask_dcunet,
        mask_frcrn=functional.mask_frcrn, mask_carn=functional.mask_carn)


class _Block(nn.Module):
    """(causal time pad) -> complex conv / convT -> complex BN -> activation.

    dccrn.py:11-57, frcrn.py:12-59 (causal=True pads `time_pad` frames on the left before the
    layer, :29 / :53), dcunet.py:13-43 (no explicit pad)."""

    def __init__(self, ops, transposed, cin, cout, kernel, act, norm=True, time_pad=0, **kw):
        super().__init__()
        layer = (ops.ComplexConvTranspose2d if transposed else ops.ComplexConv2d)(
            cin, cout, kernel, bias=not norm, **kw)
        setattr(self, 'conv_transposed' if transposed else 'conv', layer)
        self._layer_name = 'conv_transposed' if transposed else 'conv'
        self.norm = ComplexBatchNorm2d(cout) if norm else nn.Identity()
        self.act = act if act is not None else nn.Identity()
        self.time_pad = time_pad

    def forward(self, x):
        if self.time_pad:
            x = F.pad(x, (self.time_pad, 0, 0, 0))
        return self.act(self.norm(getattr(self, self._layer_name)(x)))


class _Stack(nn.Module):
    def __init__(self, layers):
        super().__init__()
        self.layers = nn.ModuleList(layers)


def _fusable(model, x, first_block):
    """the chained engine serves eval / no_grad CUDA calls of models built on this package's conv modules"""
    name = first_block._layer_name
    return (model.fused and not model.training and not torch.is_grad_enabled() and x.is_cuda
            and type(getattr(first_block, name)).__module__.startswith('speech_enhancement_b200'))


def _clamped(wav):
    """the models' final clamp_(-1, 1) (_2008_00264_dccrn.py:184); already done by the iSTFT epilogue when it wrote
    PCM16 samples (ConviSTFT.pcm16_out)"""
    return wav if wav.dtype == torch.int16 else torch.clamp_(wav, -1, 1)


def _spec_planes(spec, n_bins):
    return spec[:, :n_bins, :], spec[:, n_bins:, :]


# ------------------------------------------------------------------------------------ DCCRN
class _LSTMBlock(nn.Module):
    """dccrn.py:59-86: (num_layers) ComplexLSTM + ComplexLinear."""

    def __init__(self, cin, hidden, linear, num_layers=2, **kw):
        super().__init__()
        ndir = 2 if kw.get('bidirectional', True) else 1
        mods = [ComplexLSTM(cin, hidden, num_layers=1, **kw) for _ in range(num_layers - 1)]
        mods += [ComplexLSTM(ndir * hidden, hidden, num_layers=1, **kw), ComplexLinear(ndir * hidden, linear)]
        self.layers = nn.ModuleList(mods)

    def forward(self, x):
        for layer in self.layers:
            x = layer(x)
        return x


class DCCRN(nn.Module):
    """DCCRN-R/C/E/CL (dccrn.py:123-212)."""

    def __init__(self, config='dccrn-CL', window_size=400, hop_size=100, fft_size=512, lstm_channels=256,
                 linear_channels=1024, bidirectional=False, is_complex=True, ops=None):
        super().__init__()
        if not is_complex:
            raise RuntimeError('DCCRN host: only the complex variant is on the accelerated path')
        ops = ops or cuda_ops()
        self._ops = ops
        if config in ('dccrn-C', 'dccrn-R', 'dccrn-E'):
            enc, self.masking = [32, 64, 128, 128, 256, 256], config[-1]
        else:
            enc, self.masking = [32, 64, 128, 256, 256, 256], 'E'
        dec = enc[:-1][::-1] + [2]
        self.fft_size = fft_size
        self.stft = ops.ConvSTFT(window_size, hop_size, fft_size)
        self.istft = ops.ConviSTFT(window_size, hop_size, fft_size)
        chans = [2] + enc
        self.encoder = _Stack([_Block(ops, False, chans[i], chans[i + 1], (5, 2), nn.PReLU(), time_pad=1,
                                      stride=(2, 1), padding=(2, 0)) for i in range(len(enc))])
        cin, layers = 256, []
        for cout in dec:
            layers.append(_Block(ops, True, cin * 2, cout, (5, 2), nn.PReLU(), time_pad=0,
                                 stride=(2, 1), padding=(2, 0), output_padding=(1, 0)))
            cin = cout
        self.decoder = _Stack(layers)
        freq = fft_size // 2 // (2 ** len(enc)) * enc[-1]
        self.lstm = _LSTMBlock(freq, lstm_channels, linear_channels, num_layers=2, batch_first=True,
                               bidirectional=bidirectional)
        self._fb = None

    # eval-mode / no-grad forwards on CUDA run the chained engine (chain.py): BN, activation, causal
    # pad and skip concat are folded into the conv kernels and activations stay in CL2 between layers
    fused = True

    def _use_fused(self, x):
        return (self.fused and not self.training and not torch.is_grad_enabled() and x.is_cuda
                and hasattr(self.encoder.layers[0], 'conv')
                and type(self.encoder.layers[0].conv).__module__.startswith('speech_enhancement_b200'))

    def _fused_blocks(self):
        if self._fb is None:
            from .chain import FusedBlock
            enc = [FusedBlock(l.conv, l.norm, l.act) for l in self.encoder.layers]
            dec = []
            for l in self.decoder.layers:
                cin2 = 2 * l.conv_transposed.real_conv.weight.shape[0]
                dec.append(FusedBlock(l.conv_transposed, l.norm, l.act, src_channels=[cin2 // 2, cin2 // 2],
                                      concat='complex'))
            self._fb = (enc, dec)
        return self._fb

    def _forward_fused(self, x):
        from .chain import cl2_from_lstm_rows
        n_bins = self.fft_size // 2 + 1
        spec = self.stft(x)                                               # [B, 2F, T] fp32
        B, C2, T = spec.shape
        noisy_real, noisy_imag = _spec_planes(spec, n_bins)
        enc, dec = self._fused_blocks()
        # enc0 reads the real / imag row blocks of `spec` in place (DC bin skipped, causal frame pad)
        act, _ = enc[0].first_layer(spec[:, 1:], C2 * T, n_bins * T, T, B, n_bins - 1, T, pad_left=1)
        skips = [act]
        rows = None
        for i in range(1, len(enc)):
            if i == len(enc) - 1:
                # the last encoder layer also writes its output as the [B, T, C*F] rows the LSTM block reads
                # (dccrn.py:161-163: column c*F + f), next to the CL2 skip
                c = enc[i].cout2
                f, t = enc[i].out_size(act.H, act.W, 0, 1)
                rows = torch.empty((B, t, c * f), device=spec.device, dtype=torch.float32)
                act, _ = enc[i]([act], pad_left=1, out_cl2=True, y_out=rows, y_strides=(t * c * f, f, 1, c * f))
            else:
                act, _ = enc[i]([act], pad_left=1, out_cl2=True)
            skips.append(act)
        act = cl2_from_lstm_rows(self.lstm(rows), B, c, f, t)
        y = None
        for i, blk in enumerate(dec):
            skip = skips.pop()
            if act.W > skip.W:
                act = act.trimmed(W=skip.W)
            last = i == len(dec) - 1
            act, y = blk([act, skip], out_cl2=not last, out_nchw=last)
        y = F.pad(y, (0, 0, 1, 0))                                        # zero DC row (:169)
        mask_real, mask_imag = y[:, 0], y[:, 1]
        if mask_real.shape[-1] > noisy_real.shape[-1]:
            mask_real, mask_imag = mask_real[..., :-1], mask_imag[..., :-1]
        est_spec = self._ops.mask_dccrn(self.masking, noisy_real, noisy_imag, mask_real, mask_imag)
        wav = self.istft(est_spec)
        return est_spec, _clamped(wav)

    def forward(self, x):
        if self._use_fused(x):
            return self._forward_fused(x)
        n_bins = self.fft_size // 2 + 1
        spec = self.stft(x)                                               # [B, 2F, T]
        noisy_real, noisy_imag = _spec_planes(spec, n_bins)
        x = torch.stack([noisy_real, noisy_imag], dim=1)[:, :, 1:, :]      # drop DC (:155-156)
        skips = []
        for layer in self.encoder.layers:
            x = layer(x)
            skips.append(x)
        b, c, f, t = x.shape
        x = self.lstm(x.reshape(b, c * f, t).permute(0, 2, 1))
        x = x.permute(0, 2, 1).reshape(b, c, f, t)
        for layer in self.decoder.layers:
            skip = skips.pop()
            if x.shape[-1] > skip.shape[-1]:
                x = x[..., :-1]
            x = layer(complex_concat([x, skip], dim=1))
        x = F.pad(x, (0, 0, 1, 0))                                        # zero DC row (:169)
        mask_real, mask_imag = x[:, 0], x[:, 1]
        if mask_real.shape[-1] > noisy_real.shape[-1]:
            mask_real, mask_imag = mask_real[..., :-1], mask_imag[..., :-1]
        est_spec = self._ops.mask_dccrn(self.masking, noisy_real, noisy_imag, mask_real, mask_imag)
        wav = self.istft(est_spec)
        return est_spec, _clamped(wav)


# ------------------------------------------------------------------------------------ DCUNet
# (complex channels, kernel, stride, padding) per encoder layer, architectures.py:53-99
_K75, _K53 = (7, 5), (5, 3)
DCUNET_LAYERS = {
    'dcunet10': [(32, _K75, (2, 2), (3, 2)), (64, _K75, (2, 2), (3, 2)), (64, _K53, (2, 2), (2, 1)),
                 (64, _K53, (2, 2), (2, 1)), (64, _K53, (2, 1), (2, 1))],
    'dcunet16': [(32, _K75, (2, 2), (3, 2)), (32, _K75, (2, 1), (3, 2)), (64, _K75, (2, 2), (3, 2))]
                + [(64, _K53, s, (2, 1)) for s in ((2, 1), (2, 2), (2, 1), (2, 2), (2, 1))],
    'dcunet20': [(32, (7, 1), (1, 1), (3, 0)), (32, (1, 7), (1, 1), (0, 3)), (64, _K75, (2, 2), (3, 2)),
                 (64, _K75, (2, 1), (3, 2))]
                + [(64, _K53, s, (2, 1)) for s in ((2, 2), (2, 1), (2, 2), (2, 1), (2, 2))]
                + [(90, _K53, (2, 1), (2, 1))],
    'dcunet20-large': [(45, (7, 1), (1, 1), (3, 0)), (45, (1, 7), (1, 1), (0, 3)), (90, _K75, (2, 2), (3, 2)),
                       (90, _K75, (2, 1), (3, 2))]
                      + [(90, _K53, s, (2, 1)) for s in ((2, 2), (2, 1), (2, 2), (2, 1), (2, 2))]
                      + [(128, _K53, (2, 1), (2, 1))],
}


class DCUNet(nn.Module):
    """DCUNet-10/16/20/20-large (dcunet.py:98-189)."""

    def __init__(self, config, window_size=512, hop_size=128, fft_size=512, normalize=False, is_complex=True,
                 ops=None):
        super().__init__()
        if not is_complex:
            raise RuntimeError('DCUNet host: only the complex variant is on the accelerated path')
        ops = ops or cuda_ops()
        self._ops = ops
        arch = DCUNET_LAYERS[config]
        self.window_size, self.hop_size, self.fft_size, self.normalize = window_size, hop_size, fft_size, normalize
        self.stft = ops.ConvSTFT(window_size, hop_size, fft_size)
        self.istft = ops.ConviSTFT(window_size, hop_size, fft_size)
        lrelu = lambda: nn.LeakyReLU(0.01)
        c0 = arch[0][0] * 2
        self.first_conv = _Block(ops, False, 2, c0, 3, lrelu(), padding=1)
        enc, cin = [], c0
        for ch, k, s, p in arch:
            enc.append(_Block(ops, False, cin, ch * 2, k, lrelu(), stride=s, padding=p))
            cin = ch * 2
        self.encoder = _Stack(enc)
        dec = []
        for i in range(len(arch) - 1):                                    # dcunet.py:72-78
            _, k, s, p = arch[-i - 1]
            cout = arch[-i - 2][0] * 2
            dec.append(_Block(ops, True, cin * 2, cout, k, lrelu(), stride=s, padding=p))
            cin = cout
        _, k, s, p = arch[0]
        dec.append(_Block(ops, True, cin * 2, 2, k, None, stride=s, padding=p))
        self.decoder = _Stack(dec)

    fused = True
    _fb = None

    def _fused_blocks(self):
        if self._fb is None:
            from .chain import FusedBlock
            first = FusedBlock(self.first_conv.conv, self.first_conv.norm, self.first_conv.act)
            enc = [FusedBlock(l.conv, l.norm, l.act) for l in self.encoder.layers]
            dec = []
            for l in self.decoder.layers:
                cin2 = 2 * l.conv_transposed.real_conv.weight.shape[0]
                dec.append(FusedBlock(l.conv_transposed, l.norm, l.act, src_channels=[cin2 // 2, cin2 // 2],
                                      concat='plain'))
            self._fb = (first, enc, dec)
        return self._fb

    def _forward_fused(self, x):
        """same network with every layer chained in the engine's CL2 format: BN + LeakyReLU folded, the
        F.pad to the skip's size (dcunet.py:89-92) done by out-of-bounds fill, torch.cat as a second source."""
        n_bins = self.fft_size // 2 + 1
        spec = self.stft(x)                                               # [B, 2F, T]
        B, C2, T = spec.shape
        if self.normalize:
            # dcunet.py:127-130 on the [B, 2F, T] view (same element set as [B, 2, F, T] over dims 1-3); the mask is
            # applied to the NORMALISED spectrum (`identity`, :134), so `spec` is simply replaced
            mean, std = spec.mean(dim=(1, 2), keepdim=True), spec.std(dim=(1, 2), keepdim=True)
            spec = (spec - mean) / (std + 1e-8)
        first, enc, dec = self._fused_blocks()
        act, _ = first.first_layer(spec, C2 * T, n_bins * T, T, B, n_bins, T)
        skips = []
        for blk in enc:
            act, _ = blk([act])
            skips.append(act)
        y = None
        for i, blk in enumerate(dec):
            skip = skips.pop()
            if act.H > skip.H or act.W > skip.W:
                raise RuntimeError('DCUNet: decoder activation larger than its skip connection')
            last = i == len(dec) - 1
            act, y = blk([act, skip], out_cl2=not last, out_nchw=last)
        y = y[:, :, :n_bins, :T]                                          # dcunet.py:141-146
        real, imag = _spec_planes(spec, n_bins)
        est_spec = self._ops.mask_dcunet('bounded_tanh', real, imag, y[:, 0], y[:, 1])
        wav = self.istft(est_spec)
        return est_spec, _clamped(wav)

    def forward(self, x):
        if _fusable(self, x, self.first_conv):
            return self._forward_fused(x)
        n_bins = self.fft_size // 2 + 1
        spec = self.stft(x)
        real, imag = _spec_planes(spec, n_bins)
        cs = torch.stack([real, imag], dim=1)
        if self.normalize:                                                # dcunet.py:127-130
            mean = torch.mean(cs, dim=[1, 2, 3], keepdim=True)
            std = torch.std(cs, dim=[1, 2, 3], keepdim=True)
            cs = (cs - mean) / (std + 1e-8)
        x = self.first_conv(cs)
        skips = []
        for layer in self.encoder.layers:
            x = layer(x)
            skips.append(x)
        for layer in self.decoder.layers:
            skip = skips.pop()
            if skip.shape != x.shape:                                     # dcunet.py:89-92
                x = F.pad(x, (0, abs(skip.shape[3] - x.shape[3]), 0, abs(skip.shape[2] - x.shape[2])))
            x = layer(torch.cat([x, skip], dim=1))
        x = x[:, :, :cs.shape[2], :cs.shape[3]]                           # dcunet.py:141-146
        est_spec = self._ops.mask_dcunet('bounded_tanh', cs[:, 0], cs[:, 1], x[:, 0], x[:, 1])
        wav = self.istft(est_spec)
        return est_spec, _clamped(wav)


# ------------------------------------------------------------------------------------ FRCRN
class _LinearBlock(nn.Module):
    def __init__(self, cin, cout, act):
        super().__init__()
        self.linear = ComplexLinear(cin, cout, bias=False)
        self.act = nn.ReLU() if act else nn.Identity()

    def forward(self, x):
        return self.act(self.linear(x))


class _ChannelAttention(nn.Module):
    """ccbam.py:28-63: avg/max pooled descriptors -> shared complex MLP -> sigmoid."""

    def __init__(self, channels, r=16):
        super().__init__()
        red = channels // r if channels // r else 2
        self.shared_fc_layer = nn.Sequential(_LinearBlock(channels, red, True), _LinearBlock(red, channels, False))

    def forward(self, x):
        b, c = x.shape[:2]
        avg = self.shared_fc_layer(x.mean(dim=(2, 3)))     # [real | imag] order == channel order
        mx = self.shared_fc_layer(x.amax(dim=(2, 3)))
        return torch.sigmoid(avg + mx).view(b, c, 1, 1)


class _SpatialAttention(nn.Module):
    """ccbam.py:65-86: channel-pooled maps -> 7x7 complex conv + CBN + ReLU -> sigmoid."""

    def __init__(self, ops):
        super().__init__()
        self.conv = _Block(ops, False, 4, 2, 7, nn.ReLU(), padding=3)

    def forward(self, x):
        real, imag = split_complex(x)
        pooled = torch.cat([real.mean(1, keepdim=True), real.amax(1, keepdim=True),
                            imag.mean(1, keepdim=True), imag.amax(1, keepdim=True)], dim=1)
        return torch.sigmoid(self.conv(pooled))


class CCBAM(nn.Module):
    """complex convolutional block attention (ccbam.py:88-106)."""

    def __init__(self, feature_map_channels, reduction=16, ops=None):
        super().__init__()
        ops = ops or cuda_ops()
        self.channel_attention_branch = _ChannelAttention(feature_map_channels, reduction)
        self.spatial_attention_branch = _SpatialAttention(ops)

    def _attn_operands(self, dev):
        """dense matrices of the shared MLP (the reference's ComplexLinear is block diagonal over the real / imaginary
        halves, complex_nn.py:106-113) and the 7x7 attention conv with its eval-mode complex BN folded in; rebuilt when
        a parameter or running statistic changes"""
        from .chain import FusedBlock
        fc = self.channel_attention_branch.shared_fc_layer
        conv = self.spatial_attention_branch.conv
        blk = self.__dict__.get('_sa_block')
        if blk is None:
            blk = self.__dict__['_sa_block'] = FusedBlock(conv.conv, conv.norm, conv.act)
        ops = blk._operands()
        lins = [fc[0].linear.real_linear, fc[0].linear.imag_linear, fc[1].linear.real_linear, fc[1].linear.imag_linear]
        from . import _abi
        key = _abi.cache_key([l.weight for l in lins], id(ops['packed']))
        c = self.__dict__.get('_attn_ops')
        if c is None or c[0] != key:
            with torch.no_grad():
                w1 = torch.block_diag(lins[0].weight.float(), lins[1].weight.float()).contiguous().to(dev)
                w2 = torch.block_diag(lins[2].weight.float(), lins[3].weight.float()).contiguous().to(dev)
            c = self.__dict__['_attn_ops'] = (key, w1, w2)
        if ops['slope'] != 0.0 or ops['packed'].shape != (2, 4 * 49):
            raise RuntimeError('CCBAM.forward_cl2: the spatial attention block must be a 7x7 complex conv 4 -> 2 + ReLU')
        return c[1], c[2], ops['packed'], ops['bias']

    def forward_cl2(self, act):
        """the same attention on a CL2 activation (chain.Act): five launches of csrc/ccbam.cu -- channel statistics,
        channel attention (partial reduce + shared MLP + sigmoid), spatial pooling of x * ca, spatial attention
        (7x7 conv + folded BN + ReLU + sigmoid), apply."""
        from . import _abi
        from .chain import Act
        x = act.data
        B, _, H, W, Cp = x.shape
        if (H, W) != (act.H, act.W):
            raise RuntimeError('CCBAM.forward_cl2: trimmed activations are not supported')
        if self.training:
            raise RuntimeError('CCBAM.forward_cl2: eval mode only (the complex BN of the attention conv is folded)')
        C, P = act.C, H * W
        lib = _abi.lib()
        w1, w2, packed, bias = self._attn_operands(x.device)
        if w1.shape[1] != C or w2.shape[0] != C:
            raise RuntimeError(f'CCBAM.forward_cl2: attention MLP is built for {w1.shape[1]} channels, got {C}')
        with torch.cuda.device(x.device), _abi.stage('ccbam'):
            nb = lib.se_ccbam_num_partials(P)
            partial = torch.empty((B, nb, 2, Cp), device=x.device, dtype=torch.float32)
            _abi.check(lib.se_ccbam_channel_stats(_abi.ptr(x), B, P, C, Cp, _abi.ptr(partial), _abi.stream_ptr()))
            ca = torch.empty((B, C), device=x.device, dtype=torch.float32)
            _abi.check(lib.se_ccbam_channel_attn(_abi.ptr(partial), nb, P, _abi.ptr(w1), _abi.ptr(w2), B, C, Cp,
                                                 w1.shape[0], _abi.ptr(ca), _abi.stream_ptr()))
            pooled = torch.empty((B, 4, H, W), device=x.device, dtype=torch.float32)
            _abi.check(lib.se_ccbam_spatial_pool(_abi.ptr(x), _abi.ptr(ca), B, P, C, Cp, _abi.ptr(pooled),
                                                 _abi.stream_ptr()))
            sa = torch.empty((B, 2, H, W), device=x.device, dtype=torch.float32)
            _abi.check(lib.se_ccbam_spatial_attn(_abi.ptr(pooled), _abi.ptr(packed), _abi.ptr(bias), B, H, W,
                                                 _abi.ptr(sa), _abi.stream_ptr()))
            out = torch.empty_like(x)
            _abi.check(lib.se_ccbam_apply(_abi.ptr(x), _abi.ptr(ca), _abi.ptr(sa), B, P, C, Cp, _abi.ptr(out),
                                          _abi.stream_ptr()))
        return Act(out, C, H, W)

    def forward(self, x):
        x = x * self.channel_attention_branch(x)
        sa = self.spatial_attention_branch(x)
        real, imag = split_complex(x)
        return merge_real_imag(x, real + sa[:, 0:1], imag + sa[:, 1:2], dim=1)


class _FrcrnDecoder(nn.Module):
    def __init__(self, ops, channels=128, repeats=6, reduction=16):
        super().__init__()
        self.skip_connection_attention_layers = nn.ModuleList(
            [CCBAM(channels, reduction, ops=ops) for _ in range(repeats)])
        self.layers = nn.ModuleList(
            [_Block(ops, True, channels * 2, channels, (5, 2), nn.LeakyReLU(0.2), time_pad=0,
                    stride=(2, 1), padding=(0, 0)) for _ in range(repeats)])


class FRCRN(nn.Module):
    """FRCRN without FSMN (frcrn.py:104-155)."""

    def __init__(self, window_size=320, hop_size=160, fft_size=640, lstm_channels=256, reduction_ratio=16,
                 is_complex=True, ops=None):
        super().__init__()
        if not is_complex:
            raise RuntimeError('FRCRN host: only the complex variant is on the accelerated path')
        ops = ops or cuda_ops()
        self._ops = ops
        self.fft_size = fft_size
        self.stft = ops.ConvSTFT(window_size, hop_size, fft_size)
        self.istft = ops.ConviSTFT(window_size, hop_size, fft_size)
        chans = [2] + [128] * 6
        self.encoder = _Stack([_Block(ops, False, chans[i], chans[i + 1], (5, 2), nn.LeakyReLU(0.2), time_pad=1,
                                      stride=(2, 1), padding=(0, 0)) for i in range(6)])
        self.decoder = _FrcrnDecoder(ops, 128, 6, reduction_ratio)
        self.lstm = ComplexLSTM(256, lstm_channels, num_layers=2, bidirectional=False, batch_first=True)
        self.final_conv = nn.Conv2d(128, 2, kernel_size=(1, 2), bias=False)

    fused = True
    _fb = None
    _final_fb = None

    def _fused_blocks(self):
        if self._fb is None:
            from .chain import FusedBlock
            enc = [FusedBlock(l.conv, l.norm, l.act) for l in self.encoder.layers]
            dec = []
            for l in self.decoder.layers:
                cin2 = 2 * l.conv_transposed.real_conv.weight.shape[0]
                dec.append(FusedBlock(l.conv_transposed, l.norm, l.act, src_channels=[cin2 // 2, cin2 // 2],
                                      concat='complex'))
            self._fb = (enc, dec)
        return self._fb

    def _final_block(self):
        if self._final_fb is None:
            from .chain import RealConvBlock
            self._final_fb = RealConvBlock(self.final_conv)
        return self._final_fb

    def _forward_fused(self, x):
        """encoder / decoder convolutions chained in CL2 (BN + LeakyReLU folded, causal pad as a coordinate
        shift, complex_concat as a second source, the bottom-row F.pad by out-of-bounds fill), the CCBAM
        attention applied to the CL2 skips by csrc/ccbam.cu; the real final_conv reads the last decoder's CL2
        output as a RealConvBlock."""
        from .chain import cl2_from_lstm_rows
        n_bins = self.fft_size // 2 + 1
        spec = self.stft(x)
        B, C2, T = spec.shape
        noisy_real, noisy_imag = _spec_planes(spec, n_bins)
        enc, dec = self._fused_blocks()
        act, _ = enc[0].first_layer(spec[:, 1:], C2 * T, n_bins * T, T, B, n_bins - 1, T, pad_left=1)
        skips = [act]
        rows = None
        for i in range(1, len(enc)):
            if i == len(enc) - 1:
                # the last encoder layer also writes the [B, T, C*F] rows the LSTM reads (frcrn.py: column c*F + f)
                c = enc[i].cout2
                f, t = enc[i].out_size(act.H, act.W, 0, 1)
                rows = torch.empty((B, t, c * f), device=spec.device, dtype=torch.float32)
                act, _ = enc[i]([act], pad_left=1, out_cl2=True, y_out=rows, y_strides=(t * c * f, f, 1, c * f))
            else:
                act, _ = enc[i]([act], pad_left=1, out_cl2=True)
            skips.append(act)
        act = cl2_from_lstm_rows(self.lstm(rows), B, c, f, t)
        y = None
        n = len(dec)
        for i, (att, blk) in enumerate(zip(self.decoder.skip_connection_attention_layers, dec)):
            skip = att.forward_cl2(skips.pop())
            if act.W > skip.W:
                act = act.trimmed(W=skip.W)
            act, _ = blk([act, skip], out_cl2=True, out_nchw=False)
        _, y = self._final_block()([act], out_cl2=False, out_nchw=True)   # [B, 2, F-2, T]
        est_spec = self._ops.mask_frcrn(noisy_real, noisy_imag, y[:, 0], y[:, 1])
        wav = self.istft(est_spec)
        return est_spec, _clamped(wav)

    def forward(self, x):
        if _fusable(self, x, self.encoder.layers[0]):
            return self._forward_fused(x)
        n_bins = self.fft_size // 2 + 1
        spec = self.stft(x)
        noisy_real, noisy_imag = _spec_planes(spec, n_bins)
        x = torch.stack([noisy_real, noisy_imag], dim=1)[:, :, 1:, :]
        skips = []
        for layer in self.encoder.layers:
            x = layer(x)
            skips.append(x)
        b, c, f, t = x.shape
        x = self.lstm(x.reshape(b, c * f, t).permute(0, 2, 1))
        x = x.permute(0, 2, 1).reshape(b, c, f, t)
        for att, layer in zip(self.decoder.skip_connection_attention_layers, self.decoder.layers):
            skip = att(skips.pop())
            if x.shape[-1] > skip.shape[-1]:
                x = x[..., :-1]
            if x.shape[-2] < skip.shape[-2]:
                x = F.pad(x, (0, 0, 0, 1))
            x = layer(complex_concat([x, skip], dim=1))
        # cuDNN convolutions default to TF32 on CUDA; the parity target is the fp32 reference
        with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
            x = self.final_conv(x)                                        # [B, 2, F-2, T]
        est_spec = self._ops.mask_frcrn(noisy_real, noisy_imag, x[:, 0], x[:, 1])
        wav = self.istft(est_spec)
        return est_spec, _clamped(wav)


# ------------------------------------------------------------------------------------ CARN
class _GLU(nn.Module):
    """carn.py:9-27: conv(x) * sigmoid(conv'(x)); attribute names follow the reference per flavour."""

    def __init__(self, transposed, cin, cout, kernel, **kw):
        super().__init__()
        mk = nn.ConvTranspose2d if transposed else nn.Conv2d
        self._names = ('conv_transpose1', 'conv_transpose2') if transposed else ('conv1', 'conv2')
        setattr(self, self._names[0], mk(cin, cout, kernel, **kw))
        setattr(self, self._names[1], mk(cin, cout, kernel, **kw))

    def forward(self, x):
        return getattr(self, self._names[0])(x) * torch.sigmoid(getattr(self, self._names[1])(x))


class _RealBlock(nn.Module):
    """carn.py:30-56: real conv / convT (+ optional gate) -> BatchNorm2d -> PReLU.  Stock PyTorch: the real
    convolutions of CARN are outside the hot path."""

    def __init__(self, transposed, cin, cout, kernel, gate=False, **kw):
        super().__init__()
        name = 'conv_transposed' if transposed else 'conv'
        if gate:
            layer = _GLU(transposed, cin, cout, kernel, bias=False, **kw)
        else:
            layer = (nn.ConvTranspose2d if transposed else nn.Conv2d)(cin, cout, kernel, bias=False, **kw)
        setattr(self, name, layer)
        self._name = name
        self.norm = nn.BatchNorm2d(cout)
        self.act = nn.PReLU()

    def forward(self, x):
        return self.act(self.norm(getattr(self, self._name)(x)))


class _CarnAttention(nn.Module):
    """carn.py:59-78."""

    def __init__(self, c):
        super().__init__()
        self.conv1 = nn.Conv2d(c, c * 2, 3, padding=1, bias=False)
        self.conv2 = nn.Conv2d(c, c * 2, 3, padding=1, bias=False)
        self.conv3 = nn.Conv2d(c * 2, c, 3, padding=1, bias=False)

    def forward(self, x_u, x_c):
        gate = torch.sigmoid(self.conv3(torch.sigmoid(self.conv1(x_u) + self.conv2(x_c))))
        return gate * x_c


class _CarnDecoder(nn.Module):
    def __init__(self, cin=128, gate=False):
        super().__init__()
        self.conv_transpose_layers, self.attention_layers = nn.ModuleList(), nn.ModuleList()
        for cout in (128, 96, 64, 32, 16, 2):
            self.attention_layers.append(_CarnAttention(cin))
            self.conv_transpose_layers.append(_RealBlock(True, cin * 2, cout, (1, 3), gate=gate, stride=(2, 1),
                                                         padding=(0, 1), output_padding=(1, 0)))
            cin = cout


class CARN(nn.Module):
    """CARN / GCARN (carn.py:122-178): real-valued encoder / decoder with attention skips around a 2-layer
    LSTM; the hot-path pieces are the STFT, the complex mask with the reference's sign quirk, and the iSTFT.
    All three carry autograd, so this model trains (BASELINE config 4)."""

    def __init__(self, window_size=320, hop_size=160, fft_size=512, lstm_channels=512, gate=False, ops=None):
        super().__init__()
        ops = ops or cuda_ops()
        self._ops = ops
        self.fft_size = fft_size
        self.stft = ops.ConvSTFT(window_size, hop_size, fft_size)
        self.istft = ops.ConviSTFT(window_size, hop_size, fft_size)
        enc, cin = [], 2
        for cout in (16, 32, 64, 96, 128, 128):
            enc.append(_RealBlock(False, cin, cout, (3, 3), gate=gate, stride=(2, 1), padding=(1, 1)))
            cin = cout
        self.encoder = _Stack(enc)
        self.decoder = _CarnDecoder(128, gate)
        self.lstm = nn.LSTM(input_size=lstm_channels, hidden_size=lstm_channels, num_layers=2, batch_first=True)
        self.linear = nn.Linear(fft_size, fft_size + 2)

    def forward(self, x):
        n_bins = self.fft_size // 2 + 1
        spec = self.stft(x)
        noisy_real, noisy_imag = _spec_planes(spec, n_bins)
        x = torch.stack([noisy_real, noisy_imag], dim=1)[:, :, 1:, :]
        skips = []
        # the real convolutions / LSTM are stock cuDNN; fp32 math (no TF32) in the forward.  For the backward
        # the caller sets torch.backends.cudnn.allow_tf32 (bench / tests turn it off).
        with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
            for layer in self.encoder.layers:
                x = layer(x)
                skips.append(x)
            b, c, f, t = x.shape
            x, _ = self.lstm(x.reshape(b, c * f, t).permute(0, 2, 1))
            x = x.permute(0, 2, 1).reshape(b, c, f, t)
            for att, layer in zip(self.decoder.attention_layers, self.decoder.conv_transpose_layers):
                skip = skips.pop()
                if x.shape[2] < skip.shape[2]:
                    x = F.pad(x, (0, 0, 0, 1))
                x = layer(torch.cat([att(x, skip), skip], dim=1))
        b, c, f, t = x.shape
        x = self.linear(x.reshape(b, c * f, t).permute(0, 2, 1))
        x = x.permute(0, 2, 1).reshape(b, 2, n_bins, t)
        est_spec = self._ops.mask_carn(noisy_real, noisy_imag, x[:, 0], x[:, 1])
        wav = self.istft(est_spec)
        return est_spec, torch.clamp(wav, -1, 1)


class GCARN(CARN):
    """carn.py:174-178."""

    def __init__(self, window_size=320, hop_size=160, fft_size=512, lstm_channels=512, ops=None):
        super().__init__(window_size, hop_size, fft_size, lstm_channels, gate=True, ops=ops)
