"""Deterministic synthetic weights and 16 kHz noisy-speech inputs.

There is no network for datasets or checkpoints, so benches and parity tests use
(a) `seeded_state_dict`: a state_dict generated key-by-key from a seed, applicable to the
    reference models, the CPU oracle and the CUDA build alike (their keys are identical), and
(b) `synth_noisy_speech`: harmonic voiced-speech-like source + white noise inside (-1, 1).
"""
import math

import torch


def seeded_state_dict(model, seed=0):
    """New values for every floating-point entry of model.state_dict(), in sorted-key order.

    Weights (dim >= 2): N(0, 2/fan_in) (He); LSTM/linear biases: small normal; complex-BN
    running stats randomised so eval-mode BN is not a near-identity (RM ~ N(0,0.1),
    RVrr/RVii ~ U(0.5,1.5), RVri ~ U(-0.2,0.2)); BN affine W near identity with a random
    cross term; PReLU slopes ~ U(0.1,0.4).  STFT bases / windows (`stft.*`, `istft.*`) and
    integer buffers are left untouched.
    """
    g = torch.Generator().manual_seed(seed)
    sd = model.state_dict()
    out = {}
    for key in sorted(sd.keys()):
        v = sd[key]
        leaf = key.rsplit('.', 1)[-1]
        if not v.is_floating_point() or '.stft.' in '.' + key or '.istft.' in '.' + key \
                or key.startswith(('stft.', 'istft.')):
            out[key] = v.clone()
            continue
        shape = tuple(v.shape)
        u = lambda lo, hi: torch.rand(shape, generator=g) * (hi - lo) + lo
        n = lambda std: torch.randn(shape, generator=g) * std
        if leaf in ('RMr', 'RMi'):
            t = n(0.1)
        elif leaf in ('RVrr', 'RVii'):
            t = u(0.5, 1.5)
        elif leaf == 'RVri':
            t = u(-0.2, 0.2)
        elif leaf in ('Wrr', 'Wii'):
            t = u(0.8, 1.2)
        elif leaf == 'Wri':
            t = u(-0.3, 0.3)
        elif leaf in ('Br', 'Bi'):
            t = n(0.05)
        elif leaf in ('running_mean',):
            t = n(0.1)
        elif leaf in ('running_var',):
            t = u(0.5, 1.5)
        elif v.dim() >= 2:
            if 'conv_transposed' in key and v.dim() == 4:      # ConvTranspose2d weight [Ci, Co, kh, kw]
                fan_in = v.shape[0] * v.shape[2] * v.shape[3]
            else:
                fan_in = v[0].numel()
            t = n(math.sqrt(2.0 / max(fan_in, 1)))
        elif v.dim() == 1 and v.numel() == 1 and leaf == 'weight':
            t = u(0.1, 0.4)                        # PReLU slope
        elif leaf == 'weight':
            t = u(0.8, 1.2)                        # real BN gamma
        else:
            t = n(0.02)                            # biases
        out[key] = t.to(v.dtype)
    return out


def synth_noisy_speech(batch, length, seed=0, sample_rate=16000, return_clean=False):
    """[batch, length] fp32 in (-1, 1): 19 harmonics of f0 ~ U(100,200) Hz with 1/h amplitudes and
    random phases, syllabic envelope (0.5 + 0.5 sin 2 pi 3 t)^2, peak 0.1, plus N(0, 0.03^2) noise."""
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(length, dtype=torch.float64) / sample_rate
    f0 = 100.0 + 100.0 * torch.rand(batch, 1, generator=g, dtype=torch.float64)
    clean = torch.zeros(batch, length, dtype=torch.float64)
    for h in range(1, 20):
        ph = 2 * math.pi * torch.rand(batch, 1, generator=g, dtype=torch.float64)
        clean += torch.sin(2 * math.pi * h * f0 * t[None, :] + ph) / h
    clean *= (0.5 + 0.5 * torch.sin(2 * math.pi * 3.0 * t))[None, :] ** 2
    clean *= 0.1 / clean.abs().amax(dim=1, keepdim=True).clamp_min(1e-9)
    noise = 0.03 * torch.randn(batch, length, generator=g, dtype=torch.float64)
    noisy = (clean + noise).clamp(-0.999, 0.999).float()
    return (noisy, clean.float()) if return_clean else noisy
