"""Oracle (test infrastructure): CPU restatement of the mask / magnitude-phase math.

Reference sites (all under /root/reference/models/):
  DCCRN  _2008_00264_dccrn.py:187-212     mask_dccrn_ref  (modes 'R', 'C', 'E')
  DCUNet _1903_03107_dcunet.py:158-189    mask_dcunet_ref ('unbounded', 'bounded_sigmoid', 'bounded_tanh')
  FRCRN  _2206_07293_frcrn.py:141-152     mask_frcrn_ref
  CARN   _2104_05267_carn.py:161-169      mask_carn_ref   (keeps the sign quirk of :166)
  iSTFT  conv_stft.py:96-100              polar_to_rect_ref
All take/return [B, F, T] planes and return the [B, 2F, T] real|imag stack the
iSTFT consumes.
"""
import torch


def _mag_phase(re, im, eps=1e-8):
    return torch.sqrt(re * re + im * im + eps), torch.atan2(im, re)


def mask_dccrn_ref(mode, nr, ni, mr, mi):
    if mode == 'R':                                              # :188-190
        re, im = nr * mr, ni * mi
    elif mode == 'C':                                            # :191-193
        re, im = nr * mr - ni * mi, nr * mi + ni * mr
    elif mode == 'E':                                            # :194-205
        n_mag, n_ph = _mag_phase(nr, ni)
        m_mag, _ = _mag_phase(mr, mi)
        m_ph = torch.atan2(mi / m_mag, mr / m_mag)               # phase re-derived after the divide
        g = n_mag * torch.tanh(m_mag)
        re, im = g * torch.cos(n_ph + m_ph), g * torch.sin(n_ph + m_ph)
    else:
        raise ValueError(mode)
    return torch.cat([re, im], dim=1)


def mask_dcunet_ref(method, nr, ni, mr, mi):
    if method == 'unbounded':                                    # :159-161
        re, im = nr * mr, ni * mi
    elif method == 'bounded_sigmoid':                            # :163-165
        re, im = nr * torch.sigmoid(mr), ni * torch.sigmoid(mi)
    elif method == 'bounded_tanh':                               # :167-182
        m_mag, m_ph = _mag_phase(mr, mi)
        n_mag, n_ph = _mag_phase(nr, ni)
        m_ph = m_ph / m_mag                                      # :177 (phase divided by magnitude)
        g = n_mag * torch.tanh(m_mag)
        re, im = g * torch.cos(n_ph + m_ph), g * torch.sin(n_ph + m_ph)
    else:
        raise ValueError(method)
    return torch.cat([re, im], dim=1)


def mask_frcrn_ref(nr, ni, mr, mi):
    """nr, ni: noisy [B,F,T] incl. DC row; mr, mi: final_conv output [B,F-2,T].

    :141-152 -- the net output gets a zero row on top (F.pad, :141), tanh, is multiplied
    with the DC-less spec (F-1 rows, :125) and the product gets another zero row on top
    (:146).  Net effect: out[0] = out[1] = 0 and out[r] = tanh(m[r-2]) * n[r] for r >= 2.
    """
    import torch.nn.functional as F
    mask_r = torch.tanh(F.pad(mr, (0, 0, 1, 0)))
    mask_i = torch.tanh(F.pad(mi, (0, 0, 1, 0)))
    re = F.pad(mask_r * nr[:, 1:], (0, 0, 1, 0))
    im = F.pad(mask_i * ni[:, 1:], (0, 0, 1, 0))
    return torch.cat([re, im], dim=1)


def mask_carn_ref(nr, ni, mr, mi):
    re = mr * nr - mi * ni                                       # :165
    im = mr * ni - mi * nr                                       # :166 (sign as written)
    return torch.cat([re, im], dim=1)


def polar_to_rect_ref(mags, phase):
    return torch.cat([mags * torch.cos(phase), mags * torch.sin(phase)], dim=1)
