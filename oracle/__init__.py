"""CPU oracle for the STFT / complex-conv / mask hot path.

TEST INFRASTRUCTURE ONLY.  This package is a CPU (torch fp32/fp64 + numpy)
restatement of the reference algorithm for the path in SURVEY.md section 8.  It is
imported only by ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs, always as the checker or the timed
CPU baseline and never by the product package ``speech_enhancement_b200``.

Parity pinning: the reference ships no tests or golden vectors (SURVEY.md
section 4), so the oracle is pinned against outputs of the reference itself,
imported from /root/reference in the authoring container by
``tests/golden/make_golden.py`` and committed as ``tests/golden/*.npz``.
``tests/test_oracle_golden.py`` re-checks the oracle against those vectors on
every CPU run.
"""
from .stft_ref import init_kernels_ref, stft_ref, stft_magphase_ref, istft_ref, ola_norm_ref
from .cconv_ref import cconv2d_ref, cconvT2d_ref, pack_conv_weight_ref, pack_convT_weight_ref
from .mask_ref import (mask_dccrn_ref, mask_dcunet_ref, mask_frcrn_ref, mask_carn_ref,
                       polar_to_rect_ref)
