"""Train-mode ComplexBatchNorm2d on csrc/cbn_train.cu: the fused forward / backward against the
reference's own module (oracle/_ref, CPU) where staged, else against this package's plain-PyTorch path."""
import warnings

import pytest
import torch

from oracle import ref_loader
from speech_enhancement_b200.modules.complex_aux import ComplexBatchNorm2d
from conftest import rand, rel_l2

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _reference_cbn(channels, **kw):
    if ref_loader.available():
        return ref_loader.load_original().complex_nn.ComplexBatchNorm2d(channels, **kw), 'reference'
    m = ComplexBatchNorm2d(channels, **kw)
    m.fused_train = False
    return m, 'port'


@pytest.mark.parametrize('B,C2,H,W,affine,momentum', [(3, 8, 5, 7, True, 0.1), (2, 64, 33, 70, True, 0.1),
                                                      (4, 256, 4, 61, True, None), (2, 32, 17, 9, False, 0.1),
                                                      (1, 2, 256, 643, True, 0.1)])
def test_fused_train_cbn_matches_reference_forward_backward_and_running_stats(B, C2, H, W, affine, momentum, built_lib):
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        ref, kind = _reference_cbn(C2, affine=affine, momentum=momentum)
        ours = ComplexBatchNorm2d(C2, affine=affine, momentum=momentum)
        ours.load_state_dict(ref.state_dict())
        if affine:
            with torch.no_grad():
                for p, seed in zip((ref.Wrr, ref.Wii, ref.Br, ref.Bi), range(4)):
                    p.copy_(p + 0.2 * rand(p.shape, 40 + seed))
            ours.load_state_dict(ref.state_dict())
        ours.to(DEV)
        ref.train()
        ours.train()
        # un-centred, unequal real / imaginary scales: raw second moments would cancel in fp32
        x = rand((B, C2, H, W), 7) * 1.5 + 3.0
        x[:, C2 // 2:] *= 0.3
        go = rand((B, C2, H, W), 8)
        for step in range(2):                                   # two steps: the running statistics move twice
            xr = x.clone().requires_grad_(True)
            xo = x.clone().to(DEV).requires_grad_(True)
            n0 = built_lib.se_launch_count()
            y_ref = ref(xr)
            y = ours(xo)
            assert built_lib.se_launch_count() - n0 == 2        # statistics pass + apply pass
            y_ref.backward(go)
            n0 = built_lib.se_launch_count()
            y.backward(go.to(DEV))
            assert built_lib.se_launch_count() - n0 == 2
            assert rel_l2(y.detach().cpu(), y_ref.detach()) < 2e-6, kind
            assert rel_l2(xo.grad.cpu(), xr.grad) < 2e-5, kind
            if affine:
                for name in ('Wrr', 'Wri', 'Wii', 'Br', 'Bi'):
                    a, b = getattr(ours, name).grad.cpu(), getattr(ref, name).grad
                    assert (a - b).abs().max().item() <= 2e-5 * max(b.abs().max().item(), 1.0), (name, kind)
            for name in ('RMr', 'RMi', 'RVrr', 'RVri', 'RVii'):
                a, b = getattr(ours, name).cpu(), getattr(ref, name)
                assert (a - b).abs().max().item() <= 1e-5 * max(b.abs().max().item(), 1.0), (name, kind)
            assert int(ours.num_batches_tracked) == int(ref.num_batches_tracked) == step + 1
            ref.zero_grad()
            ours.zero_grad()
        # eval mode uses the running statistics both sides just accumulated
        ref.eval()
        ours.eval()
        with torch.no_grad():
            assert rel_l2(ours(x.to(DEV)).cpu(), ref(x)) < 2e-6


def test_fused_train_cbn_switch_and_fallbacks():
    m = ComplexBatchNorm2d(8).to(DEV).train()
    x = rand((2, 8, 6, 5), 3).to(DEV)
    m2 = ComplexBatchNorm2d(8).to(DEV).train()
    m2.load_state_dict(m.state_dict())                          # Wri is drawn at random by reset_parameters
    m2.fused_train = False
    y1, y2 = m(x), m2(x)
    assert rel_l2(y1.detach().cpu(), y2.detach().cpu()) < 2e-6
    # 3-D inputs (not an NCHW image) stay on the PyTorch path
    m(torch.randn(2, 8, 6, device=DEV))
