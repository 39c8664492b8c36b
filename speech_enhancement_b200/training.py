"""One training step of a caller model, data-parallel over the GPUs of one box (BASELINE config 4,
SURVEY.md 8d/8e): forward + SI-SNR loss + backward through the hot path's autograd nodes, ONE
allreduce of a flat fp32 gradient buffer (NCCL over NVLink on GPUs, gloo in the CPU tests), gradient
clipping on the reduced gradient, AdamW.

The reference's own step is trainer.py:210-221 with SI_SNR_loss (losses.py:62-84) on the
pad_or_truncate_wav output (utils.py:111-122); it has no working multi-GPU path, the
replica / allreduce scheme here is the standard non-synchronised-BN data parallelism.
"""
import torch
import torch.distributed as dist
import torch.nn.functional as F


def si_snr_loss(estimate, target, zero_mean=False, eps=1e-8):
    """negative mean scale-invariant SNR in dB, [B, L] x [B, L] (losses.py:62-84)"""
    if zero_mean:
        estimate = estimate - estimate.mean(dim=1, keepdim=True)
        target = target - target.mean(dim=1, keepdim=True)
    dot = (estimate * target).sum(dim=1, keepdim=True)
    s_target = dot * target / target.norm(dim=1, keepdim=True) ** 2
    num = s_target.norm(dim=1) ** 2 + eps
    den = (estimate - s_target).norm(dim=1) ** 2 + eps
    return -(10 * torch.log10(num / den)).mean()


def pad_or_truncate(estimate, target):
    """bring the enhanced waveform to the target's length (utils.py:111-122)"""
    le, lt = estimate.shape[-1], target.shape[-1]
    if le < lt:
        return F.pad(estimate, (0, lt - le))
    return estimate[:, :lt]


class FlatGradients:
    """all gradients of a model as views into ONE contiguous fp32 buffer, so a data-parallel step needs a
    single allreduce (CARN: 7.6 M parameters = 30 MB; ~60 us over NVSwitch at 8 GPUs)."""

    def __init__(self, model):
        self.params = [p for p in model.parameters() if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        dev = self.params[0].device if self.params else torch.device('cpu')
        self.buffer = torch.zeros(n, dtype=torch.float32, device=dev)
        off = 0
        for p in self.params:
            p.grad = self.buffer[off:off + p.numel()].view_as(p)
            off += p.numel()

    def zero(self):
        self.buffer.zero_()

    def rebind(self):
        """optimizers / zero_grad(set_to_none=True) may drop .grad: point it back at the flat views"""
        off = 0
        for p in self.params:
            view = self.buffer[off:off + p.numel()].view_as(p)
            if p.grad is None or p.grad.data_ptr() != view.data_ptr():
                p.grad = view
            off += p.numel()

    def allreduce_mean(self):
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.buffer, op=dist.ReduceOp.SUM)
            self.buffer.div_(dist.get_world_size())


def make_optimizer(model, lr=1e-3, weight_decay=1e-2):
    """hparams/hyperparams.py:27-30, :38"""
    return torch.optim.AdamW(model.parameters(), lr=lr, weight_decay=weight_decay)


def train_step(model, optimizer, flat, mix, clean, clip_norm=0.5, reduce=True):
    """one data-parallel step on this rank's shard; returns the local loss (python float is NOT taken here:
    no host sync inside the step).  reduce=False skips the allreduce (single-replica timing legs)."""
    flat.rebind()
    flat.zero()
    _, wav = model(mix)
    loss = si_snr_loss(pad_or_truncate(wav, clean), clean)
    loss.backward()
    if reduce:
        flat.allreduce_mean()
    if clip_norm:
        torch.nn.utils.clip_grad_norm_(flat.params, clip_norm)
    optimizer.step()
    return loss.detach()
