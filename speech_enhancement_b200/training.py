"""One training step of a caller model, data-parallel over the GPUs of one box (BASELINE config 4,
SURVEY.md 8d/8e): forward + SI-SNR loss + backward through the hot path's autograd nodes, a bucketed
allreduce of a flat fp32 gradient buffer that overlaps the backward pass (NCCL over NVLink on GPUs, gloo in the
CPU tests), gradient
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
    """all gradients of a model as views into ONE contiguous fp32 buffer, exchanged in a few large buckets
    (CARN: 7.6 M parameters = 30 MB -> 4 buckets of <= 8 MB).

    Parameters are laid out in REVERSE registration order, so the buffer fills front to back while backward runs
    (the last layers' gradients are ready first).  With `arm()` before the backward pass, a post-accumulate hook
    per parameter counts each bucket down and, the moment its last gradient has been accumulated, starts that
    bucket's allreduce asynchronously (NCCL's own stream, ordered behind the kernels issued so far), so the
    exchange of bucket k travels over NVLink while backward is still producing bucket k+1 ("on a
    side stream overlapped with the tail of backward").  `finish()` waits for the buckets and applies 1/world."""

    def __init__(self, model, bucket_bytes=8 << 20):
        self.params = [p for p in model.parameters() if p.requires_grad]
        order = self.params[::-1]
        n = sum(p.numel() for p in order)
        dev = order[0].device if order else torch.device('cpu')
        self.buffer = torch.zeros(n, dtype=torch.float32, device=dev)
        self._offsets, self._bucket_of, bounds = {}, {}, [0]
        off = 0
        for p in order:
            if off - bounds[-1] > 0 and (off + p.numel() - bounds[-1]) * 4 > bucket_bytes:
                bounds.append(off)
            self._offsets[id(p)] = off
            self._bucket_of[id(p)] = len(bounds) - 1
            p.grad = self.buffer[off:off + p.numel()].view_as(p)
            off += p.numel()
        bounds.append(off)
        self.buckets = [self.buffer[bounds[i]:bounds[i + 1]] for i in range(len(bounds) - 1)]
        self._members = [0] * len(self.buckets)
        for p in order:
            self._members[self._bucket_of[id(p)]] += 1
        self._armed, self._left, self._work, self._started = False, [], [], []
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in order]

    def zero(self):
        self.buffer.zero_()

    def rebind(self):
        """optimizers / zero_grad(set_to_none=True) may drop .grad: point it back at the flat views"""
        for p in self.params:
            off = self._offsets[id(p)]
            view = self.buffer[off:off + p.numel()].view_as(p)
            if p.grad is None or p.grad.data_ptr() != view.data_ptr():
                p.grad = view

    @staticmethod
    def _world():
        return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1

    def arm(self, reduce=True):
        """call before backward(): buckets are exchanged as they complete (no-op for a single process)"""
        self._armed = bool(reduce) and self._world() > 1
        self._left = list(self._members)
        self._work, self._started = [], [False] * len(self.buckets)

    def _on_grad(self, p):
        if not self._armed:
            return
        b = self._bucket_of[id(p)]
        self._left[b] -= 1
        if self._left[b] == 0 and not self._started[b]:
            self._started[b] = True
            self._work.append(dist.all_reduce(self.buckets[b], op=dist.ReduceOp.SUM, async_op=True))

    def finish(self):
        """after backward(): exchange what the hooks have not started (parameters without a gradient this step
        never fire), wait for every bucket, scale to the mean"""
        if not self._armed:
            return
        for b, started in enumerate(self._started):
            if not started:
                self._work.append(dist.all_reduce(self.buckets[b], op=dist.ReduceOp.SUM, async_op=True))
        for w in self._work:
            w.wait()
        self.buffer.div_(self._world())
        self._armed, self._work = False, []

    def allreduce_mean(self):
        """the whole buffer in one blocking exchange (the non-overlapped form)"""
        if self._world() > 1:
            dist.all_reduce(self.buffer, op=dist.ReduceOp.SUM)
            self.buffer.div_(self._world())


def make_optimizer(model, lr=1e-3, weight_decay=1e-2):
    """hparams/hyperparams.py:27-30, :38"""
    return torch.optim.AdamW(model.parameters(), lr=lr, weight_decay=weight_decay)


def train_step(model, optimizer, flat, mix, clean, clip_norm=0.5, reduce=True):
    """one data-parallel step on this rank's shard; returns the local loss (python float is NOT taken here:
    no host sync inside the step).  reduce=False skips the allreduce (single-replica timing legs)."""
    flat.rebind()
    flat.zero()
    flat.arm(reduce)
    _, wav = model(mix)
    loss = si_snr_loss(pad_or_truncate(wav, clean), clean)
    loss.backward()
    flat.finish()
    if clip_norm:
        torch.nn.utils.clip_grad_norm_(flat.params, clip_norm)
    optimizer.step()
    return loss.detach()
