"""world_size-2 gloo run of the data-parallel training step (BASELINE config 4): the flat
gradient after the allreduce equals the mean of the per-shard gradients computed in one process, and both
replicas end the step with identical weights."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from speech_enhancement_b200 import callers, sharding, training
from speech_enhancement_b200.synth import seeded_state_dict, synth_noisy_speech
from oracle.modules_ref import OPS


def _model():
    m = callers.CARN(ops=OPS).train()            # CPU oracle ops: this test is about the host logic
    m.load_state_dict(seeded_state_dict(m, seed=12))
    return m


def _data():
    return synth_noisy_speech(4, 4000, seed=40), synth_noisy_speech(4, 4000, seed=41)


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        model = _model()
        # 1 MB buckets: several asynchronous exchanges start from the gradient hooks while backward still runs
        opt, flat = training.make_optimizer(model), training.FlatGradients(model, bucket_bytes=1 << 20)
        assert len(flat.buckets) > 4 and sum(b.numel() for b in flat.buckets) == flat.buffer.numel()
        mix, clean = _data()
        loss = training.train_step(model, opt, flat, sharding.shard_batch(mix), sharding.shard_batch(clean))
        torch.save({'grad': flat.buffer.clone(), 'loss': loss, 'w': model.linear.weight.detach().clone()},
                   f'{out}.{rank}')
    finally:
        dist.destroy_process_group()


def test_flat_gradient_allreduce_two_ranks(tmp_path):
    out = str(tmp_path / 'r')
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r0, r1 = torch.load(out + '.0'), torch.load(out + '.1')
    assert torch.equal(r0['grad'], r1['grad']) and torch.equal(r0['w'], r1['w'])
    # single process: per-shard gradients (own BN statistics, as each replica sees them), averaged
    mix, clean = _data()
    grads = []
    for r in range(2):
        model = _model()
        flat = training.FlatGradients(model)
        _, wav = model(sharding.shard_batch(mix, r, 2))
        c = sharding.shard_batch(clean, r, 2)
        training.si_snr_loss(training.pad_or_truncate(wav, c), c).backward()
        grads.append(flat.buffer.clone())
    want = (grads[0] + grads[1]) / 2
    # the worker's buffer was clipped after the reduction (clip_norm 0.5): compare directions and the clip
    norm = want.norm()
    want = want * min(1.0, 0.5 / (norm.item() + 1e-6))
    assert torch.allclose(r0['grad'], want, rtol=1e-4, atol=1e-7)


def test_flat_gradients_are_views_and_survive_zero_grad():
    model = _model()
    flat = training.FlatGradients(model)
    opt = training.make_optimizer(model)
    assert sum(p.numel() for p in model.parameters()) == flat.buffer.numel()
    opt.zero_grad(set_to_none=True)
    flat.rebind()
    p = list(model.parameters())[-1]              # reverse registration order: the last parameter leads the buffer
    assert p.grad.data_ptr() == flat.buffer.data_ptr()
    x = torch.randn(1, 4000) * 0.1
    est = model(x)[1]
    training.si_snr_loss(training.pad_or_truncate(est, x), x).backward()
    assert flat.buffer.abs().sum() > 0
