"""world_size-2 (and 3) gloo runs of the utterance-sharding host logic: the shards tile the
batch, results gathered on rank 0 equal the single-process result, the timing reduction is a max."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from speech_enhancement_b200 import callers, sharding
from speech_enhancement_b200.synth import seeded_state_dict, synth_noisy_speech
from oracle.modules_ref import OPS


def test_shard_bounds_tile_the_batch():
    for n in (0, 1, 5, 8, 64, 129):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_bounds(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        sharding.shard_bounds(4, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_utt, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        # the CPU oracle stands in for the device model here: this test is about the host logic
        model = callers.DCCRN('dccrn-CL', 400, 100, 512, ops=OPS).eval()
        model.load_state_dict(seeded_state_dict(model, seed=3))
        x = synth_noisy_speech(n_utt, 4000, seed=21)
        mine = sharding.shard_batch(x)
        assert mine.shape[0] == sharding.shard_bounds(n_utt, rank, world)[1] - sharding.shard_bounds(n_utt, rank, world)[0]
        wav = sharding.enhance_sharded(model, x, dst=0)
        slowest = sharding.reduce_max(10.0 + rank)
        assert slowest == 10.0 + world - 1
        if rank == 0:
            torch.save(wav, out)
        else:
            assert wav is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize('world,n_utt', [(2, 4), (2, 3), (3, 2)])
def test_sharded_enhancement_matches_single_process(tmp_path, world, n_utt):
    out = str(tmp_path / 'wav.pt')
    mp.spawn(_worker, args=(world, _free_port(), n_utt, out), nprocs=world, join=True)
    got = torch.load(out)
    model = callers.DCCRN('dccrn-CL', 400, 100, 512, ops=OPS).eval()
    model.load_state_dict(seeded_state_dict(model, seed=3))
    with torch.no_grad():
        want = model(synth_noisy_speech(n_utt, 4000, seed=21))[1]
    assert got.shape == want.shape
    # same per-utterance arithmetic whatever the shard it ran in (batch-size dependent BLAS blocking aside)
    assert torch.allclose(got, want, rtol=0, atol=2e-6)
