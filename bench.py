#!/usr/bin/env python
"""bench.py -- DCCRN-CL enhancement throughput (audio-seconds enhanced per second) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one forward of DCCRN-CL (400/100/512, _2008_00264_dccrn.py:124) over a batch of
`--batch` synthetic 4 s / 16 kHz noisy utterances PER GPU (utterance-sharded, weak scaling, no
collective on the data path).  Rank 0 prints ONE JSON line (contract in the task statement):
  value     audio-s/s with inputs resident in HBM (CUDA events, max over ranks)
  e2e       same metric through the public API from pinned HOST buffers: H2D of the batch and
            D2H of the enhanced waveform inside the timed region
  roofline  the complex-conv stack (dominant kernels): algorithmic FLOPs (SURVEY.md 8a,
            65.95 GFLOP/utt) / summed CUDA-event time of the conv launches, vs MEASURED_PEAKS.json
  stages    per-stage device time + HBM GB/s of STFT / iSTFT / mask against the measured copy peak
  cpu_baseline  the CPU oracle (torch fp32 restatement of the reference path) timed on this box
`--impl reference` times that CPU oracle alone (the reference is pure Python and cannot travel to
the GPU box) on all host threads, same metric/unit/config.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SR, CLIP_S = 16000, 4
L = SR * CLIP_S
UNIT = 'audio-s/s'
# BASELINE.json's metric is quoted on DCCRN (default); the two other complex families of configs[1] / [2]
# run through the same bench with --model.  conv GFLOP/utt: SURVEY.md 8a/8d (sum of 2*M*K*N per layer).
MODELS = {
    'dccrn': dict(cls='DCCRN', args=('dccrn-CL', 400, 100, 512), stft=(400, 100, 512), conv_gflop=65.95,
                  name='DCCRN-CL', metric='dccrn_cl_audio_seconds_enhanced_per_second', act_mb=268),
    'dcunet20': dict(cls='DCUNet', args=('dcunet20', 512, 128, 512), stft=(512, 128, 512), conv_gflop=152.99,
                     name='DCUNet-20', metric='dcunet20_audio_seconds_enhanced_per_second', act_mb=600),
    'frcrn': dict(cls='FRCRN', args=(320, 160, 640), stft=(320, 160, 640), conv_gflop=97.61,
                  name='FRCRN', metric='frcrn_audio_seconds_enhanced_per_second', act_mb=500),
}
# BASELINE configs[3]: CARN forward+backward+AdamW, GLOBAL batch 32 split over the GPUs (strong scaling),
# one flat-gradient allreduce per step (speech_enhancement_b200/training.py)
MODELS['carn_train'] = dict(cls='CARN', args=(320, 160, 512), stft=(320, 160, 512), conv_gflop=0.0, train=True,
                            name='CARN training step', metric='carn_train_audio_seconds_per_second', act_mb=0)
MODEL = MODELS['dccrn']


def peaks():
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            p = json.load(f)
        return {'hbm_gbs': p['hbm_gbs'], 'bf16_tflops': p['bf16_tflops'],
                'bf16_tflops_sustained': p.get('bf16_tflops_sustained', p['bf16_tflops']), 'source': 'measured'}
    except Exception:
        return {'hbm_gbs': 6650.0, 'bf16_tflops': 1590.0, 'bf16_tflops_sustained': 1400.0, 'source': 'fallback'}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--id={self.index}', f'--query-gpu={self.Q}',
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v == 'Active'})
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': reasons, 'samples': len(sm)}


def build_models(device, seed=0):
    from speech_enhancement_b200 import callers
    from speech_enhancement_b200.synth import seeded_state_dict
    model = getattr(callers, MODEL['cls'])(*MODEL['args'])
    model.load_state_dict(seeded_state_dict(model, seed=seed))
    return model.eval().to(device)


def build_oracle(seed=0):
    from speech_enhancement_b200 import callers
    from speech_enhancement_b200.synth import seeded_state_dict
    from oracle.modules_ref import OPS
    model = getattr(callers, MODEL['cls'])(*MODEL['args'], ops=OPS)
    model.load_state_dict(seeded_state_dict(model, seed=seed))
    return model.eval()


def time_cpu_oracle(batch, iters, warmup=1):
    """audio-s/s of the CPU oracle on `batch` x 4 s, all host threads; returns (value, cores, seconds/iter)."""
    from speech_enhancement_b200.synth import synth_noisy_speech
    cores = len(os.sched_getaffinity(0)) if hasattr(os, 'sched_getaffinity') else (os.cpu_count() or 1)
    torch.set_num_threads(cores)
    torch.set_flush_denormal(True)      # CPU at its best: denormals in the synthetic activations cost 3x otherwise
    model = build_oracle()
    x = synth_noisy_speech(batch, L, seed=1)
    with torch.no_grad():
        for _ in range(warmup):
            model(x)
        ts = []
        for _ in range(iters):
            t0 = time.perf_counter()
            model(x)
            ts.append(time.perf_counter() - t0)
    ts.sort()
    med = ts[len(ts) // 2]
    return batch * CLIP_S / med, torch.get_num_threads(), med


def config_dict(batch, n_gpus):
    win, hop, fft = MODEL['stft']
    return {'workload': f'{MODEL["name"]} forward (win {win} / hop {hop} / fft {fft}), {batch} x 4 s 16 kHz utterances '
                        'per GPU, fp32 I/O, eval/no_grad, seeded random-init weights',
            'per_gpu_batch': batch, 'global_batch': batch * n_gpus, 'clip_seconds': CLIP_S, 'sample_rate': SR,
            'parallelism': f'utterance-sharded x{n_gpus}, no data-path collective',
            'l2_policy': f'per-step activation working set (>= {MODEL["act_mb"]} MB/utt) exceeds the 126 MB L2; '
                         'inputs rotate over 3 buffers'}


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    batch = min(args.batch, 2)
    value, cores, _ = time_cpu_oracle(batch, iters=max(args.steps, 1), warmup=max(args.warmup, 1))
    sec = batch * CLIP_S / value
    line = {'impl': 'reference', 'metric': MODEL['metric'], 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': sec * 1e3, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': config_dict(args.batch, args.gpus),
            'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                             'sample': f'{batch} x 4 s utterances per step (bounded sample of the workload), '
                                       'CPU oracle = torch-CPU restatement of the reference path, flush-denormal on'},
            'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------ GPU arm
class StageTimer:
    """CUDA-event timing of the hot-path stages through the package's `_abi.stage` brackets (used in a
    SEPARATE profiling pass after the headline timing, so it never perturbs `value`)."""

    def __init__(self):
        from speech_enhancement_b200 import _abi
        self._abi = _abi
        self.events = {}
        _abi._timer = self

    def start(self, k):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        self.events.setdefault(k, []).append([e, None])

    def stop(self, k):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        self.events[k][-1][1] = e

    def close(self):
        self._abi._timer = None

    def totals_ms(self):
        torch.cuda.synchronize()
        return {k: sum(a.elapsed_time(b) for a, b in v) for k, v in self.events.items()}


def time_cpu_train(batch, iters):
    """CARN step on the CPU oracle (B_cpu = 4), all host threads"""
    from speech_enhancement_b200 import training
    from speech_enhancement_b200.synth import synth_noisy_speech
    cores = len(os.sched_getaffinity(0)) if hasattr(os, 'sched_getaffinity') else (os.cpu_count() or 1)
    torch.set_num_threads(cores)
    torch.set_flush_denormal(True)
    model = build_oracle().train()
    opt, flat = training.make_optimizer(model), training.FlatGradients(model)
    mix, clean = synth_noisy_speech(batch, L, seed=1), synth_noisy_speech(batch, L, seed=2)
    training.train_step(model, opt, flat, mix, clean, reduce=False)
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter()
        training.train_step(model, opt, flat, mix, clean, reduce=False)
        ts.append(time.perf_counter() - t0)
    ts.sort()
    med = ts[len(ts) // 2]
    return batch * CLIP_S / med, torch.get_num_threads(), med


def run_train(args):
    """--model carn_train: one data-parallel training step per bench step."""
    import torch.distributed as dist
    from speech_enhancement_b200 import _abi, sharding, training
    from speech_enhancement_b200.synth import synth_noisy_speech

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    GB = 32                                                     # global batch of configs[3]
    if args.impl == 'reference':
        if rank == 0:
            value, cores, sec = time_cpu_train(4, max(args.steps, 1))
            print(json.dumps({'impl': 'reference', 'metric': MODEL['metric'], 'value': value, 'unit': UNIT,
                              'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': sec * 1e3,
                              'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f32',
                              'data': 'synthetic', 'config': {'workload': 'CARN fwd+bwd+AdamW step', 'global_batch': GB},
                              'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                                               'sample': '4 x 4 s utterances per step on the CPU oracle'},
                              'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
                              'gpu_launches': 0}), flush=True)
        return
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py: no CUDA device (speech_enhancement_b200 has no CPU path)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    torch.backends.cudnn.allow_tf32 = False                     # fp32 math in the stock convolutions too
    lib = _abi.lib()
    model = build_models(dev).train()
    opt, flat = training.make_optimizer(model), training.FlatGradients(model)
    b0, b1 = sharding.shard_bounds(GB, rank, world)
    host = [(synth_noisy_speech(GB, L, seed=200 + i)[b0:b1].pin_memory(),
             synth_noisy_speech(GB, L, seed=300 + i)[b0:b1].pin_memory()) for i in range(3)]
    dev_in = [(m.to(dev), c.to(dev)) for m, c in host]

    def step(i):
        return training.train_step(model, opt, flat, *dev_in[i % 3])

    def step_e2e(i):
        m, c = host[i % 3]
        loss = training.train_step(model, opt, flat, m.to(dev, non_blocking=True), c.to(dev, non_blocking=True))
        return loss.item()                                      # D2H of the step's result

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        return sharding.reduce_max(e0.elapsed_time(e1), device=dev)

    for i in range(max(args.warmup, 3)):
        step(i)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = lib.se_launch_count()
    ms = timed(step, args.steps)
    launches = lib.se_launch_count() - n0
    step_e2e(0)
    ms_e2e = timed(step_e2e, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    # per-stage profiling pass: EVERY rank steps (the step contains the allreduce), only rank 0 records
    st = StageTimer() if rank == 0 else None
    nprof = min(args.steps, 3)
    for i in range(nprof):
        step(i)
    barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    pk = peaks()
    tot = st.totals_ms()
    st.close()
    win, hop, fft = MODEL['stft']
    T = (L + fft - win) // hop + 1
    nb = b1 - b0
    alg = {'stft': 4 * (L + (fft + 2) * T), 'istft': 4 * ((fft + 2) * T + L), 'mask': 4 * 6 * (fft // 2 + 1) * T}
    alg.update(stft_bwd=alg['stft'], istft_bwd=alg['istft'], mask_bwd=4 * 10 * (fft // 2 + 1) * T)
    stages = {}
    for k, byt in alg.items():
        t = tot.get(k, 0.0) / nprof / 1e3
        stages[k] = {'ms': t * 1e3, 'GB/s': byt * nb / t / 1e9 if t else None,
                     'frac_hbm': byt * nb / t / 1e9 / pk['hbm_gbs'] if t else None}
    hot = sum(v['ms'] for v in stages.values())
    stages['step_ms'] = ms / args.steps
    stages['hot_path_share'] = hot / (ms / args.steps)
    dom = max(alg, key=lambda k: stages[k]['ms'])
    roofline = {'bound': 'hbm', 'kernel': f'{dom} (largest of this package\'s kernels in the step; the real convolutions '
                                          'and the LSTM of CARN are stock cuDNN, out of scope)',
                'achieved': stages[dom]['GB/s'], 'peak': pk['hbm_gbs'], 'unit': 'GB/s',
                'frac': stages[dom]['frac_hbm'], 'traffic': None, 'peak_source': pk['source'] + ' HBM copy'}
    audio_s = GB * CLIP_S * args.steps
    cpu_val, cores, cpu_sec = time_cpu_train(4, 1)
    grad_bytes = flat.buffer.numel() * 4
    line = {'metric': MODEL['metric'], 'value': audio_s / (ms / 1e3), 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': max(args.warmup, 3), 'ms_per_step': ms / args.steps, 'higher_is_better': True, 'scaling': 'strong',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': f'CARN (320/160/512, lstm 512) forward + SI-SNR + backward + clip 0.5 + AdamW, global '
                                   f'batch {GB} x 4 s split {nb} per GPU, one flat-gradient allreduce '
                                   f'({grad_bytes / 1e6:.1f} MB) per step', 'global_batch': GB, 'per_gpu_batch': nb,
                       'parallelism': f'data-parallel x{world}, NCCL allreduce of the flat gradient',
                       'l2_policy': 'inputs rotate over 3 buffers; activations exceed L2'},
            'e2e': {'value': audio_s / (ms_e2e / 1e3), 'unit': UNIT, 'h2d_bytes_per_step': 2 * nb * L * 4 * world,
                    'd2h_bytes_per_step': 4 * world, 'ms_per_step': ms_e2e / args.steps},
            'gpu_launches': int(launches), 'clocks': clocks, 'roofline': roofline, 'stages': stages,
            'cpu_baseline': {'value': cpu_val, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                             'sample': f'one CARN step at 4 x 4 s on the CPU oracle ({cpu_sec:.2f} s)'}}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_ours(args):
    import torch.distributed as dist
    from speech_enhancement_b200 import _abi, sharding
    from speech_enhancement_b200.synth import synth_noisy_speech

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py: no CUDA device (speech_enhancement_b200 has no CPU path)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    lib = _abi.lib()
    if lib.se_device_ok() != 1:
        raise RuntimeError('bench.py: device is not compute capability 10.x')

    B = args.batch
    model = build_models(dev)
    host = [synth_noisy_speech(B, L, seed=100 + 7 * rank + i).pin_memory() for i in range(3)]
    dev_in = [h.to(dev) for h in host]
    out_len = model.istft(model.stft(dev_in[0][:1])).shape[1] if False else None

    def step(i):
        with torch.no_grad():
            return model(dev_in[i % 3])[1]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, after=None):
        """K steps bracketed by barrier + synchronize; device time via CUDA events; max over ranks."""
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        if after is not None:
            after()
        e1.record()
        barrier()
        return sharding.reduce_max(e0.elapsed_time(e1), device=dev)      # the slowest rank defines the step

    for i in range(max(args.warmup, 3)):
        step(i)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = lib.se_launch_count()
    ms = timed(step, args.steps)
    launches = lib.se_launch_count() - n0

    # end-to-end: pinned host -> device -> model -> host, every step, through the public API
    # (speech_enhancement_b200.pipeline.HostPipeline: the copies of neighbouring steps overlap the kernels; every
    # step's H2D and D2H are inside the timed region, the closing event is ordered behind the last D2H)
    from speech_enhancement_b200.pipeline import HostPipeline
    pipe = HostPipeline(model, dev)
    out_len = pipe.output_length(host[0])
    host_outs = [torch.empty((B, out_len), dtype=torch.float32).pin_memory() for _ in range(2)]
    host_out = host_outs[0]

    def step_e2e(i):
        pipe.submit(host[i % 3], host_outs[i & 1])

    step_e2e(0)
    step_e2e(1)
    pipe.synchronize()
    ms_e2e = timed(step_e2e, args.steps, after=pipe.join)
    pipe.synchronize()
    clocks = sampler.stop() if rank == 0 else None

    audio_s = B * CLIP_S * world * args.steps
    value, e2e_value = audio_s / (ms / 1e3), audio_s / (ms_e2e / 1e3)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- per-stage profiling pass (rank 0 only, hooks + events; not part of `value`) ----
    pk = peaks()
    st = StageTimer()
    nprof = min(args.steps, 5)
    for i in range(nprof):
        step(i)
    tot = st.totals_ms()
    st.close()
    for k in ('stft', 'istft', 'mask', 'cconv', 'cl2', 'lstm', 'linear', 'ccbam'):
        tot.setdefault(k, 0.0)
    win, hop, fft = MODEL['stft']
    T = (L + fft - win) // hop + 1
    Lp = out_len or (T - 1) * hop + win - fft
    alg = {'stft': 4 * (L + (fft + 2) * T), 'istft': 4 * ((fft + 2) * T + Lp), 'mask': 4 * 6 * (fft // 2 + 1) * T}
    stages = {}
    for k in ('stft', 'istft', 'mask'):
        t = tot[k] / nprof / 1e3
        gbs = alg[k] * B / t / 1e9
        stages[k] = {'ms': t * 1e3, 'GB/s': gbs, 'frac_hbm': gbs / pk['hbm_gbs']}
    t_conv = tot['cconv'] / nprof / 1e3
    tflops = MODEL['conv_gflop'] * B / t_conv / 1e3
    stages['cconv'] = {'ms': t_conv * 1e3, 'TFLOP/s': tflops, 'launches': len(st.events.get('cconv', [])) // nprof}
    stages['cl2_convert_ms'] = tot['cl2'] / nprof
    stages['lstm_recurrence_ms'] = tot['lstm'] / nprof
    stages['lstm_linear_projections_ms'] = tot['linear'] / nprof
    stages['ccbam_ms'] = tot['ccbam'] / nprof
    stages['step_ms'] = ms / args.steps
    stages['hot_path_share'] = (tot['stft'] + tot['istft'] + tot['mask'] + tot['cconv']) / nprof / (ms / args.steps)
    roofline = {'bound': 'tensor', 'kernel': f'complex conv stack ({len(st.events.get("cconv", [])) // nprof} conv launches per step)',
                'achieved': tflops, 'peak': pk['bf16_tflops_sustained'], 'unit': 'TFLOP/s',
                'frac': tflops / pk['bf16_tflops_sustained'],
                # dram__bytes_read + write summed over the 11 conv GEMM launches of one step from the ncu --set full
                # captures (profiles/r1c_gemm_full_summary.txt, B=64: 25.46 GB; the eight launches re-captured in
                # profiles/r1e_gemm_full_summary.txt / r1g_gemm_full_summary.txt agree within 2 %), scaled per utterance
                'traffic': 25.458e9 / 64 * B if args.model == 'dccrn' else None,
                'traffic_unit': 'B per step (all conv GEMM launches)',
                'peak_source': pk['source'] + ' bf16 dense (sustained)',
                'note': f'algorithmic FLOPs ({MODEL["conv_gflop"]} GFLOP/utt); the tensor engine issues 3 bf16 passes per product'}

    cpu_val, cores, cpu_sec = time_cpu_oracle(2, iters=5, warmup=1)
    stages['stft_engine'] = model.stft._pick_engine('ConvSTFT', model.stft.weight[0, 0], False)
    stages['istft_engine'] = model.istft._pick_engine('ConviSTFT', model.istft.window, True)
    line = {'metric': MODEL['metric'], 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': max(args.warmup, 3), 'ms_per_step': ms / args.steps, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32 (bf16x3 split operands, fp32 accumulate)',
            'data': 'synthetic', 'config': config_dict(B, world),
            'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': B * L * 4 * world,
                    'd2h_bytes_per_step': int(host_out.numel()) * 4 * world, 'ms_per_step': ms_e2e / args.steps},
            'gpu_launches': int(launches), 'clocks': clocks, 'roofline': roofline, 'stages': stages,
            'cpu_baseline': {'value': cpu_val, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                             'sample': f'2 x 4 s utterances, median of 5 ({cpu_sec:.2f} s each); CPU oracle = '
                                       'torch-CPU restatement of the reference path on the host cores, flush-denormal on'}}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--batch', type=int, default=int(os.environ.get('SE_BENCH_BATCH', '64')),
                    help='4 s utterances per GPU per step')
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--model', default=os.environ.get('SE_BENCH_MODEL', 'dccrn'), choices=sorted(MODELS))
    args = ap.parse_args()
    global MODEL
    MODEL = MODELS[args.model]
    if MODEL.get('train'):
        return run_train(args)
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
