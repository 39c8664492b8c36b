"""The driver's reference arm (`bench.py --impl reference`) on the CPU: one JSON line with the contract's keys, timing the
staged, unmodified reference (`kind: "reference"`) when oracle/_ref exists, the oracle port otherwise."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_contract_line():
    res = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--steps', '1', '--warmup', '1'],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [l for l in res.stdout.splitlines() if l.startswith('{')]
    assert len(lines) == 1
    d = json.loads(lines[0])
    from oracle import ref_loader
    assert d['impl'] == 'reference' and d['metric'] == 'dccrn_cl_audio_seconds_enhanced_per_second'
    assert d['unit'] == 'audio-s/s' and d['higher_is_better'] is True and d['value'] > 0
    assert d['cpu_baseline']['kind'] == ('reference' if ref_loader.available() else 'port')
    assert d['cpu_baseline']['cores'] >= 1 and d['cpu_baseline']['value'] == d['value']
    assert d['cpu_baseline']['batch1']['value'] > 0 and d['cpu_baseline']['batch8']['value'] > 0
    assert d['e2e'] == {'value': d['value'], 'unit': 'audio-s/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}
    assert d['config']['per_gpu_batch'] == 64 and 'DCCRN-CL' in d['config']['workload']
    assert d['gpu_launches'] == 0


def test_reference_arm_is_silent_on_other_ranks():
    env = dict(os.environ, RANK='1', WORLD_SIZE='2', LOCAL_RANK='1')
    res = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--gpus', '2'],
                         capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert res.returncode == 0 and res.stdout.strip() == ''
