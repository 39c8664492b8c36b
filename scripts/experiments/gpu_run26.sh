#!/bin/bash
# FFT engine v4: full parity suite, benches, config-5 sweep, ncu of the inverse
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r30_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/r30_tests.log
for b in 64 256; do timeout 300 python scripts/stft_bench.py --batch $b 2>&1 | tail -1; done
for m in dccrn dcunet20; do
  timeout 600 python bench.py --steps 10 --warmup 3 --model $m > gpurun_out/r30_bench_$m.log 2>&1
  echo "bench $m rc=$?"; tail -1 gpurun_out/r30_bench_$m.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], {k:(round(v['ms'],3), round(v.get('frac_hbm',0) or 0,3)) for k,v in d['stages'].items() if isinstance(v,dict) and 'ms' in v})"
done
timeout 900 python scripts/stft_sweep.py --engine auto > gpurun_out/r30_stft_sweep.txt 2>&1; tail -3 gpurun_out/r30_stft_sweep.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:istft_fft_kernel -s 8 -c 1 \
  -o gpurun_out/r30_fft_inv -f python scripts/stft_bench.py --batch 256 > gpurun_out/r30_ncu2.log 2>&1
