#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "col2im or fused_chain or dccrn or dcunet" > gpurun_out/r46_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r46_tests.log | cut -c1-220
for g in 1 0; do
  SE_COL2IM_GATHER=$g timeout 600 python bench.py --steps 10 --warmup 3 --model dccrn > gpurun_out/r46_bench_g$g.log 2>&1
  echo "bench gather=$g rc=$?"; tail -1 gpurun_out/r46_bench_g$g.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['clocks']['sm_mhz'])" || tail -5 gpurun_out/r46_bench_g$g.log
  SE_COL2IM_GATHER=$g timeout 300 python scripts/chain_layers.py --model dccrn 2>&1 | tail -3
  SE_COL2IM_GATHER=$g timeout 300 python scripts/chain_layers.py --model dcunet20 2>&1 | tail -2
done
