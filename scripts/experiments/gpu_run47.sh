#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r53_tests.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/r53_tests.log | cut -c1-220
timeout 300 python scripts/chain_layers.py --model dccrn 2>&1 | grep -E "^ 1 |^11 |TOTAL"
timeout 300 python scripts/chain_layers.py --model dcunet20 2>&1 | grep -E "^11 |TOTAL"
timeout 600 python bench.py --steps 10 --warmup 3 --model dccrn > gpurun_out/r53_bench.log 2>&1
echo "bench rc=$?"; tail -1 gpurun_out/r53_bench.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['clocks']['sm_mhz'])" || tail -5 gpurun_out/r53_bench.log
