#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "first or small_cin or fused_chain or dccrn or dcunet or frcrn or conv" > gpurun_out/r47_tests.log 2>&1
echo "tests rc=$?"; tail -5 gpurun_out/r47_tests.log | cut -c1-220
for m in dccrn dcunet20 frcrn; do timeout 300 python scripts/chain_layers.py --model $m 2>&1 | grep -E "^ 0 |TOTAL"; done
timeout 600 python bench.py --steps 10 --warmup 3 --model dccrn > gpurun_out/r47_bench.log 2>&1
echo "bench rc=$?"; tail -1 gpurun_out/r47_bench.log | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['clocks']['sm_mhz'])" || tail -5 gpurun_out/r47_bench.log
