#!/bin/bash
mkdir -p gpurun_out
for f in default "16,2,0" "32,2,0" "64,2,0"; do
  for m in frcrn dcunet20; do
    if [ "$f" = default ]; then timeout 300 python scripts/chain_layers.py --model $m > gpurun_out/r56_${m}_default.txt 2>&1
    else SE_GEMM_FORCE="$f" timeout 300 python scripts/chain_layers.py --model $m > gpurun_out/r56_${m}_$f.txt 2>&1; fi
    echo "$f $m $(grep TOTAL gpurun_out/r56_${m}_$f.txt)"
  done
done
