#!/bin/bash
mkdir -p gpurun_out
out=gpurun_out/r54_force_sweep_dccrn.txt; : > $out
echo "== default" >> $out; timeout 300 python scripts/chain_layers.py --model dccrn 2>&1 | grep -E "^ *[0-9]+ conv|TOTAL" >> $out
for tw in 128 64 32 16; do for n in 1 2 4; do for m in 0 1; do
  echo "== $tw,$n,$m" >> $out
  SE_GEMM_FORCE="$tw,$n,$m" timeout 300 python scripts/chain_layers.py --model dccrn 2>&1 | grep -E "^ *[0-9]+ conv|TOTAL" >> $out
done; done; done
wc -l $out
