#!/bin/bash
mkdir -p gpurun_out
for pen in 1.3 1.6; do
for m in dccrn dcunet20 frcrn; do
  SE_GEMM_SPLIT_PENALTY=$pen SE_GEMM_DEBUG=4 timeout 300 python scripts/chain_layers.py --model $m > gpurun_out/r58_${m}_$pen.txt 2>&1
  echo "pen=$pen $m $(grep TOTAL gpurun_out/r58_${m}_$pen.txt)"
  grep "N=128" gpurun_out/r58_${m}_$pen.txt | sed 's/.*-> //' | cut -d' ' -f1-4 | sort | uniq -c | sort -rn | head -4
done; done
grep -E "^ 8 " gpurun_out/r58_dccrn_1.3.txt gpurun_out/r58_dccrn_1.6.txt | cut -c1-160
