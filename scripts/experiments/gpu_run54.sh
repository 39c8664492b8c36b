#!/bin/bash
mkdir -p gpurun_out
for m in dccrn dcunet20 frcrn; do
  SE_GEMM_DEBUG=4 timeout 300 python scripts/chain_layers.py --model $m > gpurun_out/r57_$m.txt 2>&1
  grep -E "^ *[0-9]+ conv.*cout2= *128|TOTAL" gpurun_out/r57_$m.txt | cut -c1-140
  grep "N=128" gpurun_out/r57_$m.txt | sed 's/.*-> //' | sort | uniq -c | head -8
done
