#!/bin/bash
# forced-plan sweep of the chained DCCRN layers: per-layer ms for every (TW, nacc, mode)
out=gpurun_out/plan_sweep_dccrn.txt
: > $out
echo "### auto" >> $out; python scripts/chain_layers.py --model dccrn --iters 3 2>/dev/null | grep -E "^ ?[0-9]+ |TOTAL" >> $out
for mode in 0 1; do for tw in 16 32 64 128; do for nacc in 1 2 4; do
  echo "### force $tw,$nacc,$mode" >> $out
  SE_GEMM_FORCE=$tw,$nacc,$mode python scripts/chain_layers.py --model dccrn --iters 3 2>/dev/null | grep -E "^ ?[0-9]+ |TOTAL" >> $out
done; done; done
