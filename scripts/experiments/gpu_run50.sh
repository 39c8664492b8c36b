#!/bin/bash
for o in 800 1500 2500 4000; do
  echo "== SE_GEMM_TILE_OVH=$o"
  for m in dccrn dcunet20 frcrn; do SE_GEMM_TILE_OVH=$o timeout 300 python scripts/chain_layers.py --model $m 2>&1 | grep -E "TOTAL"; done
  SE_GEMM_TILE_OVH=$o timeout 300 python scripts/chain_layers.py --model dcunet20 2>&1 | grep -E "^ (1|2) "
done
