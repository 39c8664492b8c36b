#!/bin/bash
for f in "16,2,0" "32,2,0" "64,1,0" "64,2,0" "128,1,0" "32,1,1" "64,2,1"; do
  echo "== SE_GEMM_FORCE=$f"
  SE_GEMM_FORCE=$f timeout 300 python scripts/chain_layers.py --model dcunet20 2>&1 | grep -E "^ (1|2) |^19 "
done
echo "== default"; timeout 300 python scripts/chain_layers.py --model dcunet20 2>&1 | grep -E "^ (1|2) |^19 "
