#!/bin/bash
mkdir -p gpurun_out
SE_GEMM_DEBUG=4 SE_GEMM_PROF=1 timeout 300 python scripts/chain_layers.py --model dcunet20 > gpurun_out/r52_dcunet_prof.log 2>&1
grep -E "k=7x1|k=1x7" gpurun_out/r52_dcunet_prof.log | sort | uniq -c | sort -rn | head -20
SE_GEMM_DEBUG=4 SE_GEMM_PROF=1 timeout 300 python scripts/chain_layers.py --model dccrn > gpurun_out/r52_dccrn_prof.log 2>&1
grep -E "se_cconv_gemm" gpurun_out/r52_dccrn_prof.log | tail -26
