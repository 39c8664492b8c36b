#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r21_tests.log 2>&1
echo "tests rc=$?"; tail -12 gpurun_out/r21_tests.log
timeout 600 python scripts/layer_bench.py --model dcunet --batch 16 > gpurun_out/r21_layers_dcunet.log 2>&1; tail -3 gpurun_out/r21_layers_dcunet.log
