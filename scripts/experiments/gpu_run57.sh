#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q -k "lstm or dccrn or frcrn or pipeline" > gpurun_out/r59_tests.log 2>&1
echo "tests rc=$?"; tail -3 gpurun_out/r59_tests.log | cut -c1-200
