#!/bin/bash
# 2-GPU check of the default bench (utterance-sharded inference) and the reference arm under torchrun
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r33_bench_2gpu.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/r33_bench_2gpu.log | cut -c1-700
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r33_ref_2gpu.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/r33_ref_2gpu.log | cut -c1-500
