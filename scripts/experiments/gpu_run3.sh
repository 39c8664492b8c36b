set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r3_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r3_pytest.log
tail -15 gpurun_out/r3_pytest.log
timeout 300 python scripts/layer_bench.py --model dccrn --batch 16 > gpurun_out/r3_layers_dccrn_b16.log 2>&1
cat gpurun_out/r3_layers_dccrn_b16.log
timeout 300 python scripts/layer_bench.py --model dcunet --batch 16 --iters 5 > gpurun_out/r3_layers_dcunet_b16.log 2>&1
timeout 300 python scripts/layer_bench.py --model frcrn --batch 16 --iters 5 > gpurun_out/r3_layers_frcrn_b16.log 2>&1
cat gpurun_out/r3_layers_dcunet_b16.log gpurun_out/r3_layers_frcrn_b16.log
