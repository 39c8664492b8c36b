set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest.log
python scripts/layer_bench.py --model dccrn --batch 16 > gpurun_out/r2_layers_dccrn_b16.log 2>&1
python scripts/layer_bench.py --model dccrn --batch 64 --iters 5 > gpurun_out/r2_layers_dccrn_b64.log 2>&1
python scripts/layer_bench.py --model dcunet --batch 16 --iters 5 > gpurun_out/r2_layers_dcunet_b16.log 2>&1
python scripts/layer_bench.py --model frcrn --batch 16 --iters 5 > gpurun_out/r2_layers_frcrn_b16.log 2>&1
python bench.py --steps 5 --warmup 3 --batch 64 > gpurun_out/r2_bench_b64.log 2>&1
python scripts/layer_bench.py --model dccrn --batch 16 --iters 1 --only enc1,enc3,dec1,dec3 > gpurun_out/r2_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:cconv_umma_kernel -s 8 -c 4 -o gpurun_out/r2_conv_full python scripts/layer_bench.py --model dccrn --batch 16 --iters 1 --only enc1,enc3,dec1,dec3 > gpurun_out/r2_ncu.log 2>&1
tail -3 gpurun_out/r2_pytest.log; cat gpurun_out/r2_layers_dccrn_b16.log; tail -1 gpurun_out/r2_bench_b64.log | cut -c1-600
