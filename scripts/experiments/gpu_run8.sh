mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -k "cconv" 2>&1 | tail -5
echo "== plan"; SE_GEMM_DEBUG=4 python scripts/layer_bench.py --model dccrn --batch 16 --iters 1 2>&1 | grep -E "se_cconv_gemm" | sort | uniq
python scripts/layer_bench.py --model dccrn --batch 16
python scripts/layer_bench.py --model dcunet --batch 16 --iters 5
python scripts/layer_bench.py --model frcrn --batch 16 --iters 5
