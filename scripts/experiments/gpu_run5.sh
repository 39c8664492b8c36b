mkdir -p gpurun_out
L="enc2,enc3,dec1,dec2,dec3,dec4"
for m in 0 8 16; do
echo "== mode $m"; SE_GEMM_DEBUG=$m python scripts/layer_bench.py --model dccrn --batch 16 --only $L
SE_GEMM_DEBUG=$m timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -k "layer_shapes or golden_umma" 2>&1 | tail -3
done
