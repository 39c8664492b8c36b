mkdir -p gpurun_out
L="enc1,enc2,enc3,dec1,dec2,dec3,dec4"
echo "== plan"; SE_GEMM_DEBUG=4 python scripts/layer_bench.py --model dccrn --batch 16 --iters 1 --only $L 2>&1 | grep -E "se_cconv_gemm" | sort | uniq
echo "== baseline"; python scripts/layer_bench.py --model dccrn --batch 16 --only $L
echo "== no loads (MMA rate)"; SE_GEMM_DEBUG=1 python scripts/layer_bench.py --model dccrn --batch 16 --only $L
echo "== no MMA (fill rate)"; SE_GEMM_DEBUG=2 python scripts/layer_bench.py --model dccrn --batch 16 --only $L
for f in "32,1,32" "32,2,32" "64,2,32" "128,1,32" "32,1,64" "64,2,64" "32,4,64" "32,4,32" "16,4,32"; do echo "== force $f"; SE_GEMM_FORCE=$f python scripts/layer_bench.py --model dccrn --batch 16 --only $L; done
