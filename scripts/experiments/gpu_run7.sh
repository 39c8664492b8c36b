L="dec3,dec4"
for dbg in 1 0; do for f in "32,2,32" "32,2,64" "32,1,32" "32,1,64"; do echo "== dbg $dbg force $f"; SE_GEMM_DEBUG=$dbg SE_GEMM_FORCE=$f python scripts/layer_bench.py --model dccrn --batch 16 --only $L; done; done
