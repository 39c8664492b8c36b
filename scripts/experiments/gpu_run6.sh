mkdir -p gpurun_out
python scripts/layer_bench.py --model dccrn --batch 16 --iters 1 --only dec1,dec3 > gpurun_out/r6_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:cconv_gemm_kernel -s 2 -c 1 -o gpurun_out/r6_dec1 python scripts/layer_bench.py --model dccrn --batch 16 --iters 1 --only dec1 > gpurun_out/r6_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:cconv_gemm_kernel -s 2 -c 1 -o gpurun_out/r6_dec3 python scripts/layer_bench.py --model dccrn --batch 16 --iters 1 --only dec3 > gpurun_out/r6_ncu3.log 2>&1
cat gpurun_out/r6_plain.log
