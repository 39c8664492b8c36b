mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --deselect "tests/test_gpu_parity.py::test_models_vs_reference_golden[frcrn-FRCRN-args4]" 2>&1 | tail -5
SE_GEMM_DEBUG=4 python scripts/layer_bench.py --model dccrn --batch 16 --iters 1 2>&1 | grep -E "se_cconv_gemm" | sort | uniq | cut -c1-200
python scripts/layer_bench.py --model dccrn --batch 16
python bench.py --steps 5 --warmup 3 --batch 64 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], json.dumps(d['stages']))"
