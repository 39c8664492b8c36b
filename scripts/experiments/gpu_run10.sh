mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --deselect "tests/test_gpu_parity.py::test_models_vs_reference_golden[frcrn-FRCRN-args4]" 2>&1 | tail -15
python bench.py --steps 5 --warmup 3 --batch 16 2>&1 | tail -1 | cut -c1-2500
python bench.py --steps 5 --warmup 3 --batch 64 2>&1 | tail -1 | cut -c1-2500
