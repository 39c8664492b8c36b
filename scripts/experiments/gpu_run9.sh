mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python scripts/layer_bench.py --model dccrn --batch 16
python scripts/layer_bench.py --model dcunet --batch 16 --iters 5 --only first,dec9
python scripts/layer_bench.py --model frcrn --batch 16 --iters 5 --only enc0
