set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r1_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1_launches.csv python bench.py --steps 5 --warmup 3 > gpurun_out/r1_ncu.log 2>&1
tail -3 gpurun_out/r1_pytest.log; tail -2 gpurun_out/r1_bench.log
