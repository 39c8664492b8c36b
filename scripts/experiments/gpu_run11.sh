mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --batch 16 > gpurun_out/r11_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv --log-file gpurun_out/r11_launches.csv python bench.py --steps 2 --warmup 3 --batch 16 > gpurun_out/r11_ncu.log 2>&1
tail -2 gpurun_out/r11_ncu.log | cut -c1-300
