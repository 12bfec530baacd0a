mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain25.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches25.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu25.log 2>&1
echo done
