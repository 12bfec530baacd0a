mkdir -p gpurun_out
ECHR_ATT_V2_MAXLEN=1024 timeout 300 python bench.py --workload cfg4 --steps 5 --no-cpu-baseline > gpurun_out/b56_cfg4_v2.json 2> gpurun_out/b56_cfg4_v2.err
timeout 300 python bench.py --workload cfg4 --steps 5 --no-cpu-baseline > gpurun_out/b56_cfg4.json 2> gpurun_out/b56_cfg4.err
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b56_cfg2.json 2> gpurun_out/b56_cfg2.err
echo done
