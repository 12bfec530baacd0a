mkdir -p gpurun_out
timeout 300 python bench.py --precision bf16 --no-cpu-baseline > gpurun_out/r60_cfg2_bf16.json 2> gpurun_out/r60_cfg2_bf16.err
timeout 300 python bench.py --precision bf16 --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/r60_cfg5_bf16.json 2> gpurun_out/r60_cfg5_bf16.err
timeout 300 python bench.py --precision bf16 --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/r60_cfg3_bf16.json 2> gpurun_out/r60_cfg3_bf16.err
