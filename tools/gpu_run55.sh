mkdir -p gpurun_out
ECHR_ATT_V2_ROWS=1000000 timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b55_cfg5_v2.json 2> gpurun_out/b55_cfg5_v2.err
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b55_cfg5.json 2> gpurun_out/b55_cfg5.err
ECHR_ATT_V2_ROWS=1000000 timeout 300 python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b55_cfg3_v2.json 2> gpurun_out/b55_cfg3_v2.err
timeout 300 python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b55_cfg3.json 2> gpurun_out/b55_cfg3.err
ECHR_ATT_V2_ROWS=1000000 timeout 300 python bench.py --workload cfg5 --videos 1024 --beam 3 --steps 3 --no-cpu-baseline > gpurun_out/b55_cfg5b3_v2.json 2> gpurun_out/b55_cfg5b3_v2.err
echo done
