mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest46.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest46.log
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b46_cfg5.json 2> gpurun_out/b46_cfg5.err
ECHR_CRIT_MAIN=0 timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b46_cfg5_old.json 2> gpurun_out/b46_cfg5_old.err
timeout 300 python bench.py --workload cfg4 --steps 5 --no-cpu-baseline > gpurun_out/b46_cfg4.json 2> gpurun_out/b46_cfg4.err
timeout 300 python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b46_cfg3.json 2> gpurun_out/b46_cfg3.err
timeout 300 python bench.py --workload cfg5 --videos 1024 --beam 3 --steps 3 --no-cpu-baseline > gpurun_out/b46_cfg5b3.json 2> gpurun_out/b46_cfg5b3.err
echo done
