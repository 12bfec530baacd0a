mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest28.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest28.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b28_cfg2.json 2> gpurun_out/b28_cfg2.err
timeout 300 python bench.py --workload cfg1 --no-cpu-baseline > gpurun_out/b28_cfg1.json 2> gpurun_out/b28_cfg1.err
ECHR_ATT_STREAM=1 timeout 300 python bench.py --workload cfg1 --no-cpu-baseline > gpurun_out/b28_cfg1_nostage.json 2> gpurun_out/b28_cfg1_nostage.err
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b28_cfg5.json 2> gpurun_out/b28_cfg5.err
echo done
