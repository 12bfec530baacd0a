mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest57.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest57.log
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b57_cfg5.json 2> gpurun_out/b57_cfg5.err
ECHR_HOIST_CTX=1 timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b57_cfg5_hoist.json 2> gpurun_out/b57_cfg5_hoist.err
timeout 300 python bench.py --workload cfg4 --steps 5 --no-cpu-baseline > gpurun_out/b57_cfg4.json 2> gpurun_out/b57_cfg4.err
timeout 300 python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b57_cfg3.json 2> gpurun_out/b57_cfg3.err
ECHR_HOIST_CTX=0 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b57_cfg2_nohoist.json 2> gpurun_out/b57_cfg2_nohoist.err
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b57_cfg2.json 2> gpurun_out/b57_cfg2.err
echo done
