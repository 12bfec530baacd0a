mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest19.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest19.log
timeout 300 python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b19_cfg3.json 2> gpurun_out/b19_cfg3.err
echo done
