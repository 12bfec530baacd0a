mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest44.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest44.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b44_cfg2.json 2> gpurun_out/b44_cfg2.err
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b44_cfg5.json 2> gpurun_out/b44_cfg5.err
timeout 600 python tools/stress_determinism.py 150 > gpurun_out/stress44.log 2>&1
echo done
