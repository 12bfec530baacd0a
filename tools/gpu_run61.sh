mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -x -p no:cacheprovider -s > gpurun_out/r61_pytest.log 2>&1; echo "pytest exit=$?" >> gpurun_out/r61_pytest.log
timeout 600 python tools/compare_precisions.py f16 cfg2 > gpurun_out/r61_cmp_f16.log 2>&1
timeout 600 python tools/compare_precisions.py bf16w cfg2 > gpurun_out/r61_cmp_bf16w.log 2>&1
timeout 300 python bench.py --precision f16 --no-cpu-baseline > gpurun_out/r61_cfg2_f16.json 2> gpurun_out/r61_cfg2_f16.err
timeout 300 python bench.py --precision f16 --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/r61_cfg5_f16.json 2> gpurun_out/r61_cfg5_f16.err
timeout 300 python bench.py --precision f16 --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/r61_cfg3_f16.json 2> gpurun_out/r61_cfg3_f16.err
