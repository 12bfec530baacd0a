mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -x -p no:cacheprovider > gpurun_out/r58_pytest.log 2>&1; echo "pytest exit=$?" >> gpurun_out/r58_pytest.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r58_cfg2.json 2> gpurun_out/r58_cfg2.err
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/r58_cfg5.json 2> gpurun_out/r58_cfg5.err
bash tools/gpu_profile.sh
