mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gemm_gpu.py -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest17a.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest17a.log
ECHR_GEMM2_MIN_ROWS=1 timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest17b.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest17b.log
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest17c.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest17c.log
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b17_cfg5.json 2> gpurun_out/b17_cfg5.err
ECHR_GEMM2_MIN_ROWS=100000000 timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b17_cfg5_v1.json 2> gpurun_out/b17_cfg5_v1.err
timeout 300 python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b17_cfg3.json 2> gpurun_out/b17_cfg3.err
timeout 300 python bench.py --workload cfg4 --steps 3 --no-cpu-baseline > gpurun_out/b17_cfg4.json 2> gpurun_out/b17_cfg4.err
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b17_cfg2.json 2> gpurun_out/b17_cfg2.err
echo done
