mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest26.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest26.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b26_cfg2.json 2> gpurun_out/b26_cfg2.err
ECHR_ENCODE_GRAPH=0 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b26_cfg2_eager.json 2> gpurun_out/b26_cfg2_eager.err
timeout 300 python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b26_cfg3.json 2> gpurun_out/b26_cfg3.err
timeout 300 python bench.py --workload cfg4 --steps 5 --no-cpu-baseline > gpurun_out/b26_cfg4.json 2> gpurun_out/b26_cfg4.err
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b26_cfg5.json 2> gpurun_out/b26_cfg5.err
echo done
