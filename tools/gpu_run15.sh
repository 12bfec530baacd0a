mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest15.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest15.log
python tools/trace_decode.py cfg2 bf16w > gpurun_out/trace15.txt 2>&1
python bench.py --no-cpu-baseline > gpurun_out/b15.json 2> gpurun_out/b15.err
python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b15_cfg5.json 2> gpurun_out/b15_cfg5.err
python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b15_cfg3.json 2> gpurun_out/b15_cfg3.err
python bench.py --workload cfg4 --steps 3 --no-cpu-baseline > gpurun_out/b15_cfg4.json 2> gpurun_out/b15_cfg4.err
echo done
