mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest5.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest5.log
python bench.py --no-cpu-baseline > gpurun_out/b5_cfg2.json 2> gpurun_out/b5_cfg2.err
python bench.py --no-cpu-baseline --precision bf16 > gpurun_out/b5_cfg2_bf16.json 2> gpurun_out/b5_cfg2_bf16.err
python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b5_cfg5_1024.json 2> gpurun_out/b5_cfg5_1024.err
echo done
