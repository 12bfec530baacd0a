mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest4.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest4.log
python bench.py --no-cpu-baseline > gpurun_out/b4_cfg2.json 2> gpurun_out/b4_cfg2.err
python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b4_cfg5_1024.json 2> gpurun_out/b4_cfg5_1024.err
python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b4_cfg3.json 2> gpurun_out/b4_cfg3.err
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain4.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 300 --csv --log-file gpurun_out/launches4.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu4.log 2>&1
echo done
