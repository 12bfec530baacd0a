mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest2.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest2.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?" >> gpurun_out/smoke.log
python bench.py > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err; echo "rc=$?" >> gpurun_out/bench_cfg2.err
python bench.py --precision fp32 --no-cpu-baseline > gpurun_out/bench_cfg2_fp32.json 2> gpurun_out/bench_cfg2_fp32.err
python bench.py --precision bf16 --no-cpu-baseline > gpurun_out/bench_cfg2_bf16.json 2> gpurun_out/bench_cfg2_bf16.err
python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/bench_cfg5_1024.json 2> gpurun_out/bench_cfg5_1024.err
python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err
python bench.py --workload cfg4 --steps 3 --no-cpu-baseline > gpurun_out/bench_cfg4.json 2> gpurun_out/bench_cfg4.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
echo done
