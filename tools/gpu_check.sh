# Standard round check on one B200: GPU parity tests, default bench line, the other BASELINE configs, determinism.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/check_pytest.log 2>&1; echo "pytest exit=$?" >> gpurun_out/check_pytest.log
timeout 600 python bench.py > gpurun_out/check_cfg2.json 2> gpurun_out/check_cfg2.err
timeout 300 python bench.py --workload cfg1 --no-cpu-baseline > gpurun_out/check_cfg1.json 2> gpurun_out/check_cfg1.err
timeout 300 python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/check_cfg3.json 2> gpurun_out/check_cfg3.err
timeout 300 python bench.py --workload cfg4 --steps 5 --no-cpu-baseline > gpurun_out/check_cfg4.json 2> gpurun_out/check_cfg4.err
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/check_cfg5.json 2> gpurun_out/check_cfg5.err
timeout 300 python bench.py --workload cfg5 --videos 1024 --beam 3 --steps 3 --no-cpu-baseline > gpurun_out/check_cfg5b3.json 2> gpurun_out/check_cfg5b3.err
timeout 600 python tools/stress_determinism.py 100 > gpurun_out/check_stress.log 2>&1
ECHR_TRACE=1 timeout 200 python tools/trace_decode.py > gpurun_out/check_trace_cfg2.txt 2>&1
echo done
