mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest48.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest48.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b48_cfg2.json 2> gpurun_out/b48_cfg2.err
ECHR_ARGMAX_KEYS=0 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b48_cfg2_nokeys.json 2> gpurun_out/b48_cfg2_nokeys.err
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b48_cfg5.json 2> gpurun_out/b48_cfg5.err
timeout 300 python bench.py --workload cfg1 --no-cpu-baseline > gpurun_out/b48_cfg1.json 2> gpurun_out/b48_cfg1.err
timeout 600 python tools/stress_determinism.py 150 > gpurun_out/stress48.log 2>&1
ECHR_TRACE=1 timeout 200 python tools/trace_decode.py > gpurun_out/trace48.txt 2>&1
timeout 200 python tools/time_forward.py fp32 > gpurun_out/forward48.txt 2>&1
echo done
