mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest20.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest20.log
timeout 400 python bench.py > gpurun_out/b20_cfg2.json 2> gpurun_out/b20_cfg2.err
timeout 300 python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b20_cfg3.json 2> gpurun_out/b20_cfg3.err
timeout 300 python bench.py --workload cfg4 --steps 5 --no-cpu-baseline > gpurun_out/b20_cfg4.json 2> gpurun_out/b20_cfg4.err
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b20_cfg5.json 2> gpurun_out/b20_cfg5.err
timeout 300 python bench.py --workload cfg5 --videos 1024 --beam 3 --steps 3 --no-cpu-baseline > gpurun_out/b20_cfg5b3.json 2> gpurun_out/b20_cfg5b3.err
ECHR_TRACE=1 timeout 200 python tools/trace_decode.py > gpurun_out/trace20.txt 2>&1
echo done
