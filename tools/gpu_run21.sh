mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest21.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest21.log
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/b21_cfg2.json 2> gpurun_out/b21_cfg2.err
ECHR_H2ATT_BN=64 timeout 400 python bench.py --no-cpu-baseline > gpurun_out/b21_cfg2_bn64.json 2> gpurun_out/b21_cfg2_bn64.err
ECHR_H2ATT_BN=128 timeout 400 python bench.py --no-cpu-baseline > gpurun_out/b21_cfg2_bn128.json 2> gpurun_out/b21_cfg2_bn128.err
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b21_cfg5.json 2> gpurun_out/b21_cfg5.err
timeout 300 python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b21_cfg3.json 2> gpurun_out/b21_cfg3.err
ECHR_TRACE=1 timeout 200 python tools/trace_decode.py > gpurun_out/trace21.txt 2>&1
echo done
