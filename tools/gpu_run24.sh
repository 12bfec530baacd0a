mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest24.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest24.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b24_cfg2.json 2> gpurun_out/b24_cfg2.err
ECHR_KSPLIT=1 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b24_cfg2_k1.json 2> gpurun_out/b24_cfg2_k1.err
ECHR_KSPLIT=2 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b24_cfg2_k2.json 2> gpurun_out/b24_cfg2_k2.err
ECHR_KSPLIT=0 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b24_cfg2_k0.json 2> gpurun_out/b24_cfg2_k0.err
ECHR_TRACE=1 timeout 200 python tools/trace_decode.py > gpurun_out/trace24.txt 2>&1
echo done
