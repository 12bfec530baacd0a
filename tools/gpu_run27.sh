mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest27.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest27.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b27_cfg2.json 2> gpurun_out/b27_cfg2.err
ECHR_ATT_V2_THREADS=512 timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b27_cfg2_t512.json 2> gpurun_out/b27_cfg2_t512.err
ECHR_TRACE=1 timeout 200 python tools/trace_decode.py > gpurun_out/trace27.txt 2>&1
ECHR_ATT_V2_THREADS=512 ECHR_TRACE=1 timeout 200 python tools/trace_decode.py > gpurun_out/trace27_t512.txt 2>&1
echo done
