mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest23.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest23.log
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/b23_cfg2.json 2> gpurun_out/b23_cfg2.err
ECHR_ATT_V2_ROWS=0 timeout 400 python bench.py --no-cpu-baseline > gpurun_out/b23_cfg2_v1.json 2> gpurun_out/b23_cfg2_v1.err
ECHR_TRACE=1 timeout 200 python tools/trace_decode.py > gpurun_out/trace23.txt 2>&1
echo done
