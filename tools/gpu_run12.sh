mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest12.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest12.log
python tools/trace_decode.py cfg2 bf16w > gpurun_out/trace12_staged.txt 2>&1
ECHR_ATT_STREAM=1 python tools/trace_decode.py cfg2 bf16w > gpurun_out/trace12_stream.txt 2>&1
python bench.py --no-cpu-baseline > gpurun_out/b12_staged.json 2> gpurun_out/b12_staged.err
ECHR_ATT_STREAM=1 python bench.py --no-cpu-baseline > gpurun_out/b12_stream.json 2> gpurun_out/b12_stream.err
echo done
