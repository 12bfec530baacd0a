mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest13.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest13.log
python tools/trace_decode.py cfg2 bf16w > gpurun_out/trace13.txt 2>&1
python bench.py --no-cpu-baseline > gpurun_out/b13.json 2> gpurun_out/b13.err
echo done
