mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest9.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest9.log
python bench.py --no-cpu-baseline > gpurun_out/b9_staged.json 2> gpurun_out/b9_staged.err
ECHR_ATT_STREAM=1 python bench.py --no-cpu-baseline > gpurun_out/b9_stream.json 2> gpurun_out/b9_stream.err
python bench.py --no-cpu-baseline --precision bf16 > gpurun_out/b9_bf16.json 2> gpurun_out/b9_bf16.err
echo done
