mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider > gpurun_out/pytest14.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest14.log
python bench.py --no-cpu-baseline > gpurun_out/b14.json 2> gpurun_out/b14.err
echo done
