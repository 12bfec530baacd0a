mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest51.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest51.log
timeout 200 python tools/time_forward.py fp32 > gpurun_out/forward51.txt 2>&1
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/b51_cfg2.json 2> gpurun_out/b51_cfg2.err
echo done
