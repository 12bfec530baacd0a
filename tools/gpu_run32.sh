mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest32.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest32.log
for mask in 23 31 22 30 7 15; do
ECHR_PDL_MASK=$mask timeout 300 python bench.py --no-cpu-baseline --steps 10 > gpurun_out/b32_m$mask.json 2> gpurun_out/b32_m$mask.err
done
ECHR_CRIT_MAIN=0 timeout 300 python bench.py --no-cpu-baseline --steps 10 > gpurun_out/b32_old.json 2> gpurun_out/b32_old.err
ECHR_PDL_MASK=23 ECHR_TRACE=1 timeout 200 python tools/trace_decode.py > gpurun_out/trace32.txt 2>&1
echo done
