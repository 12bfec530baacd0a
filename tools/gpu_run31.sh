mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest31.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest31.log
for mask in 30 28 22 14 12 20 6; do
ECHR_PDL_MASK=$mask timeout 300 python bench.py --no-cpu-baseline --steps 10 > gpurun_out/b31_m$mask.json 2> gpurun_out/b31_m$mask.err
done
ECHR_TRACE=1 timeout 200 python tools/trace_decode.py > gpurun_out/trace31.txt 2>&1
timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b31_cfg5.json 2> gpurun_out/b31_cfg5.err
ECHR_PDL_MASK=0 timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b31_cfg5_m0.json 2> gpurun_out/b31_cfg5_m0.err
timeout 300 python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b31_cfg3.json 2> gpurun_out/b31_cfg3.err
ECHR_PDL_MASK=0 timeout 300 python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/b31_cfg3_m0.json 2> gpurun_out/b31_cfg3_m0.err
echo done
