mkdir -p gpurun_out
for mask in 0 8 24 26 2 10 30 31; do
ECHR_PDL_MASK=$mask timeout 300 python bench.py --no-cpu-baseline --steps 10 > gpurun_out/b30_m$mask.json 2> gpurun_out/b30_m$mask.err
done
ECHR_PDL_MASK=26 ECHR_PDL_LATE=0 timeout 300 python bench.py --no-cpu-baseline --steps 10 > gpurun_out/b30_m26_early.json 2> gpurun_out/b30_m26_early.err
ECHR_PDL_MASK=26 ECHR_TRACE=1 timeout 200 python tools/trace_decode.py > gpurun_out/trace30_m26.txt 2>&1
echo done
