mkdir -p gpurun_out
for mask in 23 19 7 3 22; do
ECHR_PDL_MASK=$mask timeout 300 python bench.py --no-cpu-baseline --steps 10 > gpurun_out/b49_k_m$mask.json 2> gpurun_out/b49_k_m$mask.err
ECHR_ARGMAX_KEYS=0 ECHR_PDL_MASK=$mask timeout 300 python bench.py --no-cpu-baseline --steps 10 > gpurun_out/b49_n_m$mask.json 2> gpurun_out/b49_n_m$mask.err
done
ECHR_PDL_MASK=7 ECHR_TRACE=1 timeout 200 python tools/trace_decode.py > gpurun_out/trace49_m7.txt 2>&1
echo done
