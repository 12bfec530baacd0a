mkdir -p gpurun_out
for m in 0 27 24 16 8 3 11 19 31; do
ECHR_PDL_MASK=$m python bench.py --no-cpu-baseline --steps 20 > gpurun_out/b8_m$m.json 2> gpurun_out/b8_m$m.err
done
echo done
