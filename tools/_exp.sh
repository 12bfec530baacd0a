mkdir -p gpurun_out/r02zD
for bv in 820 936 1024 1366 2048; do python bench.py --batch-videos $bv --steps 10 --no-cpu-baseline --no-second-beam > gpurun_out/r02zD/bv$bv.json 2>/dev/null; done
