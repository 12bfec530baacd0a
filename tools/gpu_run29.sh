mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain29.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 60 -c 9 -o gpurun_out/prof29_gemm -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu29.log 2>&1
echo done
