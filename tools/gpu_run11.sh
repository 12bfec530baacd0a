mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain11.log 2>&1 && \
ncu --set full --cache-control none --clock-control none --import-source on -k regex:attention -s 40 -c 2 -o gpurun_out/prof_attn_warm -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_attn_warm.log 2>&1 && \
ncu --set full --cache-control none --clock-control none --import-source on -k regex:gemm_tc_kernel -s 100 -c 3 -o gpurun_out/prof_gemm_warm -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_gemm_warm.log 2>&1
echo done
