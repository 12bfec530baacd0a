mkdir -p gpurun_out
python bench.py --workload cfg3 --steps 3 --no-cpu-baseline > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 40 -c 5 -o gpurun_out/prof_gemm -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_gemm.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:attention_kernel -s 12 -c 1 -o gpurun_out/prof_attn -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_attn.log 2>&1
echo done
