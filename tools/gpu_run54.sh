mkdir -p gpurun_out
ECHR_GEMM2_MIN_ROWS=1000000 timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b54_cfg5_nop.json 2> gpurun_out/b54_cfg5_nop.err
ECHR_GEMM2_MIN_ROWS=1000000 ECHR_ATT_V2_ROWS=100000 timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b54_cfg5_nop_v2.json 2> gpurun_out/b54_cfg5_nop_v2.err
ECHR_GEMM2_MIN_ROWS=1000000 ECHR_PDL_MASK=0 timeout 300 python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b54_cfg5_nop_nopdl.json 2> gpurun_out/b54_cfg5_nop_nopdl.err
ECHR_GEMM2_MIN_ROWS=1000000 ECHR_TRACE=1 timeout 300 python tools/trace_decode.py cfg5 bf16w 1024 > gpurun_out/trace54_cfg5_nop.txt 2>&1
echo done
