# Round profile set, part 2: ncu --set full of the decode-step kernels inside a steady-state step.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 2 --no-cpu-baseline"
$CMD > gpurun_out/prof_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gemm_tc_kernel|attention_v2" -s 400 -c 8 -o gpurun_out/prof_step -f $CMD > gpurun_out/prof_ncu2.log 2>&1
echo done
