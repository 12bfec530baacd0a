# Round profile set: plain bench run, ncu launch list of the same command, ncu --set full of the step kernels.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/prof_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/prof_launches.csv $CMD > gpurun_out/prof_ncu1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gemm_tc_kernel|attention_v2" -s 72 -c 8 -o gpurun_out/prof_step -f $CMD > gpurun_out/prof_ncu2.log 2>&1
echo done
