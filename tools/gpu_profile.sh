# Round profile set, part 1: plain bench run, then the ncu launch list of the same command.
# (warm-up 2: a shape runs eagerly on first sight and is captured into its graph on the second.)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 2 --no-cpu-baseline"
$CMD > gpurun_out/prof_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/prof_launches.csv $CMD > gpurun_out/prof_ncu1.log 2>&1
echo done
