mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain22.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:attention_kernel -s 20 -c 2 -o gpurun_out/prof22_att -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu22.log 2>&1
echo done
