mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$T --master-port 29521 bench.py --gpus 8 --no-cpu-baseline > gpurun_out/scale8b_k20.json 2> gpurun_out/scale8b_k20.err
ECHR_BENCH_NO_GATHER=1 $T --master-port 29522 bench.py --gpus 8 --no-cpu-baseline > gpurun_out/scale8b_nogather.json 2> gpurun_out/scale8b_nogather.err
$T --master-port 29523 bench.py --gpus 8 --steps 100 --no-cpu-baseline > gpurun_out/scale8b_k100.json 2> gpurun_out/scale8b_k100.err
nproc > gpurun_out/scale8b_nproc.txt; uptime >> gpurun_out/scale8b_nproc.txt
