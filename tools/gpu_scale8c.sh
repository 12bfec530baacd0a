mkdir -p gpurun_out
( cat /sys/fs/cgroup/cpu.max; cat /sys/fs/cgroup/cpu.stat | head -8; nproc; lscpu | grep -i "model name\|thread\|socket\|^CPU(s)" ) > gpurun_out/scale8c_host.txt 2>&1
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$T --master-port 29521 bench.py --gpus 8 --no-cpu-baseline > gpurun_out/scale8c_k20.json 2> gpurun_out/scale8c_k20.err
ECHR_BENCH_NO_SAMPLER=1 $T --master-port 29522 bench.py --gpus 8 --no-cpu-baseline > gpurun_out/scale8c_nosampler.json 2> gpurun_out/scale8c_nosampler.err
( cat /sys/fs/cgroup/cpu.stat | head -8 ) >> gpurun_out/scale8c_host.txt 2>&1
