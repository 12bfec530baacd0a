mkdir -p gpurun_out
python bench.py --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
for n in 2 4 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --no-cpu-baseline > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --workload cfg5 --videos 512 --steps 5 --no-cpu-baseline > gpurun_out/scale_cfg5_n8.json 2> gpurun_out/scale_cfg5_n8.err
python bench.py --workload cfg5 --videos 512 --steps 5 --no-cpu-baseline > gpurun_out/scale_cfg5_n1.json 2> gpurun_out/scale_cfg5_n1.err
echo done
