mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --no-cpu-baseline > gpurun_out/scale8_cfg2.json 2> gpurun_out/scale8_cfg2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 8 --workload cfg5 --videos 512 --steps 5 --no-cpu-baseline > gpurun_out/scale8_cfg5.json 2> gpurun_out/scale8_cfg5.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --workload cfg5 --videos 512 --beam 3 --steps 3 --no-cpu-baseline > gpurun_out/scale8_cfg5b3.json 2> gpurun_out/scale8_cfg5b3.err
echo done
