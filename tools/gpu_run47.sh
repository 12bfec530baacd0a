mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/b47_n2.json 2> gpurun_out/b47_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --impl reference --steps 2 --warmup 1 > gpurun_out/b47_n2_ref.json 2> gpurun_out/b47_n2_ref.err
python bench.py --steps 5 > gpurun_out/b47_n1_full.json 2> gpurun_out/b47_n1_full.err
echo done
