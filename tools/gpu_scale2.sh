mkdir -p gpurun_out
python bench.py --no-cpu-baseline > gpurun_out/scale2_n1.json 2> gpurun_out/scale2_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/scale2_n2.json 2> gpurun_out/scale2_n2.err
echo done
