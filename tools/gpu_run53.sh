mkdir -p gpurun_out
python bench.py --no-cpu-baseline > gpurun_out/b53_n1.json 2> gpurun_out/b53_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/b53_n2.json 2> gpurun_out/b53_n2.err
python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b53_cfg5.json 2> gpurun_out/b53_cfg5.err
echo done
