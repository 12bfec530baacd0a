mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest6.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest6.log
python bench.py --no-cpu-baseline > gpurun_out/b6_n1.json 2> gpurun_out/b6_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/b6_n2.json 2> gpurun_out/b6_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --impl reference --steps 2 --warmup 1 > gpurun_out/b6_n2_ref.json 2> gpurun_out/b6_n2_ref.err
echo done
