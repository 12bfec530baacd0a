mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$T --master-port 29521 bench.py --gpus 8 --no-cpu-baseline > gpurun_out/scale8d_k20.json 2> gpurun_out/scale8d_k20.err
$T --master-port 29522 bench.py --gpus 8 --in-flight 1 --no-cpu-baseline > gpurun_out/scale8d_d1.json 2> gpurun_out/scale8d_d1.err
CUDA_VISIBLE_DEVICES=5 python bench.py --no-cpu-baseline > gpurun_out/scale8d_n1_gpu5.json 2> gpurun_out/scale8d_n1_gpu5.err
