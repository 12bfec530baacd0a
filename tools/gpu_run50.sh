mkdir -p gpurun_out
timeout 200 python tools/time_forward.py fp32 > gpurun_out/forward50.txt 2>&1
echo done
