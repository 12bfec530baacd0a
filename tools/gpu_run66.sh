mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_sst.py tests/test_gemm_gpu.py -m gpu -q --tb=short -x -s -p no:cacheprovider > gpurun_out/r66_pytest.log 2>&1; echo "pytest exit=$?" >> gpurun_out/r66_pytest.log
timeout 300 python tools/bench_sst.py f16 64 100 > gpurun_out/r66_sst_f16.json 2> gpurun_out/r66_sst.err
timeout 300 python tools/bench_sst.py fp32 64 100 > gpurun_out/r66_sst_fp32.json 2>> gpurun_out/r66_sst.err
timeout 300 python tools/bench_sst.py f16 1 100 > gpurun_out/r66_sst_f16_v1.json 2>> gpurun_out/r66_sst.err
