mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_proposals.py tests/test_abi.py -m gpu -q --tb=short -x -p no:cacheprovider > gpurun_out/r65_pytest.log 2>&1; echo "pytest exit=$?" >> gpurun_out/r65_pytest.log
timeout 300 python tools/bench_select.py > gpurun_out/r65_select.json 2> gpurun_out/r65_select.err
timeout 300 python tools/bench_select.py 64 400 256 > gpurun_out/r65_select_T400.json 2>> gpurun_out/r65_select.err
