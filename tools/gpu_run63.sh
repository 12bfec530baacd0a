mkdir -p gpurun_out
timeout 600 python tools/two_in_flight.py f16 cfg2 > gpurun_out/r63_tif_cfg2.log 2>&1
timeout 600 python tools/two_in_flight.py f16 cfg5 1024 > gpurun_out/r63_tif_cfg5.log 2>&1
timeout 600 python tools/two_in_flight.py bf16w cfg2 > gpurun_out/r63_tif_cfg2_bf16w.log 2>&1
