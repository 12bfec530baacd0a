mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --tb=short -x -p no:cacheprovider > gpurun_out/r64_pytest.log 2>&1; echo "pytest exit=$?" >> gpurun_out/r64_pytest.log
for p in f16 bf16w; do for d in 1 2 3; do
timeout 300 python bench.py --precision $p --in-flight $d --no-cpu-baseline > gpurun_out/r64_cfg2_${p}_d$d.json 2> gpurun_out/r64_cfg2_${p}_d$d.err
done; done
for d in 1 2; do
timeout 300 python bench.py --precision f16 --in-flight $d --workload cfg5 --videos 1024 --steps 6 --no-cpu-baseline > gpurun_out/r64_cfg5_f16_d$d.json 2> gpurun_out/r64_cfg5_f16_d$d.err
timeout 300 python bench.py --precision f16 --in-flight $d --workload cfg3 --steps 4 --no-cpu-baseline > gpurun_out/r64_cfg3_f16_d$d.json 2> gpurun_out/r64_cfg3_f16_d$d.err
done
