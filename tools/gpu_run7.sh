mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -x > gpurun_out/pytest7.log 2>&1; echo "pytest exit=$?" >> gpurun_out/pytest7.log
python bench.py --no-cpu-baseline > gpurun_out/b7_pdl.json 2> gpurun_out/b7_pdl.err
ECHR_NO_PDL=1 python bench.py --no-cpu-baseline > gpurun_out/b7_nopdl.json 2> gpurun_out/b7_nopdl.err
python bench.py --workload cfg5 --videos 1024 --steps 5 --no-cpu-baseline > gpurun_out/b7_cfg5.json 2> gpurun_out/b7_cfg5.err
echo done
