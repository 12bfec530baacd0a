# One GPU-box pass of the round: parity tests, the driver's bench line, the other configurations, a PCIe probe and
# the ncu launch list of a (shortened) bench command.   usage: bash tools/gpu_round.sh <tag> [tests|regimes|bench|q5|quick|trace|ncu ...]
TAG=${1:-run}; shift
WHAT=${@:-tests bench ncu}
O=gpurun_out/$TAG
mkdir -p $O
for w in $WHAT; do case $w in
tests)
  # (the regime tests in their own process: a crashed kernel there must not poison the rest of the suite)
  timeout 1200 python -m pytest tests/test_0_regimes_gpu.py tests/test_kernels_gpu.py -m gpu -q --tb=short -p no:cacheprovider -rA > $O/pytest_regimes.log 2>&1; echo "pytest exit=$?" >> $O/pytest_regimes.log
  timeout 1500 python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider -rA --deselect tests/test_0_regimes_gpu.py --deselect tests/test_kernels_gpu.py > $O/pytest.log 2>&1; echo "pytest exit=$?" >> $O/pytest.log ;;
regimes)
  timeout 1200 python -m pytest tests/test_0_regimes_gpu.py tests/test_kernels_gpu.py -m gpu -q --tb=short -p no:cacheprovider -rA > $O/pytest_regimes.log 2>&1; echo "pytest exit=$?" >> $O/pytest_regimes.log ;;
q5)
  timeout 600 python bench.py --steps 10 --no-cpu-baseline > $O/bench_cfg5.json 2> $O/bench_cfg5.err ;;
trace)
  # debug-knob build only: kernel timeline of one decode of the default bench's batch shape
  timeout 300 python tools/trace_decode.py cfg5 f16 1024 > $O/trace_cfg5.txt 2>&1 ;;
bench)
  timeout 900 python bench.py > $O/bench_cfg5.json 2> $O/bench_cfg5.err
  timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
  timeout 600 python bench.py --workload cfg2 --batch-videos 64 --no-cpu-baseline > $O/bench_cfg2.json 2> $O/bench_cfg2.err
  timeout 600 python bench.py --workload cfg3 --steps 5 --no-cpu-baseline > $O/bench_cfg3.json 2> $O/bench_cfg3.err
  timeout 600 python bench.py --workload cfg4 --steps 5 --no-cpu-baseline > $O/bench_cfg4.json 2> $O/bench_cfg4.err
  python - > $O/pcie.txt 2>&1 <<'PY'
import torch, time
h = torch.empty(256 << 20, dtype=torch.uint8).pin_memory(); d = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for name, a, b in (("h2d", d, h), ("d2h", h, d)):
    for _ in range(2): a.copy_(b, non_blocking=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(10): a.copy_(b, non_blocking=True)
    torch.cuda.synchronize(); print(name, "%.1f GB/s" % (10 * 0.268435456 / (time.perf_counter() - t0)))
PY
  ;;
stress)
  # race hunt in place of compute-sanitizer (closed on this pool): bitwise-identical replays of five batch shapes, greedy and beam
  timeout 900 python tools/stress_determinism.py 100 f16 > $O/stress_f16.txt 2>&1; echo "exit=$?" >> $O/stress_f16.txt
  timeout 900 python tools/stress_determinism.py 60 fp32 > $O/stress_fp32.txt 2>&1; echo "exit=$?" >> $O/stress_fp32.txt ;;
quick)
  timeout 600 python bench.py --steps 10 --no-cpu-baseline --no-second-beam > $O/bench_cfg5.json 2> $O/bench_cfg5.err
  timeout 600 python bench.py --workload cfg2 --batch-videos 64 --no-cpu-baseline --no-second-beam > $O/bench_cfg2.json 2> $O/bench_cfg2.err
  timeout 600 python bench.py --workload cfg3 --steps 5 --no-cpu-baseline --no-second-beam > $O/bench_cfg3.json 2> $O/bench_cfg3.err
  timeout 600 python bench.py --workload cfg4 --steps 5 --no-cpu-baseline --no-second-beam > $O/bench_cfg4.json 2> $O/bench_cfg4.err ;;
attw)
  # debug-knob build only: warps per row of the 16-bit attention forced from the environment
  for w in 2 4 8; do for c in cfg5 cfg3 cfg4 cfg2; do
    ECHR_ATT_W=$w timeout 600 python bench.py --workload $c --steps 5 --no-cpu-baseline --no-second-beam $( [ $c = cfg2 ] && echo --batch-videos 64 ) > $O/attw${w}_$c.json 2> $O/attw${w}_$c.err
  done; done ;;
sweep)
  for bv in 128 256 1024; do timeout 600 python bench.py --batch-videos $bv --steps 10 --no-cpu-baseline --no-second-beam > $O/sweep_bv$bv.json 2> $O/sweep_bv$bv.err; done
  for fl in 2 3; do timeout 600 python bench.py --in-flight $fl --steps 10 --no-cpu-baseline --no-second-beam > $O/sweep_fl$fl.json 2> $O/sweep_fl$fl.err; done ;;
ncu)
  CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-second-beam"
  $CMD > $O/ncu_plain.json 2> $O/ncu_plain.err &&
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1 ;;
scale)
  # strong-scaling line of the default workload on NG GPUs (gpurun --gpus NG), NCCL ring / NVLS evidence in the .err
  NG=${NG:-2}
  NCCL_DEBUG=INFO timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 2951$NG \
      bench.py --gpus $NG --steps 10 > $O/scale_n$NG.json 2> $O/scale_n$NG.err
  grep -E "NCCL INFO (Channel|Connected|comm 0x|NVLS|ncclCommInit)" $O/scale_n$NG.err | head -40 > $O/scale_n$NG.nccl.txt ;;
ncufull)
  # ncu --set full of every kernel family (tools/profile_targets.py); the raw page is exported on the box, the
  # report itself only comes back when it fits the 64 MiB return channel
  CMD="python tools/profile_targets.py"
  $CMD > $O/targets_plain.log 2>&1 &&
  timeout 2400 ncu --set full --clock-control none -k regex:"gemm_tc|attention|pair_mlp|relation_attn|finalize|select_|rerank|sst_|event_feats|decode_init|split_planes|gather_tap" \
      -c 500 -o $O/targets -f $CMD > $O/targets_ncu.log 2>&1
  ncu -i $O/targets.ncu-rep --page raw --csv > $O/targets_raw.csv 2> $O/targets_export.log
  if [ $(stat -c %s $O/targets.ncu-rep 2>/dev/null || echo 0) -gt 40000000 ]; then rm -f $O/targets.ncu-rep; fi ;;
ncustep)
  # ncu --set full of the decode-step kernels at the default bench's batch shape (cfg5, 1024 videos, ~4070 rows)
  CMD="python tools/profile_targets.py --step-only"
  $CMD > $O/step_plain.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none -k regex:"gemm_tc|attention_f16|finalize_greedy" -c 60 -o $O/step -f $CMD > $O/step_ncu.log 2>&1
  ncu -i $O/step.ncu-rep --page raw --csv > $O/step_raw.csv 2> $O/step_export.log
  if [ $(stat -c %s $O/step.ncu-rep 2>/dev/null || echo 0) -gt 40000000 ]; then rm -f $O/step.ncu-rep; fi ;;
sanitize)
  # compute-sanitizer over smoke(): encode, greedy decode, SST encoder, selection, caption of selected proposals
  timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 7 python -c "import __graft_entry__ as g; g.smoke()" > $O/sanitizer_memcheck.log 2>&1; echo "exit=$?" >> $O/sanitizer_memcheck.log
  timeout 1500 compute-sanitizer --tool racecheck --error-exitcode 7 python -c "import __graft_entry__ as g; g.smoke()" > $O/sanitizer_racecheck.log 2>&1; echo "exit=$?" >> $O/sanitizer_racecheck.log
  timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 7 python tools/profile_targets.py > $O/sanitizer_memcheck_targets.log 2>&1; echo "exit=$?" >> $O/sanitizer_memcheck_targets.log ;;
esac; done
echo done
