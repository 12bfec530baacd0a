mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 300 python bench.py --precision f16 --no-cpu-baseline $EXTRA > gpurun_out/r62_$tag.json 2> gpurun_out/r62_$tag.err; }
EXTRA=""
run base X=1
run ks0 ECHR_KSPLIT=0
run ks3 ECHR_KSPLIT=3
run hoist1 ECHR_HOIST_CTX=1
run h2bn64 ECHR_H2ATT_BN=64
run h2bn16 ECHR_H2ATT_BN=16
run pdl15 ECHR_PDL_MASK=15
run pdl3 ECHR_PDL_MASK=3
EXTRA="--workload cfg4 --steps 5"
run c4_base X=1
run c4_g2_512 ECHR_GEMM2_MIN_ROWS=512
run c4_g2_2048 ECHR_GEMM2_MIN_ROWS=2048
EXTRA="--workload cfg5 --videos 1024 --steps 5"
run c5_base X=1
run c5_hoist1 ECHR_HOIST_CTX=1
