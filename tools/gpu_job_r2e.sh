#!/bin/bash
# ncu evidence for the round: launch list of one timed step (cfg3, 672 slides, bf16x3) and --set full captures of the new kernels
O=gpurun_out/r2e; mkdir -p $O
export PATHS_BENCH_CUPROF=1
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
$B > $O/plain.log 2>&1 && \
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches.csv $B > $O/ncu1.log 2>&1
echo "launch-list rc=$?"; tail -2 $O/ncu1.log
$B > $O/plain2.log 2>&1 && \
ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:attn_tc" -c 4 -o $O/prof_attn $B > $O/ncu2.log 2>&1
echo "full(attn) rc=$?"; tail -2 $O/ncu2.log
$B > $O/plain3.log 2>&1 && \
ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:gemm_tma_kernel" -c 6 -o $O/prof_gemm $B > $O/ncu3.log 2>&1
echo "full(gemm) rc=$?"; tail -2 $O/ncu3.log
B4="python bench.py --workload cfg4 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
$B4 > $O/plain4.log 2>&1 && \
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_cfg4.csv $B4 > $O/ncu4.log 2>&1
echo "launch-list cfg4 rc=$?"; tail -2 $O/ncu4.log
ls -la $O
