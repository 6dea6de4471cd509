#!/bin/bash
# after the CTA-pair contractions: full GPU suite (plain and with pairs forced), launch list, ncu --set full of the pair kernels
O=gpurun_out/r2o; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.txt 2>&1; echo "suite rc=$?"; tail -3 $O/pytest_gpu.txt
PATHS_B200_FORCE_PAIR=1 timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu_forced_pairs.txt 2>&1; echo "suite(forced pairs) rc=$?"; tail -3 $O/pytest_gpu_forced_pairs.txt
export PATHS_BENCH_CUPROF=1
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
$B > $O/plain.log 2>&1 && \
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_672.csv $B > $O/ncu1.log 2>&1
echo "launch-list rc=$?"
python tools/launch_summary.py $O/launches_672.csv > $O/launches_672.summary.txt; head -30 $O/launches_672.summary.txt
$B > $O/plain3.log 2>&1 && \
ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:gemm_tma_kernel" -c 14 -o $O/prof_gemm $B > $O/ncu3.log 2>&1
echo "full(gemm) rc=$?"
