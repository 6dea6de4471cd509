#!/bin/bash
O=gpurun_out/r2g; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.txt 2>&1; echo "pytest rc=$?"; tail -8 $O/pytest.txt
timeout 600 python tools/prof_e2e_steps.py 672 > $O/e2e_steps.txt 2>&1; echo "e2e steps rc=$?"; cat $O/e2e_steps.txt | cut -c1-200
export PATHS_BENCH_CUPROF=1
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
$B > $O/plain.log 2>&1 && \
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches.csv $B > $O/ncu1.log 2>&1
echo "launch-list rc=$?"; tail -2 $O/ncu1.log
unset PATHS_BENCH_CUPROF
python tools/launch_summary.py $O/launches.csv > $O/launches.summary.txt 2>&1; head -40 $O/launches.summary.txt
