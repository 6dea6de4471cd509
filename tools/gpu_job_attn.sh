#!/bin/bash
# quick attention iteration: numerics, then the ncu launch list of one step (attention kernels only)
O=gpurun_out/attn; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_attention.py -q -x > $O/pytest_attn.txt 2>&1; rc=$?; echo "pytest(attn) rc=$rc"; tail -6 $O/pytest_attn.txt
if [ $rc -ne 0 ]; then exit 0; fi
export PATHS_BENCH_CUPROF=1
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
$B > $O/plain.log 2>&1 && \
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -k regex:attn -c 200 --csv --log-file $O/launches.csv $B > $O/ncu1.log 2>&1
echo "launch-list rc=$?"
python tools/launch_summary.py $O/launches.csv | head -12
