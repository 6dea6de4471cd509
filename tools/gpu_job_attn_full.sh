#!/bin/bash
# ncu --set full of the short-bag tcgen05 attention kernels (one step of cfg3, 672 slides); run only after gpu_job_attn.sh passed
O=gpurun_out/attn; mkdir -p $O
export PATHS_BENCH_CUPROF=1
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
$B > $O/plain2.log 2>&1 && \
ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:attn_tc_(fwd|bwd)_kernel" -c 8 -o $O/prof_attn $B > $O/ncu2.log 2>&1
echo "full rc=$?"; ls -la $O
