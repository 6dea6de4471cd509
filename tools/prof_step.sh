set -x
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --slides-per-gpu 256"
$B > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu1.log 2>&1
echo "launch-list rc=$?"
$B > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm_tma_kernel -s 30 -c 3 -o gpurun_out/prof_gate $B > gpurun_out/ncu2.log 2>&1
echo "full rc=$?"
tail -3 gpurun_out/plain.log
