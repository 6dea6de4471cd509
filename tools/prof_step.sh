# Step profile on one B200 (run under gpurun): plain run, then the ncu launch list of the TIMED steps only
# (bench.py brackets them with cudaProfilerStart/Stop when PATHS_BENCH_CUPROF=1), then optionally one --set full capture.
#   bash tools/prof_step.sh [full-regex] [extra bench args]
set -x
FULL="$1"; shift
export PATHS_BENCH_CUPROF=1
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode $*"
$B > gpurun_out/plain.log 2>&1 && \
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu1.log 2>&1
echo "launch-list rc=$?"
if [ -n "$FULL" ]; then
$B > gpurun_out/plain2.log 2>&1 && \
ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:$FULL" -c 12 -o gpurun_out/prof_full $B > gpurun_out/ncu2.log 2>&1
echo "full rc=$?"
fi
tail -1 gpurun_out/plain.log
