# Sweep of the aggregator level sets (PATHS_B200_DEC_SETS_FWD / _BWD) on the bench workload; run under gpurun.
for cfg in "01234:01234" "0123,4:0123,4" "012,34:012,34" "01,23,4:01,23,4" "0,1,2,3,4:0,1,2,3,4" "0123,4:01234" "01234:0123,4" "012,34:0123,4" "0123,4:012,34"; do
  F=${cfg%%:*}; B=${cfg##*:}
  PATHS_B200_DEC_SETS_FWD=$F PATHS_B200_DEC_SETS_BWD=$B python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-fast-mode 2>/dev/null \
    | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$F | $B', round(d['ms_per_step'],3), {k: round(v,2) for k,v in d['roofline']['region_ms_per_step'].items()})"
done
