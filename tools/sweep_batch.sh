# Throughput vs slides per GPU (wave quantisation of the 128-row tiles over 148 SMs); run under gpurun.
for B in ${SWEEP:-256 296 320 336 352 384 448 512}; do
  python bench.py --slides-per-gpu $B --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-fast-mode 2>/dev/null \
    | python -c "import sys,json; d=json.loads(sys.stdin.read()); print($B, 'ms', round(d['ms_per_step'],3), 'slides/s', round(d['value']), 'rows', d['config']['rows_per_level'], 'gate frac', round(d['roofline']['frac'],3))"
done
