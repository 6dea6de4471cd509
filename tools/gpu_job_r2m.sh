#!/bin/bash
# programmatic launches on by default: full GPU suite, graph variants, small batch
O=gpurun_out/r2m; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$?"; tail -2 $O/$name.err; }
run bench_cfg3 --steps 20 --warmup 5 --no-cpu-baseline
run bench_cfg3_graph --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --graph
run bench_cfg3_b32 --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline
run bench_cfg3_b32_graph --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --graph --no-e2e
PATHS_B200_PDL=0 run bench_cfg3_b32_graph_nopdl --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --graph --no-e2e
run bench_cfg2 --workload cfg2 --steps 20 --warmup 5 --no-cpu-baseline
run bench_cfg2_graph --workload cfg2 --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --graph
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1].split('/')[-1], 'ms/step', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value']) if d.get('e2e') else None, 'fast', round(d['fast_mode']['ms_per_step'],3) if d.get('fast_mode') else None)
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
