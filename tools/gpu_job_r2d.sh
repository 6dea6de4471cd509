#!/bin/bash
# the other BASELINE.json configurations and the small-batch regime (N=1): one JSON line each, kept under profiles/
O=gpurun_out/r2d; mkdir -p $O
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$?"; tail -2 $O/$name.err; }
run bench_cfg2 --workload cfg2 --steps 20 --warmup 5
run bench_cfg2_graph --workload cfg2 --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --graph
run bench_cfg2_b1_graph --workload cfg2 --slides-per-gpu 1 --steps 100 --warmup 5 --no-cpu-baseline --no-e2e --graph
run bench_cfg2_b1 --workload cfg2 --slides-per-gpu 1 --steps 50 --warmup 5 --no-cpu-baseline --no-fast-mode
run bench_cfg5 --workload cfg5 --steps 10 --warmup 3
run bench_cfg4 --workload cfg4 --steps 5 --warmup 3 --no-cpu-baseline
run bench_cfg3_b32 --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline
run bench_cfg3_b32_graph --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --graph --no-e2e
PATHS_B200_PDL=0 run bench_cfg3_b32_nopdl --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --no-e2e
run bench_cfg3_b64 --slides-per-gpu 64 --steps 50 --warmup 5 --no-cpu-baseline --no-e2e
run bench_cfg3_b64_graph --slides-per-gpu 64 --steps 50 --warmup 5 --no-cpu-baseline --graph --no-e2e
run bench_cfg3_b256 --slides-per-gpu 256 --steps 20 --warmup 5 --no-cpu-baseline
run bench_cfg3_b256_graph --slides-per-gpu 256 --steps 20 --warmup 5 --no-cpu-baseline --graph --no-e2e
run bench_cfg3_dropout005 --dropout 0.05 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e
PATHS_B200_GRAPHS=1 timeout 600 python bench.py --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --no-fast-mode > $O/bench_cfg3_b32_e2egraph.json 2> $O/bench_cfg3_b32_e2egraph.err; echo "e2e graph rc=$?"
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1].split('/')[-1], 'ms/step', round(d['ms_per_step'],3), 'value', round(d['value']), 'e2e', round(d['e2e']['value']) if d.get('e2e') else None, 'host ms', round(d['host_ms_per_step'],3), 'step_frac', round(d['roofline']['step_frac'],3), 'launches', d['gpu_launches'])
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
