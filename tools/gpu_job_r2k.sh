#!/bin/bash
O=gpurun_out/r2k; mkdir -p $O
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_cfg3.json 2> $O/bench_cfg3.err; echo "bench rc=$?"; tail -2 $O/bench_cfg3.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_reference_arm.json 2> $O/ref.err; echo "ref rc=$?"
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$?"; tail -2 $O/$name.err; }
run bench_cfg4 --workload cfg4 --steps 5 --warmup 3 --no-cpu-baseline
run bench_cfg2 --workload cfg2 --steps 20 --warmup 5 --no-cpu-baseline
run bench_cfg2_graph --workload cfg2 --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --graph
run bench_cfg3_b32 --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline
run bench_cfg3_b32_graph --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --graph --no-e2e
PATHS_B200_GRAPHS=1 run bench_cfg3_b32_e2egraph --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --no-fast-mode
PARITY_REPORT_OUT=$O/parity_grads.txt timeout 600 python tools/parity_report.py > $O/parity.log 2>&1; echo "parity rc=$?"; tail -2 $O/parity.log | cut -c1-400
bash tools/gpu_job_r2e.sh
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1].split('/')[-1], 'ms/step', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value']) if d.get('e2e') else None, 'e2e ms', round(d['e2e']['ms_per_step'],3) if d.get('e2e') else None, d['e2e'].get('host_wall_ms_per_step') if d.get('e2e') else None, 'fast', round(d['fast_mode']['ms_per_step'],3) if d.get('fast_mode') else None, 'step_frac', round(d['roofline']['step_frac'],3) if d.get('roofline') else None)
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
