#!/bin/bash
O=gpurun_out/r2i; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_attention.py tests/test_gpu_dropout.py tests/test_gpu_train.py -q > $O/pytest_sel.txt 2>&1; echo "pytest(sel) rc=$?"; tail -3 $O/pytest_sel.txt
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_cfg3.json 2> $O/bench_cfg3.err; echo "bench rc=$?"; tail -2 $O/bench_cfg3.err
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$?"; tail -2 $O/$name.err; }
PATHS_B200_PDL=1 run bench_cfg3_b32_graph_pdl --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --graph --no-e2e
PATHS_B200_PDL=1 run bench_cfg3_b32_pdl --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --no-fast-mode
PATHS_B200_GRAPHS=1 run bench_cfg3_b32_e2egraph --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --no-fast-mode
PATHS_B200_GRAPHS=1 PATHS_B200_PDL=1 run bench_cfg3_b32_e2egraph_pdl --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --no-fast-mode
run bench_cfg3_bf16store --store-dtype bf16 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode
export PATHS_BENCH_CUPROF=1
for cfgname in "672:" "32:--slides-per-gpu 32" "cfg4:--workload cfg4" "cfg2:--workload cfg2"; do
  tag=${cfgname%%:*}; extra=${cfgname#*:}
  B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode $extra"
  $B > $O/plain_$tag.log 2>&1 && \
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_$tag.csv $B > $O/ncu_$tag.log 2>&1
  echo "launch-list $tag rc=$?"
  python tools/launch_summary.py $O/launches_$tag.csv > $O/launches_$tag.summary.txt 2>&1; head -14 $O/launches_$tag.summary.txt
done
unset PATHS_BENCH_CUPROF
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1].split('/')[-1], 'ms/step', round(d['ms_per_step'],3), 'value', round(d['value']), 'e2e', round(d['e2e']['value']) if d.get('e2e') else None, 'e2e ms', round(d['e2e']['ms_per_step'],3) if d.get('e2e') else None, 'fast', round(d['fast_mode']['ms_per_step'],3) if d.get('fast_mode') else None, 'step_frac', round(d['roofline']['step_frac'],3))
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
