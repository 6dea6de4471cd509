#!/bin/bash
# N=1 evidence after the shared-parent gate form / attention class list / programmatic launches: every config, reference arm,
# parity table, launch lists, ncu --set full of the changed kernels
O=gpurun_out/r2n; mkdir -p $O
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_cfg3.json 2> $O/bench_cfg3.err; echo "bench rc=$?"; tail -2 $O/bench_cfg3.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_reference_arm.json 2> $O/ref.err; echo "ref rc=$?"
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/$name.json 2> $O/$name.err; echo "$name rc=$?"; tail -2 $O/$name.err; }
run bench_cfg4 --workload cfg4 --steps 5 --warmup 3 --no-cpu-baseline
run bench_cfg5 --workload cfg5 --steps 10 --warmup 3 --no-cpu-baseline
run bench_cfg2 --workload cfg2 --steps 20 --warmup 5 --no-cpu-baseline
run bench_cfg2_graph --workload cfg2 --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --graph
run bench_cfg2_b1 --workload cfg2 --slides-per-gpu 1 --steps 50 --warmup 5 --no-cpu-baseline
run bench_cfg2_b1_graph --workload cfg2 --slides-per-gpu 1 --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --graph
run bench_cfg3_b256 --slides-per-gpu 256 --steps 20 --warmup 5 --no-cpu-baseline
run bench_cfg3_b256_graph --slides-per-gpu 256 --steps 20 --warmup 5 --no-cpu-baseline --graph --no-e2e
run bench_cfg3_b64 --slides-per-gpu 64 --steps 50 --warmup 5 --no-cpu-baseline
run bench_cfg3_b64_graph --slides-per-gpu 64 --steps 50 --warmup 5 --no-cpu-baseline --graph --no-e2e
run bench_cfg3_b32 --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline
run bench_cfg3_b32_graph --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --graph --no-e2e
PATHS_B200_GRAPHS=1 run bench_cfg3_b32_e2egraph --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --no-fast-mode
run bench_cfg3_dropout --dropout 0.05 --steps 20 --warmup 5 --no-cpu-baseline --no-e2e
PATHS_B200_NO_SHARED_PARENT=1 run bench_cfg3_no_shared_parent --steps 20 --warmup 5 --no-cpu-baseline --no-e2e
PARITY_REPORT_OUT=$O/parity_grads.txt timeout 600 python tools/parity_report.py > $O/parity.log 2>&1; echo "parity rc=$?"; tail -2 $O/parity.log | cut -c1-400
export PATHS_BENCH_CUPROF=1
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
for cfgname in "672:" "32:--slides-per-gpu 32" "cfg4:--workload cfg4" "cfg2:--workload cfg2"; do
  tag=${cfgname%%:*}; extra=${cfgname#*:}
  $B $extra > $O/plain_$tag.log 2>&1 && \
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_$tag.csv $B $extra > $O/ncu_$tag.log 2>&1
  echo "launch-list $tag rc=$?"
  python tools/launch_summary.py $O/launches_$tag.csv > $O/launches_$tag.summary.txt 2>&1
done
$B > $O/plain2.log 2>&1 && \
ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:attn_tc_(fwd|bwd)_kernel" -c 8 -o $O/prof_attn $B > $O/ncu2.log 2>&1
echo "full(attn) rc=$?"
$B > $O/plain3.log 2>&1 && \
ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:gemm_tma_kernel" -c 12 -o $O/prof_gemm $B > $O/ncu3.log 2>&1
echo "full(gemm) rc=$?"
unset PATHS_BENCH_CUPROF
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1].split('/')[-1], 'ms/step', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value']) if d.get('e2e') else None, 'e2e ms', round(d['e2e']['ms_per_step'],3) if d.get('e2e') else None, 'fast', round(d['fast_mode']['ms_per_step'],3) if d.get('fast_mode') else None, 'step_frac', round(d['roofline']['step_frac'],3) if d.get('roofline') else None)
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
