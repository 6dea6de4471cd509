#!/bin/bash
# epilogue change: parity, bench (both modes), per-contraction times of the first launches
O=gpurun_out/epi; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x > $O/pytest_parity.txt 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_parity.txt
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > $O/bench_cfg3.json 2> $O/bench_cfg3.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/epi/bench_cfg3.json')); r=d['roofline']
print('ms/step', round(d['ms_per_step'],3), 'fast', round(d['fast_mode']['ms_per_step'],3), 'gate fwd ms', round(r['kernel_ms_per_step'],3), {k: round(v,2) for k,v in r['region_ms_per_step'].items()}, 'fast gate frac', round(d['fast_mode']['gate_kernel_frac_of_peak'],3))
PY
export PATHS_BENCH_CUPROF=1
for prec in bf16x3 bf16; do
B="python bench.py --precision $prec --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
ncu --profile-from-start off --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k "regex:gemm_tma_kernel" -c 10 --csv --log-file $O/gemm_$prec.csv $B > $O/ncu_$prec.log 2>&1
python - $O/gemm_$prec.csv <<'PY'
import csv,io,sys
rows=[l for l in open(sys.argv[1]) if l.startswith('"')]
r=list(csv.DictReader(io.StringIO(''.join(rows))))
out={}
for x in r: out.setdefault(x['ID'],{})[x['Metric Name']]=x['Metric Value']
print(sys.argv[1], ' | '.join(f"{float(v['gpu__time_duration.sum'])/1e3:.1f}us {float(v['sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active']):.0f}%" for v in out.values()))
PY
done
