#!/bin/bash
# tensor-pipe activity of the first contractions of a step (gate / Wmo / Wuv forward of levels 0..2) + parity tests + bench line
O=gpurun_out/gemmprof; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x > $O/pytest_parity.txt 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_parity.txt
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > $O/bench_cfg3.json 2> $O/bench_cfg3.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/gemmprof/bench_cfg3.json')); r=d['roofline']
print('ms/step', round(d['ms_per_step'],3), 'fast', round(d['fast_mode']['ms_per_step'],3), 'gate fwd ms', round(r['kernel_ms_per_step'],3), {k: round(v,2) for k,v in r['region_ms_per_step'].items()}, 'fast gate frac', round(d['fast_mode']['gate_kernel_frac_of_peak'],3))
PY
export PATHS_BENCH_CUPROF=1
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
ncu --profile-from-start off --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k "regex:gemm_tma_kernel" -c 14 --csv --log-file $O/gemm.csv $B > $O/ncu.log 2>&1
python - <<'PY'
import csv,io
rows=[l for l in open('gpurun_out/gemmprof/gemm.csv') if l.startswith('"')]
r=list(csv.DictReader(io.StringIO(''.join(rows))))
out={}
for x in r: out.setdefault(x['ID'],{})[x['Metric Name']]=x['Metric Value']; out[x['ID']]['k']=x['Kernel Name'][x['Kernel Name'].index('gemm_tma'):][:38]
for i,v in out.items(): print(i, v['k'], v.get('gpu__time_duration.sum'), v.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'))
PY
