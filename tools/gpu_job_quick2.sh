#!/bin/bash
# attention + parity tests, default bench line, attention launch list
O=gpurun_out/quick; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_parity.py -q -x > $O/pytest_parity.txt 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_parity.txt
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/bench_cfg3.json 2> $O/bench_cfg3.err; echo "bench rc=$?"; tail -2 $O/bench_cfg3.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/quick/bench_cfg3.json')); r=d['roofline']
print('ms/step', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value']), 'fast', round(d['fast_mode']['ms_per_step'],3) if d.get('fast_mode') else None, 'step_frac', round(r['step_frac'],3))
print('gate fwd ms', round(r['kernel_ms_per_step'],3), 'regions', {k: round(v,2) for k,v in r['region_ms_per_step'].items()})
PY
export PATHS_BENCH_CUPROF=1
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches.csv $B > $O/ncu1.log 2>&1
python tools/launch_summary.py $O/launches.csv > $O/launches.summary.txt; grep -E "launches|attn|gather|split_planes_kernel" $O/launches.summary.txt
