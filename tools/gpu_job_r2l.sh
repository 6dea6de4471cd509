#!/bin/bash
# full GPU suite + default bench line (after the attention staging / class-list changes)
O=gpurun_out/r2l; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest_gpu.txt
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_cfg3.json 2> $O/bench_cfg3.err; echo "bench rc=$?"; tail -2 $O/bench_cfg3.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2l/bench_cfg3.json')); print('ms/step', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value']), 'fast', round(d['fast_mode']['ms_per_step'],3) if d.get('fast_mode') else None, 'step_frac', round(d['roofline']['step_frac'],3))
PY
