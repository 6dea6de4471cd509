#!/bin/bash
# round-end rehearsal: smoke(), the GPU suite, both bench arms
O=gpurun_out/final; mkdir -p $O
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.txt 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.txt
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.txt 2>&1; echo "suite rc=$?"; tail -2 $O/pytest_gpu.txt
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_reference_arm.json 2> $O/ref.err; echo "ref rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_cfg3.json 2> $O/bench_cfg3.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/final/bench_cfg3.json')); r=json.load(open('gpurun_out/final/bench_reference_arm.json'))
print('ms/step', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value']), 'fast', round(d['fast_mode']['ms_per_step'],3), 'frac', round(d['roofline']['frac'],3), 'step_frac', round(d['roofline']['step_frac'],3), 'clocks', d['clocks'], 'ref', round(r['value'],1), 'e2e ratio', round(d['e2e']['value']/r['value'],1))
PY
