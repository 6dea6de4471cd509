#!/bin/bash
# A/B on one box: plain launches vs programmatic launches with the implicit (exit-time) trigger
O=gpurun_out/pdl; mkdir -p $O
for v in 0 1 0 1; do
PATHS_B200_PDL=$v timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > $O/bench_pdl$v.json 2> $O/err$v.txt; echo "pdl=$v rc=$?"
python - $O/bench_pdl$v.json <<'PY'
import json,sys
d=json.load(open(sys.argv[1])); print('  ms/step', round(d['ms_per_step'],3), 'fast', round(d['fast_mode']['ms_per_step'],3) if d.get('fast_mode') else None)
PY
done
PATHS_B200_PDL=1 timeout 600 python -m pytest tests/test_gpu_parity.py -q -x 2>&1 | tail -2
for v in 0 1; do
PATHS_B200_PDL=$v timeout 600 python bench.py --slides-per-gpu 32 --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --no-fast-mode > $O/bench32_pdl$v.json 2> $O/err32_$v.txt; echo "b32 pdl=$v rc=$?"
python - $O/bench32_pdl$v.json <<'PY'
import json,sys
d=json.load(open(sys.argv[1])); print('  ms/step', round(d['ms_per_step'],3))
PY
done
