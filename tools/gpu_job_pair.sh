#!/bin/bash
# CTA-pair contraction: parity with every eligible contraction forced onto pairs, then A/B bench
O=gpurun_out/pair; mkdir -p $O
PATHS_B200_FORCE_PAIR=1 timeout 600 python -m pytest tests/test_gpu_parity.py -q -x > $O/pytest_forced.txt 2>&1; echo "forced pytest rc=$?"; tail -5 $O/pytest_forced.txt
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_train.py -q -x > $O/pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.txt
for v in 1 0 1 0; do
if [ $v = 1 ]; then unset PATHS_B200_NO_PAIR; else export PATHS_B200_NO_PAIR=1; fi
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > $O/bench_pair$v.json 2> $O/err$v.txt; echo "pair=$v rc=$?"; tail -1 $O/err$v.txt
python - $O/bench_pair$v.json <<'PY'
import json,sys
d=json.load(open(sys.argv[1])); r=d['roofline']; print('  ms/step', round(d['ms_per_step'],3), 'fast', round(d['fast_mode']['ms_per_step'],3), 'gate fwd ms', round(r['kernel_ms_per_step'],3), 'fast gate frac', round(d['fast_mode']['gate_kernel_frac_of_peak'],3), {k: round(v,2) for k,v in r['region_ms_per_step'].items()})
PY
done
