#!/bin/bash
# small batches: shared-parent form on/off
O=gpurun_out/small; mkdir -p $O
for b in 32 64 128; do for v in 0 1; do
if [ $v = 1 ]; then export PATHS_B200_NO_SHARED_PARENT=1; else unset PATHS_B200_NO_SHARED_PARENT; fi
timeout 600 python bench.py --slides-per-gpu $b --steps 50 --warmup 5 --no-cpu-baseline --no-e2e --no-fast-mode > $O/b${b}_noshared$v.json 2> $O/err.txt
python - $O/b${b}_noshared$v.json $b $v <<'PY'
import json,sys
d=json.load(open(sys.argv[1])); print('B', sys.argv[2], 'no_shared', sys.argv[3], 'ms/step', round(d['ms_per_step'],3), 'launches/step', d['gpu_launches']/d['steps'])
PY
done; done
