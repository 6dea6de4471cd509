#!/bin/bash
# tiled (long-bag) attention after the staging rewrite: numerics, dropout tests, cfg4 bench + launch list
O=gpurun_out/long; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_attention.py tests/test_gpu_dropout.py tests/test_zz_no_alloc.py -q -x > $O/pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest.txt
timeout 900 python bench.py --workload cfg4 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_cfg4.json 2> $O/bench_cfg4.err; echo "cfg4 rc=$?"; tail -2 $O/bench_cfg4.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/long/bench_cfg4.json')); print('cfg4 ms/step', round(d['ms_per_step'],3), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value']), 'fast', round(d['fast_mode']['ms_per_step'],3))
PY
export PATHS_BENCH_CUPROF=1
B="python bench.py --workload cfg4 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_cfg4.csv $B > $O/ncu1.log 2>&1
python tools/launch_summary.py $O/launches_cfg4.csv > $O/launches_cfg4.summary.txt; head -8 $O/launches_cfg4.summary.txt
