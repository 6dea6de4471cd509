#!/bin/bash
O=gpurun_out/r2h; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_attention.py -q > $O/pytest_attn.txt 2>&1; echo "pytest(attn) rc=$?"; tail -3 $O/pytest_attn.txt
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.txt 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.txt
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_cfg3.json 2> $O/bench_cfg3.err; echo "bench rc=$?"; tail -2 $O/bench_cfg3.err
export PATHS_BENCH_CUPROF=1
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
$B > $O/plain.log 2>&1 && \
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches.csv $B > $O/ncu1.log 2>&1
echo "launch-list rc=$?"
unset PATHS_BENCH_CUPROF
python tools/launch_summary.py $O/launches.csv > $O/launches.summary.txt 2>&1; head -24 $O/launches.summary.txt
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2h/bench_cfg3.json'))
print('ms/step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'with opt', d['e2e']['with_optimizer']['ms_per_step'], 'fast', d['fast_mode']['ms_per_step'], 'step_frac', d['roofline']['step_frac'], d['fast_mode']['step_frac'], 'cpu', d['cpu_baseline']['value'])
PY
bash tools/gpu_job_r2d.sh
