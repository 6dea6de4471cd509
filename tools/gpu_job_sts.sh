#!/bin/bash
O=gpurun_out/sts; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_parity.py tests/test_gpu_dropout.py -q -x > $O/pytest.txt 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest.txt
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > $O/bench_cfg3.json 2> $O/err.txt; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/sts/bench_cfg3.json')); r=d['roofline']
print('ms/step', round(d['ms_per_step'],3), 'fast', round(d['fast_mode']['ms_per_step'],3), 'gate fwd ms', round(r['kernel_ms_per_step'],3), {k: round(v,2) for k,v in r['region_ms_per_step'].items()}, 'fast gate frac', round(d['fast_mode']['gate_kernel_frac_of_peak'],3))
PY
export PATHS_BENCH_CUPROF=1
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches.csv $B > $O/ncu1.log 2>&1
python tools/launch_summary.py $O/launches.csv > $O/launches.summary.txt; head -12 $O/launches.summary.txt; grep attn $O/launches.summary.txt
timeout 600 python bench.py --workload cfg4 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode > $O/bench_cfg4.json 2>> $O/err.txt; python -c "
import json; d=json.load(open('$O/bench_cfg4.json')); print('cfg4 ms/step', round(d['ms_per_step'],3))"
