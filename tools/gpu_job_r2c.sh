#!/bin/bash
# staged GPU job: (1) everything with the SIMT attention and plain launches, (2) the tcgen05 attention on its own,
# (3) the full suite with tcgen05 attention, (4) the full suite with programmatic dependent launch on top, (5) parity table + bench lines
O=gpurun_out/r2c; mkdir -p $O
PATHS_B200_PDL=0 PATHS_B200_ATTN_SIMT=1 timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_attention.py > $O/pytest_simt.txt 2>&1; echo "pytest(simt attn, no pdl) rc=$?"; tail -4 $O/pytest_simt.txt
PATHS_B200_PDL=0 timeout 300 python -m pytest tests/test_gpu_attention.py -q > $O/pytest_attn.txt 2>&1; echo "pytest(attn) rc=$?"; tail -25 $O/pytest_attn.txt
PATHS_B200_PDL=0 timeout 900 python -m pytest tests -m gpu -q > $O/pytest_tc.txt 2>&1; echo "pytest(tc attn, no pdl) rc=$?"; tail -12 $O/pytest_tc.txt
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_pdl.txt 2>&1; echo "pytest(tc attn, pdl) rc=$?"; tail -12 $O/pytest_pdl.txt
PATHS_B200_PDL=0 PATHS_B200_ATTN_SIMT=1 PARITY_REPORT_OUT=$O/parity_grads_simt.txt timeout 600 python tools/parity_report.py > $O/parity.log 2>&1; echo "parity rc=$?"; tail -3 $O/parity.log
PATHS_B200_PDL=0 PATHS_B200_ATTN_SIMT=1 timeout 600 python bench.py --steps 10 --warmup 3 > $O/bench_cfg3_simt_nopdl.json 2> $O/bench_cfg3_simt_nopdl.err; echo "bench(simt,nopdl) rc=$?"; tail -3 $O/bench_cfg3_simt_nopdl.err
PATHS_B200_PDL=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_cfg3_tc_nopdl.json 2> $O/bench_cfg3_tc_nopdl.err; echo "bench(tc,nopdl) rc=$?"; tail -3 $O/bench_cfg3_tc_nopdl.err
PATHS_B200_ATTN_SIMT=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_cfg3_simt_pdl.json 2> $O/bench_cfg3_simt_pdl.err; echo "bench(simt,pdl) rc=$?"; tail -3 $O/bench_cfg3_simt_pdl.err
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_cfg3_tc_pdl.json 2> $O/bench_cfg3_tc_pdl.err; echo "bench(tc,pdl) rc=$?"; tail -3 $O/bench_cfg3_tc_pdl.err
for f in $O/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1].split('/')[-1], 'ms/step', round(d['ms_per_step'],3), 'value', round(d['value']), 'e2e', round(d['e2e']['value']) if d.get('e2e') else None, 'fast ms', round(d['fast_mode']['ms_per_step'],3) if d.get('fast_mode') else None, 'step_frac', round(d['roofline']['step_frac'],3))
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
