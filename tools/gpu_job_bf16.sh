#!/bin/bash
# bf16 (single-pass) mode: launch list of one step + tensor activity of the first contractions
O=gpurun_out/bf16; mkdir -p $O
export PATHS_BENCH_CUPROF=1
B="python bench.py --precision bf16 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-fast-mode"
$B > $O/plain.log 2>&1 && \
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches.csv $B > $O/ncu1.log 2>&1
python tools/launch_summary.py $O/launches.csv > $O/launches.summary.txt; head -42 $O/launches.summary.txt
ncu --profile-from-start off --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,lts__throughput.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed --clock-control none -k "regex:gemm_tma_kernel" -c 12 --csv --log-file $O/gemm.csv $B > $O/ncu2.log 2>&1
python - <<'PY'
import csv,io
rows=[l for l in open('gpurun_out/bf16/gemm.csv') if l.startswith('"')]
r=list(csv.DictReader(io.StringIO(''.join(rows))))
out={}
for x in r: out.setdefault(x['ID'],{})[x['Metric Name']]=x['Metric Value']; out[x['ID']]['k']=x['Kernel Name'][x['Kernel Name'].index('gemm_tma'):][:42]
for i,v in out.items(): print(i, v['k'], v.get('gpu__time_duration.sum'), 'tensor', v.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'), 'lts', v.get('lts__throughput.avg.pct_of_peak_sustained_elapsed'), 'dram', v.get('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'))
PY
