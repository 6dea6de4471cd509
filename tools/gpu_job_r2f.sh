#!/bin/bash
# multi-GPU evidence: usage  bash tools/gpu_job_r2f.sh N   (under gpurun --gpus N)
N=${1:-2}; O=gpurun_out/r2f; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
if [ "$N" = "2" ]; then
  timeout 600 $TR tools/dp_check.py > $O/dp_check_n2.txt 2> $O/dp_check_n2.err; echo "dp_check rc=$?"; tail -25 $O/dp_check_n2.txt
fi
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > $O/bench_cfg3_n$N.json 2> $O/bench_cfg3_n$N.err; echo "bench cfg3 N=$N rc=$?"; tail -2 $O/bench_cfg3_n$N.err
timeout 900 $TR bench.py --gpus $N --workload cfg5 --steps 10 --warmup 3 > $O/bench_cfg5_n$N.json 2> $O/bench_cfg5_n$N.err; echo "bench cfg5 N=$N rc=$?"; tail -2 $O/bench_cfg5_n$N.err
for f in $O/bench_*_n$N.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1].split('/')[-1], 'ms/step', round(d['ms_per_step'],3), 'value', round(d['value']), 'e2e', round(d['e2e']['value']), 'e2e ms', round(d['e2e']['ms_per_step'],3), 'exposed comm', d.get('exposed_comm_ms_per_step'), d['e2e'].get('exposed_comm_ms_per_step'))
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
