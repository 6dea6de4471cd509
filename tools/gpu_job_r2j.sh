#!/bin/bash
# 2-GPU call: full GPU test suite on the current build, then the data-parallel checks and N=2 bench lines
O=gpurun_out/r2j; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q > $O/pytest.txt 2>&1; echo "pytest rc=$?"; tail -4 $O/pytest.txt
bash tools/gpu_job_r2f.sh 2
