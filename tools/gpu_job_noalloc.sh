#!/bin/bash
# the allocation-free entry point: alone in a fresh process, then after the whole GPU suite
O=gpurun_out/noalloc; mkdir -p $O
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -x > $O/alone.txt 2>&1; echo "alone rc=$?"; tail -3 $O/alone.txt
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.txt 2>&1; echo "suite rc=$?"; tail -3 $O/pytest_gpu.txt
grep -n "no_alloc" $O/pytest_gpu.txt | head -3
