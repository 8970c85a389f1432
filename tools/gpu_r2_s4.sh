#!/bin/bash
# round 2, session 4: outer-loop drop-in tests, PCG-vs-dense diagnostic, full GPU suite
TAG=${1:-r2s4}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python tools/pcg_diag.py > $OUT/pcg_diag_$TAG.txt 2>&1; echo "pcg_diag rc=$?"; cat $OUT/pcg_diag_$TAG.txt
timeout 2700 python -m pytest tests -m gpu -q -s > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log
grep -E "^\[|passed|failed|FAILED|need the cond|^E  " $OUT/pytest_gpu_$TAG.log | tail -60
