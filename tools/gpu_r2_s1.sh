#!/bin/bash
# round 2, session 1: GPU tests (old + full-size parity), Cholesky chain / potrf phase breakdown, K4 + K5 bench lines
TAG=${1:-r2s1}; OUT=gpurun_out; mkdir -p $OUT
nvidia-smi -L > $OUT/gpu_$TAG.txt 2>&1; nproc >> $OUT/gpu_$TAG.txt; free -g >> $OUT/gpu_$TAG.txt
timeout 300 tools/_build/chol_chain 82 2 > $OUT/chol_chain_$TAG.txt 2>&1; echo "chol_chain rc=$?"
timeout 300 tools/_build/chol_prof 10368 300 > $OUT/chol_prof_band_$TAG.txt 2>&1; echo "chol_prof rc=$?"; cat $OUT/chol_prof_band_$TAG.txt
timeout 2700 python -m pytest tests -m gpu -x -q -s > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log
grep -E "^\[|landmark |passed|failed|Error|error" $OUT/pytest_gpu_$TAG.log | tail -60
timeout 600 python bench.py --no-cpu > $OUT/bench_K4_$TAG.json 2> $OUT/bench_K4_$TAG.err
echo "bench K4 rc=$?"; cut -c1-600 $OUT/bench_K4_$TAG.json
