#!/bin/bash
# the allocation-free prune entry point: alone in a fresh process, then after the whole GPU suite (collected last)
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $OUT/noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -3 $OUT/noalloc_alone.log
timeout 2400 python -m pytest tests -x -q -m gpu --durations=8 > $OUT/pytest_gpu_noalloc.log 2>&1; echo "suite rc=$?"; tail -14 $OUT/pytest_gpu_noalloc.log
