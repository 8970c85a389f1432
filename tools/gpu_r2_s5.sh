#!/bin/bash
# round 2, session 5: 64-row Cholesky tiles: chain on the K4-like band (64 vs 128), dense figure, GPU tests, K4 bench
TAG=${1:-r2s5}; OUT=gpurun_out; mkdir -p $OUT
timeout 300 tools/_build/chol_chain 82 2 0 128 > $OUT/chol_chain128_$TAG.txt 2>&1; echo "chain128 rc=$?"; head -3 $OUT/chol_chain128_$TAG.txt; tail -1 $OUT/chol_chain128_$TAG.txt
timeout 300 tools/_build/chol_chain 170 4 0 64 > $OUT/chol_chain64_$TAG.txt 2>&1; echo "chain64 rc=$?"; head -30 $OUT/chol_chain64_$TAG.txt; tail -1 $OUT/chol_chain64_$TAG.txt
timeout 300 tools/_build/chol_prof 10368 10368 128 > $OUT/chol_prof_dense128_$TAG.txt 2>&1; cat $OUT/chol_prof_dense128_$TAG.txt
timeout 300 tools/_build/chol_prof 10368 10368 64 > $OUT/chol_prof_dense64_$TAG.txt 2>&1; cat $OUT/chol_prof_dense64_$TAG.txt
timeout 2700 python -m pytest tests -m gpu -q -s > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log
grep -E "^\[K4|passed|failed|FAILED|^E  " $OUT/pytest_gpu_$TAG.log | tail -30
timeout 900 python bench.py --no-cpu --no-k5 --no-k3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; python - <<PY
import json
try:
    d = json.load(open("$OUT/bench_$TAG.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"])
    print("roofline", {k: d["roofline"][k] for k in ("achieved", "frac", "avg_launch_ms", "share_of_step", "filled_tiles", "lower_tiles", "padded_order")})
    print("dense chol", d.get("roofline_dense_cholesky")); print("all tiles", d.get("all_tiles_filled"))
except Exception as e:
    print("bench json unreadable:", e)
PY
