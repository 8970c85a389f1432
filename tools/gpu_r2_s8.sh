#!/bin/bash
# round 2, session 8: new potrf (A) (augmented elimination), restored schur_units, PCG modes: tests + breakdown + K4/K5 bench
TAG=${1:-r2s8}; OUT=gpurun_out; mkdir -p $OUT
timeout 300 tools/_build/chol_prof 10880 300 64 > $OUT/chol_prof_band64_$TAG.txt 2>&1; cat $OUT/chol_prof_band64_$TAG.txt
timeout 300 tools/_build/chol_chain 170 4 0 64 > $OUT/chol_chain64_$TAG.txt 2>&1; head -8 $OUT/chol_chain64_$TAG.txt; tail -1 $OUT/chol_chain64_$TAG.txt
timeout 2700 python -m pytest tests -m gpu -q -s > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log
grep -E "passed|failed|FAILED|^E  " $OUT/pytest_gpu_$TAG.log | tail -20
timeout 900 python bench.py --no-cpu --no-k3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; python - <<PY
import json
try:
    d = json.load(open("$OUT/bench_$TAG.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"], d["stage_seconds_per_step"])
    print("roofline", {k: d["roofline"][k] for k in ("achieved", "frac", "avg_launch_ms", "share_of_step")})
    print("dense chol", d.get("roofline_dense_cholesky", {}).get("frac"), "all tiles", d.get("all_tiles_filled"))
    print("k5", d["k5"])
except Exception as e:
    print("bench json unreadable:", e)
PY
