#!/bin/bash
# round 2, session 13: chain fusion in the dataflow Cholesky (diagonal task forms its own copy of the last off-diagonal tile)
TAG=${1:-r2s13}; OUT=gpurun_out; mkdir -p $OUT
timeout 300 tools/_build/chol_prof 10880 300 64 > $OUT/chol_prof_band64_$TAG.txt 2>&1; cat $OUT/chol_prof_band64_$TAG.txt
timeout 300 tools/_build/chol_chain 170 4 0 64 > $OUT/chol_chain64_$TAG.txt 2>&1; head -14 $OUT/chol_chain64_$TAG.txt; tail -1 $OUT/chol_chain64_$TAG.txt
timeout 1800 python -m pytest tests -m gpu -q -x > $OUT/pytest_gpu_$TAG.log 2>&1; tail -3 $OUT/pytest_gpu_$TAG.log
timeout 900 python bench.py --no-cpu --no-k3 --no-k5 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
python - <<PY
import json
d = json.load(open("$OUT/bench_$TAG.json"))
print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"], d["stage_seconds_per_step"])
print("roofline", {k: d["roofline"][k] for k in ("achieved", "frac", "avg_launch_ms", "share_of_step")})
print("dense chol", d.get("roofline_dense_cholesky", {}).get("frac"), "all tiles", d.get("all_tiles_filled"))
PY
