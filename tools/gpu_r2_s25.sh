#!/bin/bash
# round 2, session 25: fewer launches per LM / CG iteration (decide + publish, permutations inside the Cholesky, point_invert + schur_init, check + publish)
TAG=${1:-r2s25}; OUT=gpurun_out; mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -q -x > $OUT/pytest_gpu_$TAG.log 2>&1; tail -3 $OUT/pytest_gpu_$TAG.log
timeout 900 python bench.py --no-cpu --no-k3 --no-dense-chol > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
python - <<PY
import json
d = json.load(open("$OUT/bench_$TAG.json"))
print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"], d["stage_seconds_per_step"])
print("roofline", {k: d["roofline"][k] for k in ("achieved", "frac", "avg_launch_ms", "share_of_step")})
k5 = d["k5"]; print("k5", {k: k5.get(k) for k in ("value", "ms_per_step", "e2e", "roofline_frac", "gpu_launches", "error")})
PY
