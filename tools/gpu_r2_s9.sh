#!/bin/bash
TAG=${1:-r2s9}; OUT=gpurun_out; mkdir -p $OUT
timeout 300 tools/_build/chol_prof 10880 300 64 > $OUT/chol_prof_band64_$TAG.txt 2>&1; cat $OUT/chol_prof_band64_$TAG.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > $OUT/pytest_parity_$TAG.log 2>&1; tail -2 $OUT/pytest_parity_$TAG.log
timeout 900 python bench.py --no-cpu --no-k3 --no-k5 --no-dense-chol > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
python - <<PY
import json
d = json.load(open("$OUT/bench_$TAG.json"))
print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"], d["stage_seconds_per_step"])
print("roofline", {k: d["roofline"][k] for k in ("achieved", "frac", "avg_launch_ms", "share_of_step")})
PY
