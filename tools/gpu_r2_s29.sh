#!/bin/bash
# round 2, session 29: adaptive unit length of the Schur pair index (short units for small problems)
TAG=${1:-r2s29}; OUT=gpurun_out; mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -q -x > $OUT/pytest_gpu_$TAG.log 2>&1; tail -3 $OUT/pytest_gpu_$TAG.log
for W in K1 K2; do
  timeout 900 python bench.py --workload $W --steps 20 --no-k3 --no-k5 > $OUT/bench_${W}_$TAG.json 2> $OUT/bench_${W}_$TAG.err; echo "bench $W rc=$?"
  python - "$OUT/bench_${W}_$TAG.json" <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"], "chol ms", d["roofline"].get("avg_launch_ms"), d["stage_seconds_per_step"], "parity", d["parity"] and d["parity"]["ok"])
PY
done
timeout 900 python bench.py --no-cpu --no-k3 --no-k5 --no-dense-chol > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
python - <<PY
import json
d = json.load(open("$OUT/bench_$TAG.json"))
print("K4", {k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"], d["stage_seconds_per_step"])
PY
