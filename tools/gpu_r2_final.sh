#!/bin/bash
# round 2, closing session on ONE GPU: the whole GPU suite, the small named workloads (K1, K2) with their CPU samples, the
# wall-clock breakdown of the end-to-end path
TAG=${1:-r02f}; OUT=gpurun_out; mkdir -p $OUT
timeout 2400 python -m pytest tests -m gpu -q -x --durations=5 > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -8 $OUT/pytest_gpu_$TAG.log
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $OUT/pytest_noalloc_alone_$TAG.log 2>&1; echo "no-alloc alone rc=$?"; tail -1 $OUT/pytest_noalloc_alone_$TAG.log
for W in K1 K2; do
  timeout 900 python bench.py --workload $W --steps 20 --no-k3 --no-k5 > $OUT/bench_${W}_$TAG.json 2> $OUT/bench_${W}_$TAG.err; echo "bench $W rc=$?"
  python - "$OUT/bench_${W}_$TAG.json" <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"], "cpu", d["cpu_baseline"] and {k: d["cpu_baseline"].get(k) for k in ("value", "value_4_threads", "value_dense_factorisation")})
print(" parity", d["parity"] and {k: d["parity"].get(k) for k in ("ok", "max_cost_rel_diff", "final_cost_rel_diff", "max_cam_diff", "max_pt_diff", "pt_outside_cond_bound")}, "obs/s", d.get("obs_per_sec"), "lm its", d["config"]["lm_iters_per_step"], d["config"]["termination"])
PY
done
timeout 600 python tools/e2e_breakdown.py > $OUT/e2e_breakdown_$TAG.txt 2>&1; tail -4 $OUT/e2e_breakdown_$TAG.txt
