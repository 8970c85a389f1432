#!/bin/bash
# round 2, session 6: potrf (A) clock breakdown, deterministic persistent PCG matvec: tests + K5 bench + reproducibility
TAG=${1:-r2s6}; OUT=gpurun_out; mkdir -p $OUT
timeout 300 tools/_build/chol_prof 10880 300 64 > $OUT/chol_prof_band64_$TAG.txt 2>&1; cat $OUT/chol_prof_band64_$TAG.txt
timeout 2700 python -m pytest tests -m gpu -q -s > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log
grep -E "^\[K5|^\[K4/pcg|passed|failed|FAILED|^E  " $OUT/pytest_gpu_$TAG.log | tail -30
timeout 900 python bench.py --workload K5 --no-cpu --no-k3 > $OUT/bench_K5_$TAG.json 2> $OUT/bench_K5_$TAG.err
echo "bench K5 rc=$?"; tail -3 $OUT/bench_K5_$TAG.err; python - <<PY
import json
try:
    d = json.load(open("$OUT/bench_K5_$TAG.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"])
    print("roofline", {k: d["roofline"][k] for k in ("achieved", "frac", "avg_launch_ms", "share_of_step", "cg_iterations_per_step")})
    print("stages", d["stage_seconds_per_step"])
except Exception as e:
    print("bench json unreadable:", e)
PY
python - <<'PY'
# bit-reproducibility of the PCG path: two solves of K5 x0.1 from the same state must give identical trajectories and parameters
import numpy as np, bundleadjustment_b200 as bab
from bundleadjustment_b200 import synthetic as syn
pb = syn.flatten(syn.config("K5", scale=0.1))
s = bab.Solver(bab.default_options(max_num_iterations=6, linear_solver=1)); s.set_problem(pb)
s.solve(); a = ([i["cost"] for i in s.iterations()], s.cameras(), s.points())
s.reset_parameters(); s.solve(); b = ([i["cost"] for i in s.iterations()], s.cameras(), s.points())
print("PCG bit-reproducible:", a[0] == b[0], np.array_equal(a[1], b[1]), np.array_equal(a[2], b[2]))
PY
