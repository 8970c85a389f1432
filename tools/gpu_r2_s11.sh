#!/bin/bash
# round 2, session 11: potrf phase (A) as two pipelined warps + all-warp epilogue
TAG=${1:-r2s11}; OUT=gpurun_out; mkdir -p $OUT
for H in 0 1 2; do echo "## POTRF_HANDOFF=$H"; timeout 300 tools/_build/chol_prof_h$H 10880 300 64; done > $OUT/chol_prof_band64_$TAG.txt 2>&1; cat $OUT/chol_prof_band64_$TAG.txt
timeout 300 tools/_build/chol_chain 170 4 0 64 > $OUT/chol_chain64_$TAG.txt 2>&1; head -12 $OUT/chol_chain64_$TAG.txt; tail -1 $OUT/chol_chain64_$TAG.txt
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge_cases.py -m gpu -q -x > $OUT/pytest_parity_$TAG.log 2>&1; tail -3 $OUT/pytest_parity_$TAG.log
timeout 900 python bench.py --no-cpu --no-k3 --no-k5 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
python - <<PY
import json
d = json.load(open("$OUT/bench_$TAG.json"))
print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"], d["stage_seconds_per_step"])
print("roofline", {k: d["roofline"][k] for k in ("achieved", "frac", "avg_launch_ms", "share_of_step")})
print("dense chol", d.get("roofline_dense_cholesky", {}).get("frac"), "all tiles", d.get("all_tiles_filled"))
PY
