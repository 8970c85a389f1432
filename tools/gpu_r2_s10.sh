#!/bin/bash
# round 2, session 10: Cholesky chain of K4's tile pattern (64- vs 128-row tiles), PCG matvec modes on K5 (2 = atomic, 3 = folded atomic)
TAG=${1:-r2s10}; OUT=gpurun_out; mkdir -p $OUT
( echo "# tools/chol_chain 170 4 0 64  (K4's pattern: 170 tile rows of 64, band 4 tiles)"; timeout 300 tools/_build/chol_chain 170 4 0 64
  echo; echo "# tools/chol_chain 82 2 0 128  (the same matrix in 128-row tiles)"; timeout 300 tools/_build/chol_chain 82 2 0 128 ) > $OUT/chol_chain_$TAG.txt 2>&1
head -40 $OUT/chol_chain_$TAG.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "pcg" > $OUT/pytest_pcg_$TAG.log 2>&1; tail -3 $OUT/pytest_pcg_$TAG.log
for M in 2 3; do
  timeout 900 python bench.py --workload K5 --no-cpu --no-k3 --pcg-mode $M --steps 5 > $OUT/bench_K5_mode${M}_$TAG.json 2> $OUT/bench_K5_mode${M}_$TAG.err
  python - <<PY
import json
d = json.load(open("$OUT/bench_K5_mode${M}_$TAG.json"))
print("mode $M", {k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"])
print("   roofline", {k: d["roofline"].get(k) for k in ("achieved", "frac", "avg_launch_ms", "share_of_step")})
PY
done
