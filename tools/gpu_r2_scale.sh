#!/bin/bash
# round 2: strong scaling of the bench line (K4 headline + k5 leg, parity gate sharded vs single GPU) at the rank counts given
TAG=${1:-r2scale}; shift; OUT=gpurun_out; mkdir -p $OUT
for N in "$@"; do
  if [ "$N" = "1" ]; then
    timeout 900 python bench.py --no-cpu --no-k3 --no-dense-chol > $OUT/bench_${N}gpu_$TAG.json 2> $OUT/bench_${N}gpu_$TAG.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29540 + N)) bench.py --gpus $N --no-k3 > $OUT/bench_${N}gpu_$TAG.json 2> $OUT/bench_${N}gpu_$TAG.err
  fi
  echo "bench N=$N rc=$?"; grep -E "PARITY|Error|error" $OUT/bench_${N}gpu_$TAG.err | tail -3
  python - "$OUT/bench_${N}gpu_$TAG.json" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print({k: d[k] for k in ("n_gpus", "value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"])
    p = d["parity"]; print(" parity", None if not p else {k: p[k] for k in ("ok", "max_cost_rel_diff", "final_cost_rel_diff", "max_cam_diff", "max_pt_diff", "pt_outside_cond_bound")})
    print(" roofline", {k: d["roofline"][k] for k in ("frac", "avg_launch_ms", "share_of_step")})
    k5 = d["k5"]; print(" k5", {k: k5.get(k) for k in ("value", "ms_per_step", "e2e", "roofline_frac", "roofline_kernel_ms", "cg_iterations_per_step", "gpu_launches", "error")}); print(" k5 parity", k5.get("parity"))
except Exception as e:
    print("bench json unreadable:", e)
PY
done
