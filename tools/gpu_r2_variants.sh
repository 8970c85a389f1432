#!/bin/bash
# A/B runs of library variants built into tools/_build/lib_*.so (experiments only): K4 bench line per variant
OUT=gpurun_out; mkdir -p $OUT
cp bundleadjustment_b200/libba_b200.so /tmp/lib_keep.so
for L in tools/_build/lib_*.so; do
  cp $L bundleadjustment_b200/libba_b200.so
  N=$(basename $L .so)
  timeout 600 python bench.py --no-cpu --no-k3 --no-k5 --no-dense-chol ${BENCH_ARGS} > $OUT/bench_$N.json 2> $OUT/bench_$N.err
  python - "$OUT/bench_$N.json" "$N" <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print(sys.argv[2], {k: d[k] for k in ("value", "ms_per_step")}, d["stage_seconds_per_step"], "chol", d["roofline"].get("avg_launch_ms"), "parity", d["parity"])
PY
done
cp /tmp/lib_keep.so bundleadjustment_b200/libba_b200.so
