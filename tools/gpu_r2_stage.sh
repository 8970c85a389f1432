#!/bin/bash
# A/B: linearize_points with the block's distinct camera entries staged in shared memory (lib_B_stage.so, -DLIN_STAGE_CAMS) against the shipped kernel
OUT=gpurun_out; mkdir -p $OUT
cp bundleadjustment_b200/libba_b200.so /tmp/lib_keep.so
for L in tools/_build/lib_A_head.so tools/_build/lib_B_stage.so; do
  cp $L bundleadjustment_b200/libba_b200.so; N=$(basename $L .so)
  if [ "$N" = "lib_B_stage" ]; then timeout 120 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > $OUT/pytest_$N.log 2>&1; echo "$N parity tests: $(tail -1 $OUT/pytest_$N.log)"; fi
  timeout 200 python bench.py --no-cpu --no-k3 --no-k5 --no-dense-chol > $OUT/bench_K4_$N.json 2> $OUT/bench_K4_$N.err
  timeout 300 python bench.py --workload K5 --no-cpu --no-k3 --steps 3 > $OUT/bench_K5_$N.json 2> $OUT/bench_K5_$N.err
  python - "$OUT" "$N" <<'PY'
import json, sys
o, n = sys.argv[1], sys.argv[2]
for w in ("K4", "K5"):
    try:
        d = json.load(open(f"{o}/bench_{w}_{n}.json"))
        print(n, w, {k: round(d[k], 3) for k in ("value", "ms_per_step")}, {k: round(v * 1e3, 3) for k, v in d["stage_seconds_per_step"].items()})
    except Exception as e:
        print(n, w, "unreadable:", e)
PY
done
cp /tmp/lib_keep.so bundleadjustment_b200/libba_b200.so
