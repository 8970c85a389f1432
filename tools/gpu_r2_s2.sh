#!/bin/bash
# round 2, session 2: full GPU test suite, the new bench line (K4 headline + k3 + k5 + parity + cpu baselines), reference arm
TAG=${1:-r2s2}; OUT=gpurun_out; mkdir -p $OUT
timeout 2700 python -m pytest tests -m gpu -q -s > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log
grep -E "^\[|landmark|passed|failed|Error|error|FAILED" $OUT/pytest_gpu_$TAG.log | tail -70
timeout 1500 python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; tail -5 $OUT/bench_$TAG.err; python - <<PY
import json
try:
    d = json.load(open("$OUT/bench_$TAG.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"], "parity", d["parity"])
    print("roofline", {k: d["roofline"][k] for k in ("achieved", "frac", "avg_launch_ms", "share_of_step")})
    print("cpu", d["cpu_baseline"])
    print("all_tiles_filled", d.get("all_tiles_filled"))
    print("k3", d["k3"])
    print("k5", d["k5"])
except Exception as e:
    print("bench json unreadable:", e)
PY
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_reference_$TAG.json 2> $OUT/bench_reference_$TAG.err
echo "reference rc=$?"; cut -c1-900 $OUT/bench_reference_$TAG.json
