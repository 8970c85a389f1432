#!/bin/bash
# round 2, session 7: tests, K4 bench (cp.async-staged schur_units), ncu --set full of the new PCG matvec and of schur_units
TAG=${1:-r2s7}; OUT=gpurun_out; mkdir -p $OUT
timeout 2700 python -m pytest tests -m gpu -q -s > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log
grep -E "passed|failed|FAILED|^E  " $OUT/pytest_gpu_$TAG.log | tail -20
timeout 900 python bench.py --no-cpu --no-k5 --no-k3 --no-dense-chol > $OUT/bench_K4_$TAG.json 2> $OUT/bench_K4_$TAG.err
echo "bench K4 rc=$?"; python - <<PY
import json
try:
    d = json.load(open("$OUT/bench_K4_$TAG.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"], d["stage_seconds_per_step"])
except Exception as e:
    print("bench json unreadable:", e)
PY
SHORT="--steps 2 --warmup 1 --no-cpu --no-dense-chol --steps-e2e 1 --warmup-e2e 0 --lm-iters 3 --no-k5 --no-k3"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"pcg_matvec_points" --launch-skip 6 -c 1 -f -o $OUT/prof_K5_matvec_$TAG \
    python bench.py --workload K5 $SHORT > /dev/null 2> $OUT/ncu_K5_$TAG.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"schur_units" --launch-skip 4 -c 1 -f -o $OUT/prof_K4_schur_$TAG \
    python bench.py $SHORT > /dev/null 2> $OUT/ncu_K4_$TAG.err
for r in $OUT/prof_*_$TAG.ncu-rep; do [ -f "$r" ] && python tools/ncu_table.py $r > ${r%.ncu-rep}.txt 2>&1; done
ls -la $OUT | tail -8
