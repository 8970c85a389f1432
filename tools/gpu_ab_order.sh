#!/bin/bash
# A/B of BA_B200_ORDER_MASK (which kernels walk the landmark blocks in camera-sorted processing order); see ba_solver.cuh ORDER_*.
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_ab.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_gpu_ab.log
BA_B200_ORDER_MASK=31 timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_ab31.log 2>&1; echo "pytest(mask 31) rc=$?"; tail -3 $OUT/pytest_gpu_ab31.log
Q="--no-cpu --no-dense-chol --steps-e2e 1 --warmup-e2e 0"
for m in 24 0 25 26 8 16; do
  BA_B200_ORDER_MASK=$m timeout 300 python bench.py $Q > $OUT/ab_K4_m$m.json 2>/dev/null
  python - $OUT/ab_K4_m$m.json $m <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); s=d['stage_seconds_per_step']
    print(f"K4 mask {sys.argv[2]:>2}: {d['value']:.1f} it/s  lin {s['linearize']*1e3:.3f} ms  solve {s['linear_solve']*1e3:.3f} ms  rest {(s['total']-s['linearize']-s['linear_solve'])*1e3:.3f} ms  eval {d['obs_per_sec']/1e9:.2f} Gobs/s")
except Exception as e: print('K4 mask', sys.argv[2], 'failed', e)
PY
done
for m in 24 0 25 26 28; do
  BA_B200_ORDER_MASK=$m timeout 600 python bench.py --workload K5 --solver pcg --steps 3 $Q > $OUT/ab_K5_m$m.json 2>/dev/null
  python - $OUT/ab_K5_m$m.json $m <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); s=d['stage_seconds_per_step']
    print(f"K5 mask {sys.argv[2]:>2}: {d['value']:.2f} it/s  lin {s['linearize']*1e3:.2f} ms  solve {s['linear_solve']*1e3:.2f} ms  rest {(s['total']-s['linearize']-s['linear_solve'])*1e3:.2f} ms  eval {d['obs_per_sec']/1e9:.2f} Gobs/s")
except Exception as e: print('K5 mask', sys.argv[2], 'failed', e)
PY
done
