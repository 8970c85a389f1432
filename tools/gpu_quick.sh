#!/bin/bash
# Quick GPU check: tests + short K4/K5 bench lines (no CPU baseline, no dense-Cholesky leg).  Optional env assignments as
# arguments run the benches once more with each of them set (A/B of measurement-aid switches), e.g.  BA_B200_LIN_CTAS=2
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_q.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_gpu_q.log
Q="--no-cpu --no-dense-chol --steps-e2e 1 --warmup-e2e 0"
run() {
  env $1 timeout 300 python bench.py $Q > $OUT/q_K4.json 2>$OUT/q_K4.err; python -c "
import json; d=json.load(open('$OUT/q_K4.json')); s=d['stage_seconds_per_step']; print('K4 [$1]', round(d['value'],1), {k: round(v*1e3,3) for k,v in s.items()}, round(d['obs_per_sec']/1e9,2))"
  env $1 timeout 600 python bench.py --workload K5 --solver pcg --steps 3 $Q > $OUT/q_K5.json 2>$OUT/q_K5.err; python -c "
import json; d=json.load(open('$OUT/q_K5.json')); s=d['stage_seconds_per_step']; print('K5 [$1]', round(d['value'],2), {k: round(v*1e3,2) for k,v in s.items()}, round(d['obs_per_sec']/1e9,2))"
}
run "BA_B200_NOOP=1"
for a in "$@"; do run "$a"; done
