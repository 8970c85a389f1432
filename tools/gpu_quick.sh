OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_q.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_gpu_q.log
Q="--no-cpu --no-dense-chol --steps-e2e 1 --warmup-e2e 0"
timeout 300 python bench.py $Q > $OUT/q_K4.json 2>$OUT/q_K4.err; python -c "
import json; d=json.load(open('$OUT/q_K4.json')); s=d['stage_seconds_per_step']; print('K4', d['value'], s, d['obs_per_sec'])"
timeout 600 python bench.py --workload K5 --solver pcg --steps 3 $Q > $OUT/q_K5.json 2>$OUT/q_K5.err; python -c "
import json; d=json.load(open('$OUT/q_K5.json')); s=d['stage_seconds_per_step']; print('K5', d['value'], s, d['obs_per_sec'])"
