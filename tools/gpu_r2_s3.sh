#!/bin/bash
# round 2, session 3 (run with gpurun --gpus 2): tile-packed S + packed collectives: GPU tests on one GPU, Cholesky chain,
# then the bench at 1 and 2 ranks (parity gate sharded vs single, K5 PCG at 2 ranks = the stop-ring fix on > 1 rank)
TAG=${1:-r2s3}; OUT=gpurun_out; mkdir -p $OUT
timeout 300 tools/_build/chol_chain 82 2 > $OUT/chol_chain_$TAG.txt 2>&1; echo "chol_chain rc=$?"; head -3 $OUT/chol_chain_$TAG.txt
timeout 2700 python -m pytest tests -m gpu -q -s > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log
grep -E "^\[|passed|failed|FAILED|need the cond" $OUT/pytest_gpu_$TAG.log | tail -50
summ() { python - "$1" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print({k: d[k] for k in ("n_gpus", "value", "ms_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"])
    print(" parity", d["parity"])
    print(" roofline", {k: d["roofline"][k] for k in ("frac", "avg_launch_ms", "share_of_step")})
    print(" k3", None if not d.get("k3") else {k: d["k3"].get(k) for k in ("value", "ms_per_step", "error")})
    print(" k5", d["k5"])
except Exception as e:
    print("bench json unreadable:", e)
PY
}
timeout 900 python bench.py --no-cpu > $OUT/bench_1gpu_$TAG.json 2> $OUT/bench_1gpu_$TAG.err
echo "bench N=1 rc=$?"; summ $OUT/bench_1gpu_$TAG.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 > $OUT/bench_2gpu_$TAG.json 2> $OUT/bench_2gpu_$TAG.err
echo "bench N=2 rc=$?"; tail -3 $OUT/bench_2gpu_$TAG.err; summ $OUT/bench_2gpu_$TAG.json
