#!/bin/bash
# One gpurun --gpus N call.  N == 2: sharded-vs-single parity check (tools/mgpu_check.py) + bench lines; N > 2: bench lines only.
# Usage: bash tools/gpu_session_mgpu.sh <tag> <N>
TAG=${1:-vX}; N=${2:-8}
OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
if [ "$N" -le 4 ]; then      # 2 and 4 ranks: the 4-rank run is the regression test of the stop-ring fix (f6de604)
  timeout 600 $TR --nproc-per-node $N --master-port 29511 tools/mgpu_check.py > $OUT/mgpu_check_${N}gpu_$TAG.log 2>&1; echo "mgpu_check rc=$?"; grep '"case"' $OUT/mgpu_check_${N}gpu_$TAG.log | cut -c1-220
fi
Q="--no-cpu --no-dense-chol"
timeout 900 $TR --nproc-per-node $N --master-port 29512 bench.py --gpus $N $Q > $OUT/bench_K4_${N}gpu_$TAG.json 2> $OUT/bench_K4_${N}gpu_$TAG.err; echo "K4 x$N rc=$?"; cut -c1-200 $OUT/bench_K4_${N}gpu_$TAG.json
timeout 900 $TR --nproc-per-node $N --master-port 29513 bench.py --gpus $N --workload K5 --solver pcg --steps 3 $Q --no-k3 > $OUT/bench_K5_${N}gpu_$TAG.json 2> $OUT/bench_K5_${N}gpu_$TAG.err; echo "K5 x$N rc=$?"; cut -c1-200 $OUT/bench_K5_${N}gpu_$TAG.json
