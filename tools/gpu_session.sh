#!/bin/bash
# One gpurun call: GPU tests, bench lines (K4 dense, K5 PCG), ncu launch lists and full captures of the large kernels.
# Usage (from the repo root on the GPU box): bash tools/gpu_session.sh <tag> [quick]
# Every command runs under its own timeout; ncu only runs after the same program has exited 0 without it.
TAG=${1:-vX}
QUICK=${2:-}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi -L > $OUT/gpu_$TAG.txt 2>&1

timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log
tail -3 $OUT/pytest_gpu_$TAG.log

timeout 600 python bench.py > $OUT/bench_K4_$TAG.json 2> $OUT/bench_K4_$TAG.err
RC4=$?; echo "bench K4 rc=$RC4"; cat $OUT/bench_K4_$TAG.json | cut -c1-400
timeout 900 python bench.py --workload K5 --solver pcg --steps 3 > $OUT/bench_K5_$TAG.json 2> $OUT/bench_K5_$TAG.err
RC5=$?; echo "bench K5 rc=$RC5"; cat $OUT/bench_K5_$TAG.json | cut -c1-400
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_K4_reference_$TAG.json 2> /dev/null
echo "reference arm rc=$?"

[ -n "$QUICK" ] && exit 0

SHORT="--steps 2 --warmup 1 --no-cpu --no-dense-chol --steps-e2e 1 --warmup-e2e 0 --lm-iters 3"
if [ $RC4 -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/launches_K4_$TAG.csv \
      python bench.py $SHORT > /dev/null 2> $OUT/ncu_K4_$TAG.err
  python tools/launch_summary.py $OUT/launches_K4_$TAG.csv > $OUT/launch_summary_K4_$TAG.txt 2>&1; head -12 $OUT/launch_summary_K4_$TAG.txt
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"linearize_points|linearize_cams|backsub_kernel|schur_units|eval_candidate" \
      --launch-skip 10 -c 5 -f -o $OUT/prof_K4_lm_$TAG python bench.py $SHORT > /dev/null 2>> $OUT/ncu_K4_$TAG.err
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"evaluate_materialize" --launch-skip 4 -c 1 -f -o $OUT/prof_K4_eval_$TAG \
      python bench.py $SHORT > /dev/null 2>> $OUT/ncu_K4_$TAG.err
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"chol_dataflow" --launch-skip 4 -c 1 -f -o $OUT/prof_K4_chol_$TAG \
      python bench.py $SHORT > /dev/null 2>> $OUT/ncu_K4_$TAG.err
fi
if [ $RC5 -eq 0 ]; then
  K5="--workload K5 --solver pcg"
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/launches_K5_$TAG.csv \
      python bench.py $K5 $SHORT > /dev/null 2> $OUT/ncu_K5_$TAG.err
  python tools/launch_summary.py $OUT/launches_K5_$TAG.csv > $OUT/launch_summary_K5_$TAG.txt 2>&1; head -12 $OUT/launch_summary_K5_$TAG.txt
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"linearize_points|linearize_cams|pcg_setup_cams|pcg_matvec|backsub_kernel|eval_candidate" \
      --launch-skip 8 -c 6 -f -o $OUT/prof_K5_lm_$TAG python bench.py $K5 $SHORT > /dev/null 2>> $OUT/ncu_K5_$TAG.err
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"evaluate_materialize" --launch-skip 4 -c 1 -f -o $OUT/prof_K5_eval_$TAG \
      python bench.py $K5 $SHORT > /dev/null 2>> $OUT/ncu_K5_$TAG.err
fi
for r in $OUT/prof_*_$TAG.ncu-rep; do [ -f "$r" ] && python tools/ncu_table.py $r > ${r%.ncu-rep}.txt 2>&1; done
ls -la $OUT | tail -30
