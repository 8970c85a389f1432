#!/bin/bash
# round 2 evidence: bench lines (default run = K4 headline + k3 + k5 + parity + CPU baselines; reference arm), ncu launch lists
# and --set full captures of the dominant kernels.  ncu only runs after the same program has exited 0 without it.
TAG=${1:-r02}; OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python bench.py --steps 20 --warmup 3 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; RC=$?; echo "bench rc=$RC"; tail -2 $OUT/bench_$TAG.err
timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > $OUT/bench_reference_$TAG.json 2> /dev/null; echo "reference rc=$?"
timeout 600 python bench.py --workload K3 > $OUT/bench_K3_$TAG.json 2> $OUT/bench_K3_$TAG.err; echo "K3 rc=$?"
timeout 300 python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -1 $OUT/smoke_$TAG.log
[ $RC -ne 0 ] && exit 0
SHORT="--steps 2 --warmup 1 --no-cpu --no-dense-chol --steps-e2e 1 --warmup-e2e 0 --lm-iters 3 --no-k5 --no-k3"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/launches_K4_$TAG.csv \
    python bench.py $SHORT > /dev/null 2> $OUT/ncu_K4_$TAG.err
python tools/launch_summary.py $OUT/launches_K4_$TAG.csv > $OUT/launch_summary_K4_$TAG.txt 2>&1; head -12 $OUT/launch_summary_K4_$TAG.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/launches_K5_$TAG.csv \
    python bench.py --workload K5 $SHORT > /dev/null 2> $OUT/ncu_K5_$TAG.err
python tools/launch_summary.py $OUT/launches_K5_$TAG.csv > $OUT/launch_summary_K5_$TAG.txt 2>&1; head -12 $OUT/launch_summary_K5_$TAG.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"chol_dataflow" --launch-skip 4 -c 1 -f -o $OUT/prof_K4_chol_$TAG \
    python bench.py $SHORT > /dev/null 2>> $OUT/ncu_K4_$TAG.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"linearize_points|linearize_cams|backsub_kernel|schur_units|eval_candidate" \
    --launch-skip 10 -c 5 -f -o $OUT/prof_K4_lm_$TAG python bench.py $SHORT > /dev/null 2>> $OUT/ncu_K4_$TAG.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"linearize_points|pcg_setup_cams|pcg_matvec|backsub_kernel" \
    --launch-skip 6 -c 4 -f -o $OUT/prof_K5_lm_$TAG python bench.py --workload K5 $SHORT > /dev/null 2>> $OUT/ncu_K5_$TAG.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"pose_batch_kernel|prune_kernel" --launch-skip 8 -c 2 -f -o $OUT/prof_K3_$TAG \
    python bench.py --workload K3 --no-cpu --steps 2 --warmup 1 > /dev/null 2> $OUT/ncu_K3_$TAG.err
for r in $OUT/prof_*_$TAG.ncu-rep; do [ -f "$r" ] && python tools/ncu_table.py $r > ${r%.ncu-rep}.txt 2>&1; done
ls -la $OUT | tail -20
