#!/bin/bash
# where a keyframe-BA-sized LM iteration (K1, K2) spends its time: ncu launch list of a short bench run
OUT=gpurun_out; mkdir -p $OUT
SHORT="--steps 2 --warmup 1 --no-cpu --no-dense-chol --steps-e2e 1 --warmup-e2e 0 --no-k5 --no-k3"
for W in K1 K2; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/launches_${W}_small.csv python bench.py --workload $W $SHORT > /dev/null 2> $OUT/ncu_${W}_small.err
  python tools/launch_summary.py $OUT/launches_${W}_small.csv > $OUT/launch_summary_${W}_small.txt 2>&1; head -32 $OUT/launch_summary_${W}_small.txt
done
