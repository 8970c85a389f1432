#!/bin/bash
# Last call of a round: tests, smoke, the three bench lines, and the one ncu capture gpu_session.sh's skip count missed.
TAG=${1:-vX}; OUT=gpurun_out
bash tools/gpu_session.sh $TAG quick
timeout 300 python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -1 $OUT/smoke_$TAG.log
SHORT="--steps 2 --warmup 1 --no-cpu --no-dense-chol --steps-e2e 1 --warmup-e2e 0 --lm-iters 3"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"chol_dataflow" --launch-skip 4 -c 1 -f -o $OUT/prof_K4_chol_$TAG \
    python bench.py $SHORT > /dev/null 2> $OUT/ncu_chol_$TAG.err
python tools/ncu_table.py $OUT/prof_K4_chol_$TAG.ncu-rep > $OUT/prof_K4_chol_$TAG.txt 2>&1; head -8 $OUT/prof_K4_chol_$TAG.txt
