#!/bin/bash
# critical path of SMALL dense reduced systems (keyframe-BA sizes): K1 = 30 cameras (3 tile columns of 64), K2 = 200 cameras
OUT=gpurun_out; mkdir -p $OUT
( echo "# chol_chain 3 3 0 64 (K1: 30 cameras)"; timeout 120 tools/_build/chol_chain 3 3 0 64
  echo; echo "# chol_chain 2 2 0 128 (K1 in 128-row tiles)"; timeout 120 tools/_build/chol_chain 2 2 0 128
  echo; echo "# chol_chain 20 20 0 64 (K2: 200 cameras, dense co-visibility, 64-row tiles)"; timeout 120 tools/_build/chol_chain 20 20 0 64
  echo; echo "# chol_chain 10 10 0 128 (K2 in 128-row tiles)"; timeout 120 tools/_build/chol_chain 10 10 0 128 ) > $OUT/chol_chain_small.txt 2>&1
cat $OUT/chol_chain_small.txt | cut -c1-140
