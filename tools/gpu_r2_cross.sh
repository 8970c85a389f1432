#!/bin/bash
# dense systems: where does the 128-row tiling overtake the 64-row one?  (n = 2048 .. 8192)
OUT=gpurun_out; mkdir -p $OUT
for N64 in 32 48 64 96 128; do
  N128=$((N64 / 2))
  A=$(timeout 120 tools/_build/chol_chain $N64 $N64 0 64 | grep -E '^best' | head -1)
  B=$(timeout 120 tools/_build/chol_chain $N128 $N128 0 128 | grep -E '^best' | head -1)
  echo "n=$((N64 * 64)) dense:  64-row tiles $A   |   128-row tiles $B"
done | tee $OUT/chol_tile_crossover.txt
