#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT; N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
( timeout 200 $TR --master-port 29601 tools/nccl_probe.py default
  NCCL_ALGO=Ring timeout 200 $TR --master-port 29602 tools/nccl_probe.py ring
  NCCL_ALGO=NVLS timeout 200 $TR --master-port 29603 tools/nccl_probe.py nvls
  NCCL_ALGO=Tree timeout 200 $TR --master-port 29604 tools/nccl_probe.py tree
  NCCL_MIN_NCHANNELS=32 timeout 200 $TR --master-port 29605 tools/nccl_probe.py min32ch ) 2>&1 | grep '^\[' > $OUT/nccl_probe_${N}gpu.txt
cat $OUT/nccl_probe_${N}gpu.txt
