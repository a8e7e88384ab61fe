#!/bin/bash
# device ms per frame (tools/gpu_overlap_ab.py: whole frame and rank 0's shard of 8) for every lib/libraytrace_b200*.so variant
# usage: tools/gpu_shard_ab.sh TAG [workload ...]
O=gpurun_out; T=${1:-shard}; shift
mkdir -p $O
for L in rust-raytrace_b200/lib/libraytrace_b200*.so; do
  V=$(basename $L .so | sed 's/libraytrace_b200_*//'); V=${V:-default}
  RAYTRACE_B200_LIB=$PWD/$L timeout 600 python tools/gpu_overlap_ab.py "$@" 2>&1 | sed "s/^/[$V] /" | tee -a $O/${T}_shard_ab.txt
done
