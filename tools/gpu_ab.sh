#!/bin/bash
# A/B of every lib/libraytrace_b200*.so variant: per-class device ms on the 10k and 1M 4K frames, then the GPU parity
# tests through each variant named on the command line.  usage: tools/gpu_ab.sh TAG [variant ...]
O=gpurun_out; T=${1:-ab}; shift
mkdir -p $O
for L in rust-raytrace_b200/lib/libraytrace_b200*.so; do
  V=$(basename $L .so | sed 's/libraytrace_b200_*//'); V=${V:-default}
  for W in 10k 1m; do
    RAYTRACE_B200_LIB=$PWD/$L timeout 300 python tools/gpu_stats.py $W 2>&1 | grep -v "^heavy" | sed "s/^/[$V] /" | tee -a $O/${T}_ab.txt
  done
done
for V in "$@"; do
  L=$PWD/rust-raytrace_b200/lib/libraytrace_b200_$V.so; [ $V = default ] && L=$PWD/rust-raytrace_b200/lib/libraytrace_b200.so
  RAYTRACE_B200_LIB=$L timeout 900 python -m pytest tests -m gpu -x -q > $O/${T}_pytest_$V.log 2>&1; echo "pytest[$V] rc=$?"; tail -2 $O/${T}_pytest_$V.log
done
