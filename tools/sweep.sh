#!/bin/bash
# times every lib/libraytrace_b200_*.so variant (plus the default build) on the 10k workload
for L in rust-raytrace_b200/lib/libraytrace_b200*.so; do
  echo "== $L"
  RAYTRACE_B200_LIB=$PWD/$L python tools/gpu_quick.py ${1:-10k} 2>&1 | grep "bvh strict"
done
