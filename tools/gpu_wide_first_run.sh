#!/bin/bash
# First hardware run of the experimental four-wide shadow walk (first row).
#   1. the regular parity suite (wide path off): the baseline must be green before anything else is read
#   2. the opt-in wide parity tests, in their own process (a fault there cannot poison the regular run)
#   3. per-class device ms with the walk off, on for shadow queries (1) and on for closest-hit queries too (2), 10k and 1M spheres
# usage: tools/gpu_wide_first_run.sh [TAG]      outputs under gpurun_out/TAG_*
T=${1:-wide}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $O/${T}_pytest_gpu.log
RAYTRACE_B200_TEST_WIDE=1 timeout 600 python -m pytest tests/test_gpu_wide_experimental.py -m gpu -x -q > $O/${T}_pytest_wide.log 2>&1; echo "wide pytest rc=$?"; tail -3 $O/${T}_pytest_wide.log
for W in 10k 1m; do
  timeout 300 python tools/gpu_stats.py $W 2>&1 | grep -v "^heavy" | sed "s/^/[binary] /" | tee -a $O/${T}_ab.txt
  RAYTRACE_B200_BVH_WIDE=1 timeout 300 python tools/gpu_stats.py $W 2>&1 | grep -v "^heavy" | sed "s/^/[wide 1] /" | tee -a $O/${T}_ab.txt
  RAYTRACE_B200_BVH_WIDE=2 timeout 300 python tools/gpu_stats.py $W 2>&1 | grep -v "^heavy" | sed "s/^/[wide 2] /" | tee -a $O/${T}_ab.txt
done
