#!/bin/bash
# ncu --set full with source correlation for one launch of a kernel of the default bench frame.
# usage: tools/gpu_ncu_src.sh TAG KERNEL SKIP
O=gpurun_out; T=$1; K=$2; S=$3
mkdir -p $O
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c 1 -o $O/${T}_full_$K -f python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-also > $O/${T}_ncu_full_$K.log 2>&1; echo "ncu full $K rc=$?"
ls -la $O/${T}_full_$K.ncu-rep
