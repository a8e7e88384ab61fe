#!/bin/bash
# One GPU-box visit: parity tests, bench line, ncu launch list, ncu full captures of the top kernels.
# usage: tools/gpu_check.sh TAG   (outputs under gpurun_out/TAG_*)
TAG=${1:-run}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/${TAG}_smi.txt 2>&1
nproc >> $O/${TAG}_smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest_gpu.log
tail -3 $O/${TAG}_pytest_gpu.log
timeout 600 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
cat $O/${TAG}_bench.json
if [ "$2" != "noncu" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-also > $O/${TAG}_ncu_list.log 2>&1; echo "ncu list rc=$?"
# the same frame with one frame-wide shadow launch and one frame at a time (what the per-class timing pass and the full captures below see)
SERIAL="--tuning 32 --frames-in-flight 1"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches_serial.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-also $SERIAL > $O/${TAG}_ncu_list_serial.log 2>&1; echo "ncu list (serial) rc=$?"
for K in k_shadow_bvh k_closest_bvh k_shade; do
  SKIP=12; [ $K = k_shadow_bvh ] && SKIP=2   # one shadow launch per frame, ten of the others
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -o $O/${TAG}_full_$K -f python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-also $SERIAL > $O/${TAG}_ncu_full_$K.log 2>&1; echo "ncu full $K rc=$?"
done
fi
