#!/bin/bash
# multi-GPU visit: torchrun bench at N GPUs (default workload + the 1M-sphere 8K config)
N=${1:-2}; TAG=${2:-multi}
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/${TAG}_smi.txt
for WL in spheres-10k-4k-d8 spheres-1m-8k-d8; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --workload $WL > $O/${TAG}_n${N}_$WL.json 2> $O/${TAG}_n${N}_$WL.err; echo "rc=$?"; cat $O/${TAG}_n${N}_$WL.json | cut -c1-600
done
timeout 900 python bench.py --gpus 1 --steps 5 --warmup 3 --workload spheres-1m-8k-d8 --no-cpu-baseline > $O/${TAG}_n1_spheres-1m-8k-d8.json 2> $O/${TAG}_n1_1m.err; echo "rc=$?"; cat $O/${TAG}_n1_spheres-1m-8k-d8.json | cut -c1-600
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_reference.json 2>&1; cat $O/${TAG}_reference.json | cut -c1-400
