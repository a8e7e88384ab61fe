#!/bin/bash
# scaling visit: torchrun bench at N GPUs for the default workload and the 1M-sphere 8K config (no reference arm)
N=${1:-8}; TAG=${2:-scale}
O=gpurun_out; mkdir -p $O
for WL in spheres-10k-4k-d8 spheres-1m-8k-d8 skybox-100k-4k-16spp; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --workload $WL --no-cpu-baseline > $O/${TAG}_n${N}_$WL.json 2> $O/${TAG}_n${N}_$WL.err; echo "rc=$?"; grep "^{" $O/${TAG}_n${N}_$WL.json | cut -c1-330
done
