#!/bin/bash
# bench.py at N GPUs through torchrun (what the driver runs), outputs under gpurun_out/: usage tools/gpu_scale.sh TAG N
T=${1:-scale}; N=${2:-8}; O=gpurun_out; mkdir -p $O
if [ "$N" = 1 ]; then
  python bench.py > $O/${T}_n1_bench.json 2> $O/${T}_n1_bench.err
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > $O/${T}_n${N}_bench.json 2> $O/${T}_n${N}_bench.err
fi
echo "rc=$?"; tail -c 300 $O/${T}_n${N}_bench.err
python - <<PY
import json
d=json.loads(open("$O/${T}_n${N}_bench.json").read().strip().splitlines()[-1])
a=d.pop("also",{})
fc=d.get("frame_check") or {}
print("$N GPUs:", round(d["value"],1), "Mrays/s", round(d["ms_per_step"],3), "ms | frame_check", fc.get("max_lsb"), fc.get("identical"), "| e2e cold", round(d["e2e"]["ms_per_step"],2), "warm", round(d["e2e_warm"]["ms_per_step"],2), "| nccl", (d.get("gather_nccl") or {}).get("ms_per_step"))
for k,v in a.items():
    fc=v.get("frame_check") or {}
    print("   ", k, round(v["value"],1), round(v["ms_per_step"],3), "ms | frame_check", fc.get("max_lsb"), fc.get("identical"), "| e2e cold", round(v["e2e"]["ms_per_step"],2), "warm", round(v["e2e_warm"]["ms_per_step"],2), "|", v["wavefront"], "| cpu", (v.get("cpu_baseline") or {}).get("value"))
PY
