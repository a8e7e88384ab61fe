#!/bin/bash
# every bench workload on one GPU (the BASELINE.md table): usage tools/gpu_workloads.sh TAG
T=${1:-wl}; O=gpurun_out; mkdir -p $O
for W in test-scene-800-1024spp cornell-4k-1spp spheres-1m-4k-d8 skybox-100k-4k-16spp spheres-10k-4k-d8-linear; do
  S=5; [ $W = spheres-10k-4k-d8-linear ] && S=2; [ $W = test-scene-800-1024spp ] && S=3
  timeout 900 python bench.py --workload $W --steps $S --warmup 3 > $O/${T}_n1_$W.json 2> $O/${T}_n1_$W.err; echo "$W rc=$?"
  python - <<PY
import json
d=json.loads(open("$O/${T}_n1_$W.json").read().strip().splitlines()[-1])
r=d["roofline"]
print("  ", d["value"], "Mrays/s", d["ms_per_step"], "ms  e2e", d.get("e2e",{}).get("ms_per_step"), "warm", d.get("e2e_warm",{}).get("ms_per_step"), "|", r["kernel"], r["bound"], round(r["frac"],4), r.get("unit"), "cpu", d.get("cpu_baseline",{}).get("value"))
PY
done
