#!/bin/bash
# One GPU-box visit : parity of the new build, A/B of the shadow stack, context-creation phases at 1M spheres.
O=gpurun_out; T=s16
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/${T}_smi.txt 2>&1; nproc >> $O/${T}_smi.txt
echo "== A/B shadow stack (per-class ms, 10k and 1m)"
for L in "" _stack64; do
  for W in 10k 1m; do
    RAYTRACE_B200_LIB=$PWD/rust-raytrace_b200/lib/libraytrace_b200$L.so timeout 300 python tools/gpu_stats.py $W 2>&1 | sed "s/^/[lib$L] /" | tee -a $O/${T}_ab.txt
  done
done
echo "== context creation phases"
timeout 300 python tools/gpu_build_probe.py 1m 6 2>&1 | tee $O/${T}_build_probe_1m.txt
timeout 300 python tools/gpu_build_probe.py 10k 6 2>&1 | tee $O/${T}_build_probe_10k.txt
timeout 300 python tools/gpu_e2e_phases.py 1m 2>&1 | tee $O/${T}_e2e_phases_1m.txt
timeout 300 python tools/gpu_e2e_phases.py 10k 2>&1 | tee $O/${T}_e2e_phases_10k.txt
echo "== build kernels under ncu (launch list)"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 150 --csv --log-file $O/${T}_build_launches_1m.csv python tools/gpu_build_probe.py 1m 1 > $O/${T}_build_ncu.log 2>&1; echo "ncu rc=$?"
echo "== pytest -m gpu"
timeout 1200 python -m pytest tests -m gpu -x -q > $O/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${T}_pytest_gpu.log
tail -5 $O/${T}_pytest_gpu.log
echo "== bench"
timeout 600 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench rc=$?"
cat $O/${T}_bench.json
timeout 600 python bench.py --workload spheres-1m-4k-d8 --no-cpu-baseline > $O/${T}_bench_1m4k.json 2> $O/${T}_bench_1m4k.err; echo "bench 1m rc=$?"
cat $O/${T}_bench_1m4k.json
