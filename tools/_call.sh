bash tools/gpu_check.sh s16h
for W in spheres-1m-4k-d8 spheres-1m-8k-d8 skybox-100k-4k-16spp test-scene-800-1024spp cornell-4k-1spp; do
  timeout 600 python bench.py --workload $W --no-cpu-baseline > gpurun_out/s16h_bench_$W.json 2> gpurun_out/s16h_bench_$W.err; echo "bench $W rc=$?"; cut -c1-400 gpurun_out/s16h_bench_$W.json
done
