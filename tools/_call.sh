for L in rust-raytrace_b200/lib/libraytrace_b200*.so; do
  V=$(basename $L .so | sed 's/libraytrace_b200_*//'); V=${V:-default}
  for W in 10k 1m; do
    RAYTRACE_B200_LIB=$PWD/$L timeout 300 python tools/gpu_stats.py $W 2>&1 | grep "render" | sed "s/^/[$V] /" | cut -c1-150
  done
done
