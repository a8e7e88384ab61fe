#!/bin/bash
# Build variants queued for the next measurement visit (run here, where nvcc is free; the .so files travel with gpurun):
#   sh64   k_shade in 64-thread CTAs, 8 per SM (same 128 registers / 16 warps, finer-grained release of an SM's slots)
#   sh64b6 the same at 6 CTAs per SM (160 registers: no spills, 12 warps)
#   s12    walks of 12 nodes per look at the warp (3 % better at 1M spheres, 5 % worse at 10k when last swept)
# then on the GPU box: tools/gpu_ab.sh TAG [variant ...]
cd "$(dirname "$0")/.."
bash tools/build_variant.sh sh64 -DRT_SHADE_THREADS=64 -DRT_SHADE_MIN_BLOCKS=8
bash tools/build_variant.sh sh64b6 -DRT_SHADE_THREADS=64 -DRT_SHADE_MIN_BLOCKS=6
bash tools/build_variant.sh s12 -DRT_DESCEND_STEPS=12
ls -la rust-raytrace_b200/lib/
