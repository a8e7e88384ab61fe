#!/bin/bash
# builds lib/libraytrace_b200_<name>.so with extra -D flags for kernels.cu (kernel tuning sweeps)
# usage: tools/build_variant.sh NAME -DRT_X=1 -DRT_Y=2        (tools/build_variant.sh checked -DRT_CHECK_BOUNDS: see rt_device.cuh)
NAME=$1; shift
cd "$(dirname "$0")/../rust-raytrace_b200"
B=build/var_$NAME; mkdir -p $B lib
NV="nvcc -std=c++17 -O3 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden,-ffp-contract=off -Wno-deprecated-gpu-targets -gencode arch=compute_100a,code=sm_100a -I ../include"
$NV -DRT_NS=strict -fmad=false "$@" -c csrc/kernels.cu -o $B/ks.o &
$NV -DRT_NS=fast -fmad=true "$@" -c csrc/kernels.cu -o $B/kf.o &
wait
nvcc -shared -o lib/libraytrace_b200_$NAME.so $B/ks.o $B/kf.o build/bvh_build.o build/context.o build/peaks.o build/noalloc.o build/host_scene.o build/serialize.o -gencode arch=compute_100a,code=sm_100a -cudart static -Xlinker --no-undefined -lpthread -ldl -lrt && echo built $NAME
