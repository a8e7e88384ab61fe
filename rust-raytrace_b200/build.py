"""Build recipe for libraytrace_b200.so (sm_100a only; nvcc cross-compiles without a GPU).

    python rust-raytrace_b200/build.py [--force] [--verbose]

kernels.cu is compiled twice from the same source:
    -DRT_NS=strict -fmad=false   no FMA contraction: the reference's unfused x86-64 arithmetic, bit for bit
    -DRT_NS=fast   -fmad=true    contraction allowed
The CUDA runtime is linked statically, so the library loads (and exports every symbol of include/raytrace_b200.h)
on a machine without a GPU or libcudart; render entry points then fail with RT_ERR_CUDA.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libraytrace_b200.so")
CLI = os.path.join(LIBDIR, "raytrace_b200")  # the reference's binary (src/main.rs) over the C ABI: csrc/cli.cpp
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_COMMON = ["-std=c++17", "-O3", "-lineinfo", "-Xcompiler", "-fPIC,-fvisibility=hidden,-ffp-contract=off", "-Xptxas", "-v",
               "-Wno-deprecated-gpu-targets"] + ARCH
HOST_FLAGS = ["-std=c++17", "-O2", "-fPIC", "-fvisibility=hidden", "-ffp-contract=off", "-fno-fast-math", "-Wall", "-Wextra",
              "-Wno-unused-parameter", "-Wno-missing-field-initializers"]


def _nvcc():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found")
    return p


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd, verbose, log):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    log.append("$ " + " ".join(cmd) + "\n" + r.stdout)
    if verbose:
        print("$ " + " ".join(cmd))
        print(r.stdout)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("build step failed: " + " ".join(cmd))


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    nvcc = _nvcc()
    headers = [os.path.join(CSRC, h) for h in ("rt_types.h", "rt_device.cuh", "kernels.h", "host_scene.h")] + [
        os.path.join(INCLUDE, "raytrace_b200.h"), os.path.join(INCLUDE, "rt_srgb_tables.inc"), os.path.abspath(__file__)]
    log = []
    objs = []
    steps = [
        ("kernels_strict.o", "kernels.cu", [nvcc] + NVCC_COMMON + ["-DRT_NS=strict", "-fmad=false"]),
        ("kernels_fast.o", "kernels.cu", [nvcc] + NVCC_COMMON + ["-DRT_NS=fast", "-fmad=true"]),
        ("bvh_build.o", "bvh_build.cu", [nvcc] + NVCC_COMMON + ["-fmad=false"]),
        ("context.o", "context.cu", [nvcc] + NVCC_COMMON),
        ("peaks.o", "peaks.cu", [nvcc] + NVCC_COMMON + ["-fmad=true"]),
        ("noalloc.o", "noalloc.cu", [nvcc] + NVCC_COMMON + ["-DRT_NS=strict", "-fmad=false"]),
        ("host_scene.o", "host_scene.cpp", ["g++"] + HOST_FLAGS),
        ("serialize.o", "serialize.cpp", ["g++"] + HOST_FLAGS),
    ]
    rebuilt = False
    for obj, src, cmd in steps:
        o = os.path.join(OBJ, obj)
        s = os.path.join(CSRC, src)
        objs.append(o)
        if force or _newer(o, [s] + headers):
            _run(cmd + ["-I", INCLUDE, "-c", s, "-o", o], verbose, log)
            rebuilt = True
    if rebuilt or not os.path.exists(LIB):
        _run([nvcc, "-shared", "-o", LIB] + objs + ARCH + ["-cudart", "static", "-Xlinker", "--no-undefined", "-lpthread", "-ldl", "-lrt"], verbose, log)
    cli_src = os.path.join(CSRC, "cli.cpp")
    if rebuilt or _newer(CLI, [cli_src, os.path.join(INCLUDE, "raytrace_b200.h")]):
        _run(["g++"] + [f for f in HOST_FLAGS if f != "-fPIC"] + ["-I", INCLUDE, cli_src, "-o", CLI, "-L", LIBDIR, "-lraytrace_b200",
                                                                  "-Wl,-rpath,$ORIGIN"], verbose, log)
    with open(os.path.join(OBJ, "build.log"), "a" if not rebuilt else "w") as f:
        f.write("\n".join(log))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
