"""Test-only CPU emulation of the wavefront formulation (see emu.cpp header and README.md in this directory).

Builds tests/emu/libemu.so with g++ from the product's headers + host sources and exposes the same call shapes
as the product binding, so a test can run `emu.render(desc)` next to `oracle.render(...)` without a GPU.
"""
import ctypes as C
import importlib
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(os.path.dirname(_HERE))
_LIB = os.path.join(_HERE, "libemu.so")
_CSRC = os.path.join(_ROOT, "rust-raytrace_b200", "csrc")

B = importlib.import_module("rust-raytrace_b200._binding")

_lib = None


def _sources():
    return [os.path.join(_HERE, "emu.cpp"), os.path.join(_CSRC, "host_scene.cpp"), os.path.join(_CSRC, "serialize.cpp")]


def _deps():
    return _sources() + [os.path.join(_CSRC, h) for h in ("rt_device.cuh", "rt_types.h", "host_scene.h")] + [
        os.path.join(_ROOT, "include", "raytrace_b200.h")]


def build(force=False):
    if force or not os.path.exists(_LIB) or any(os.path.getmtime(d) > os.path.getmtime(_LIB) for d in _deps()):
        cmd = ["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-Wl,-Bsymbolic", "-I",
               os.path.join(_ROOT, "include")] + _sources() + ["-o", _LIB]
        subprocess.check_call(cmd)
    return _LIB


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB)
        B.declare(L, B.SCENE_API)
        dp, i32p, u8p = C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_uint8)
        L.emu_render_shard.argtypes = [C.c_void_p, C.POINTER(B.RenderOpts), u8p, i32p, dp, C.POINTER(B.Stats)]
        L.emu_trace_rays.argtypes = [C.c_void_p, C.POINTER(B.RenderOpts), C.c_uint64, dp, i32p, dp, dp]
        L.emu_shadow_rays.argtypes = [C.c_void_p, C.POINTER(B.RenderOpts), C.c_uint64, dp, dp, u8p]
        L.emu_untile.argtypes = [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)] + [C.c_uint32] * 5
        L.emu_untile.restype = None
        L.emu_quantize_bgr.argtypes = [dp, C.c_uint64, u8p]
        L.emu_quantize_bgr.restype = None
        L.emu_bvh_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]
        L.emu_flat_cache_probe.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_uint64)]
        _lib = L
    return _lib


class Scene(B.Scene):
    def __init__(self, desc=None, handle=None):
        super().__init__(lib(), desc=desc, handle=handle)

    @classmethod
    def parse(cls, text):
        s = B.Scene.parse(lib(), text)
        out = cls(handle=s._h.value)
        s._h = None
        return out


def render_shard(scene, **opts):
    L = lib()
    o = B.make_opts(L, **opts)
    w, h, _ = scene.options()
    width, height = o.width or w, o.height or h
    band = o.band_rows or 4
    n_bands = (height + band - 1) // band
    local_rows = ((n_bands + o.world - 1) // o.world) * band
    bw = B.bmp_bytewidth(L, width)
    rows = np.zeros((local_rows, bw), dtype=np.uint8)
    ids = np.zeros((local_rows, width), dtype=np.int32)
    acc = np.zeros((local_rows, width, 3), dtype=np.float64)
    st = B.Stats()
    B.check(L, L.emu_render_shard(scene.handle, C.byref(o), B.ptr(rows, C.c_uint8), B.ptr(ids, C.c_int32), B.ptr(acc, C.c_double), C.byref(st)))
    return dict(rows=rows, ids=ids, accum=acc, stats=st.as_dict(), width=width, height=height, local_rows=local_rows, bytewidth=bw)


def render(scene, **opts):
    """Whole frame (world = 1); arrays cropped to the image height."""
    r = render_shard(scene, **opts)
    h = r["height"]
    return dict(bgr=r["rows"][:h], ids=r["ids"][:h], accum=r["accum"][:h], stats=r["stats"])


def trace_rays(scene, origin_dir, **opts):
    L = lib()
    o = B.make_opts(L, **opts)
    od = np.ascontiguousarray(origin_dir, dtype=np.float64).reshape(-1, 6)
    n = od.shape[0]
    ids = np.zeros(n, dtype=np.int32)
    t = np.zeros(n)
    nrm = np.zeros((n, 3))
    B.check(L, L.emu_trace_rays(scene.handle, C.byref(o), n, B.ptr(od, C.c_double), B.ptr(ids, C.c_int32), B.ptr(t, C.c_double), B.ptr(nrm, C.c_double)))
    return ids, t, nrm


def shadow_rays(scene, origin_dir, r2, **opts):
    L = lib()
    o = B.make_opts(L, **opts)
    od = np.ascontiguousarray(origin_dir, dtype=np.float64).reshape(-1, 6)
    rr = np.ascontiguousarray(r2, dtype=np.float64)
    occ = np.zeros(od.shape[0], dtype=np.uint8)
    B.check(L, L.emu_shadow_rays(scene.handle, C.byref(o), od.shape[0], B.ptr(od, C.c_double), B.ptr(rr, C.c_double), B.ptr(occ, C.c_uint8)))
    return occ.astype(bool)


def untile(gathered, height, bytewidth, band_rows, world, local_rows):
    g = np.ascontiguousarray(gathered, dtype=np.uint8).view(np.uint32)
    out = np.zeros(height * bytewidth // 4, dtype=np.uint32)
    lib().emu_untile(B.ptr(g, C.c_uint32), B.ptr(out, C.c_uint32), height, bytewidth // 4, band_rows, world, local_rows)
    return out.view(np.uint8).reshape(height, bytewidth)


def quantize_bgr(rgb):
    a = np.ascontiguousarray(rgb, dtype=np.float64).reshape(-1, 3)
    out = np.zeros((a.shape[0], 3), dtype=np.uint8)
    lib().emu_quantize_bgr(B.ptr(a, C.c_double), a.shape[0], B.ptr(out, C.c_uint8))
    return out


def bvh_stats(scene):
    n, d = C.c_uint64(), C.c_uint32()
    B.check(lib(), lib().emu_bvh_stats(scene.handle, C.byref(n), C.byref(d)))
    return n.value, d.value


def bvh_node_bytes(scene, tuning=0):
    """bytes per BVH node the product would use for this scene: 32 = grid form, 64 = f32 planes"""
    b = C.c_uint32()
    B.check(lib(), lib().emu_bvh_node_bytes(scene.handle, C.c_uint32(tuning), C.byref(b)))
    return b.value
