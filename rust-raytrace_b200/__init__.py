"""raytrace-b200: B200-native tracing path of j-dong/rust-raytrace behind the crate's own API surface.

The module names mirror the upstream crate's `lib.rs` (src/lib.rs:4-13):

    serialize.deserialize(text)          -> Scene           src/serialize.rs:427
    scene.*  (description helpers)                          src/scene.rs, src/shapes.rs, src/camera.rs
    raytrace.render(scene, ...)          -> Frame           src/main.rs:39-59 over src/raytrace.rs:270
    raytrace.intersect(scene, rays)                         src/scene.rs:247
    bmp.write(path, w, h, rows) / bmp.header(w, h)          src/bmp.rs:10

Everything below these calls runs in `lib/libraytrace_b200.so` (hand-written sm_100a CUDA).  There is no CPU
fallback: without the built library the import fails, and without a GPU every render call raises RtError
(RT_ERR_CUDA).

The directory is called `rust-raytrace_b200` (not an identifier); import it with
`importlib.import_module("rust-raytrace_b200")` or through the `raytrace_b200` alias module at the repo root.
"""
import ctypes as C
import os

import numpy as np

from . import _binding as B
from ._binding import (ACCEL_AUTO, ACCEL_BVH, ACCEL_LINEAR, FP_FAST, FP_STRICT, GATHER_AUTO, GATHER_DIRECT, GATHER_NCCL,  # noqa: F401
                       GATHER_PEER, NOALLOC_MAX_SPHERES, NOALLOC_SCRATCH_BYTES, TUNE_BINARY_SHADOW, TUNE_HOST_LEVELS, TUNE_MORTON_AXIS, TUNE_NODES_F32, TUNE_NODES_Q16, TUNE_FRAME_GRAPH, TUNE_SERIAL_SHADOW, RenderOpts,
                       RtError, Stats, SyntaxError_)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RAYTRACE_B200_LIB") or os.path.join(_HERE, "lib", "libraytrace_b200.so")  # env: development builds only

_lib = None


def lib():
    """The loaded C-ABI library.  Fails loudly when it has not been built (python rust-raytrace_b200/build.py)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libraytrace_b200.so is missing: run `python rust-raytrace_b200/build.py` (needs nvcc); "
                              "there is no pure-Python or CPU fallback")
        _lib = B.declare(C.CDLL(LIB_PATH))
    return _lib


def device_count():
    return int(lib().rt_device_count())


def release_cached_memory():
    """Returns the device blocks cached from destroyed contexts to the driver (rt_release_cached_memory)."""
    lib().rt_release_cached_memory()


def measure_peaks(device=0):
    """FP64 / FP32 FMA throughput and L2 read bandwidth of `device`, measured now (rt_measure_peaks)."""
    p = B.Peaks()
    B.check(lib(), lib().rt_measure_peaks(device, C.byref(p)))
    return {"fp64_tflops": p.fp64_tflops, "fp32_tflops": p.fp32_tflops, "l2_read_gbs": p.l2_read_gbs, "l2_buffer_bytes": int(p.l2_buffer_bytes),
            "sm_count": int(p.sm_count)}


def version():
    return lib().rt_version().decode()


class Scene(B.Scene):
    """`scene::Scene` (src/scene.rs:201-212): built from a description dict or parsed from scene text."""

    def __init__(self, desc=None, handle=None):
        super().__init__(lib(), desc=desc, handle=handle)

    @classmethod
    def parse(cls, text):
        s = B.Scene.parse(lib(), text)
        out = cls(handle=s._h.value)
        s._h = None
        return out


class Frame:
    """Result of a render: BGR rows in BMP order (row 0 = bottom), primary hit ids, linear accumulators, stats."""

    def __init__(self, width, height, rows, ids, accum, stats):
        self.width, self.height, self.rows, self.ids, self.accum, self.stats = width, height, rows, ids, accum, stats

    @property
    def bgr(self):
        """(height, width, 3) view of the pixel bytes (padding stripped)."""
        return self.rows[:, : 3 * self.width].reshape(self.height, self.width, 3)

    def save(self, path):
        bmp.write(path, self.width, self.height, self.rows)


class Context:
    """Device-resident scene + buffers (`rt_context`): render repeatedly without re-uploading."""

    def __init__(self, scene, **opts):
        self._lib = lib()
        self.scene = scene
        self.opts = B.make_opts(self._lib, **opts)
        h = C.c_void_p()
        B.check(self._lib, self._lib.rt_context_create(scene.handle, C.byref(self.opts), C.byref(h)))
        self._h = h
        st = Stats()
        B.check(self._lib, self._lib.rt_context_stats(self._h, C.byref(st)))
        w, hh, _ = scene.options()
        self.width = self.opts.width or w
        self.height = self.opts.height or hh
        self.local_rows, self.bytewidth = st.local_rows, st.bytewidth
        self.create_stats = st.as_dict()

    def close(self):
        if getattr(self, "_h", None):
            self._lib.rt_context_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def render(self):
        st = Stats()
        B.check(self._lib, self._lib.rt_context_render(self._h, C.byref(st)))
        return st.as_dict()

    def render_begin(self):
        """Queue one frame and return while the GPU works; `render_end` collects it (several contexts = several frames in flight)."""
        B.check(self._lib, self._lib.rt_context_render_begin(self._h))

    def render_end(self):
        st = Stats()
        B.check(self._lib, self._lib.rt_context_render_end(self._h, C.byref(st)))
        return st.as_dict()

    def framebuffer_ptr(self):
        return int(self._lib.rt_context_framebuffer(self._h)), int(self._lib.rt_context_framebuffer_bytes(self._h))

    def download(self, want_rows=True, want_ids=False, want_accum=False):
        rows = np.zeros((self.local_rows, self.bytewidth), dtype=np.uint8) if want_rows else None
        ids = np.zeros((self.local_rows, self.width), dtype=np.int32) if want_ids else None
        acc = np.zeros((self.local_rows, self.width, 3), dtype=np.float64) if want_accum else None
        B.check(self._lib, self._lib.rt_context_download(self._h, B.ptr(rows, C.c_uint8), B.ptr(ids, C.c_int32), B.ptr(acc, C.c_double)))
        return rows, ids, acc

    def download_rows_to(self, host_ptr):
        """Copies the local framebuffer (local_rows * bytewidth bytes) to caller-owned host memory (e.g. a pinned buffer)."""
        B.check(self._lib, self._lib.rt_context_download(self._h, C.cast(C.c_void_p(host_ptr), C.POINTER(C.c_uint8)), None, None))

    def untile(self, gathered_ptr, out_ptr):
        B.check(self._lib, self._lib.rt_context_untile(self._h, C.c_void_p(gathered_ptr), C.c_void_p(out_ptr)))

    def synchronize(self):
        B.check(self._lib, self._lib.rt_context_synchronize(self._h))

    def frame_create(self):
        """Whole-frame buffer on this context's device; this context now resolves its rows into it.  Returns its address."""
        p = C.c_void_p()
        B.check(self._lib, self._lib.rt_context_frame_create(self._h, C.byref(p)))
        return int(p.value)

    def frame_export(self):
        """64-byte CUDA IPC handle of the frame (for the other ranks' frame_import)."""
        h = np.zeros(64, dtype=np.uint8)
        B.check(self._lib, self._lib.rt_context_frame_export(self._h, B.ptr(h, C.c_uint8)))
        return h

    def frame_import(self, handle):
        """Resolve into another process's frame (rank 0's): the gather becomes this GPU's own stores over NVLink."""
        h = np.ascontiguousarray(handle, dtype=np.uint8)
        assert h.size == 64
        B.check(self._lib, self._lib.rt_context_frame_import(self._h, B.ptr(h, C.c_uint8)))

    def frame_set(self, dev_ptr):
        B.check(self._lib, self._lib.rt_context_frame_set(self._h, C.c_void_p(dev_ptr or 0)))

    def set_stream(self, cuda_stream):
        """Run on a caller-owned stream (int handle, e.g. torch.cuda.current_stream().cuda_stream); None = own stream."""
        B.check(self._lib, self._lib.rt_context_set_stream(self._h, C.c_void_p(cuda_stream or 0)))

    def set_profiling(self, on=True):
        B.check(self._lib, self._lib.rt_context_set_profiling(self._h, 1 if on else 0))


class MultiContext:
    """`rt_multi`: one frame sharded by row bands over several GPUs of the node, inside the library (one host thread and
    stream per device); `gather` picks how the BGR bands become one frame (GATHER_DIRECT / GATHER_PEER / GATHER_NCCL)."""

    def __init__(self, scene, devices, gather=GATHER_AUTO, **opts):
        self._lib = lib()
        self.scene = scene
        self.opts = B.make_opts(self._lib, **opts)
        self.devices = list(range(devices)) if isinstance(devices, int) else list(devices)
        arr = (C.c_int * len(self.devices))(*self.devices)
        h = C.c_void_p()
        B.check(self._lib, self._lib.rt_multi_create(scene.handle, C.byref(self.opts), len(self.devices), arr, int(gather), C.byref(h)))
        self._h = h
        w, hh, _ = scene.options()
        self.width = self.opts.width or w
        self.height = self.opts.height or hh
        self.bytewidth = B.bmp_bytewidth(self._lib, self.width)
        self.gather = int(self._lib.rt_multi_gather_used(self._h))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.rt_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def render(self, want_ids=False):
        rows = np.zeros((self.height, self.bytewidth), dtype=np.uint8)
        ids = np.zeros((self.height, self.width), dtype=np.int32) if want_ids else None
        st = Stats()
        B.check(self._lib, self._lib.rt_multi_render(self._h, B.ptr(rows, C.c_uint8), B.ptr(ids, C.c_int32), C.byref(st)))
        return Frame(self.width, self.height, rows, ids, None, st.as_dict())

    def render_to(self, host_ptr):
        """One frame into caller-owned host memory (bytewidth*height bytes, e.g. a pinned buffer); host_ptr None: the frame
        stays in device memory (GATHER_PEER / GATHER_NCCL, see frame_device_ptr)."""
        st = Stats()
        dst = C.cast(C.c_void_p(host_ptr), C.POINTER(C.c_uint8)) if host_ptr else None
        B.check(self._lib, self._lib.rt_multi_render(self._h, dst, None, C.byref(st)))
        return st.as_dict()

    def device_stats(self, index):
        st = Stats()
        B.check(self._lib, self._lib.rt_multi_device_stats(self._h, index, C.byref(st)))
        return st.as_dict()

    def frame_device_ptr(self, index=0):
        return int(self._lib.rt_multi_frame_device(self._h, index) or 0)


class _Serialize:
    """`serialize` module (src/serialize.rs)."""

    SyntaxError = SyntaxError_

    @staticmethod
    def deserialize(text):
        return Scene.parse(text)


class _Raytrace:
    """`raytrace` module: the frame loop of src/main.rs:39-59 over raytrace::raytrace (src/raytrace.rs:270)."""

    MIN_SIGNIFICANCE = 1.0 / 256.0 / 2.0  # src/raytrace.rs:17
    MAX_DEPTH = 4  # src/raytrace.rs:18

    @staticmethod
    def render(scene, want_ids=True, want_accum=False, **opts):
        L = lib()
        o = B.make_opts(L, **opts)
        w, h, _ = scene.options()
        width, height = o.width or w, o.height or h
        bw = B.bmp_bytewidth(L, width)
        rows = np.zeros((height, bw), dtype=np.uint8)
        ids = np.zeros((height, width), dtype=np.int32) if want_ids else None
        acc = np.zeros((height, width, 3), dtype=np.float64) if want_accum else None
        st = Stats()
        B.check(L, L.rt_render(scene.handle, C.byref(o), B.ptr(rows, C.c_uint8), B.ptr(ids, C.c_int32), B.ptr(acc, C.c_double), C.byref(st)))
        return Frame(width, height, rows, ids, acc, st.as_dict())

    @staticmethod
    def render_into(scene, host_ptr, devices=None, gather=GATHER_AUTO, **opts):
        """rt_render / rt_render_multi with the BGR rows going to caller-owned host memory (bytewidth*height bytes at
        `host_ptr`, e.g. a pinned buffer).  Returns the stats."""
        L = lib()
        o = B.make_opts(L, **opts)
        st = Stats()
        dst = C.cast(C.c_void_p(host_ptr), C.POINTER(C.c_uint8))
        if devices is None:
            B.check(L, L.rt_render(scene.handle, C.byref(o), dst, None, None, C.byref(st)))
        else:
            devs = list(range(devices)) if isinstance(devices, int) else list(devices)
            B.check(L, L.rt_render_multi(scene.handle, C.byref(o), len(devs), (C.c_int * len(devs))(*devs), int(gather), dst, None, C.byref(st)))
        return st.as_dict()

    @staticmethod
    def render_multi(scene, devices, gather=GATHER_AUTO, want_ids=True, **opts):
        """rt_render_multi: the same frame, sharded over `devices` (a count or a list of ordinals)."""
        L = lib()
        o = B.make_opts(L, **opts)
        devs = list(range(devices)) if isinstance(devices, int) else list(devices)
        w, h, _ = scene.options()
        width, height = o.width or w, o.height or h
        bw = B.bmp_bytewidth(L, width)
        rows = np.zeros((height, bw), dtype=np.uint8)
        ids = np.zeros((height, width), dtype=np.int32) if want_ids else None
        st = Stats()
        B.check(L, L.rt_render_multi(scene.handle, C.byref(o), len(devs), (C.c_int * len(devs))(*devs), int(gather), B.ptr(rows, C.c_uint8),
                                     B.ptr(ids, C.c_int32), C.byref(st)))
        return Frame(width, height, rows, ids, None, st.as_dict())

    @staticmethod
    def intersect(scene, origin_dir, **opts):
        """Scene::intersect (src/scene.rs:247) for a batch of rays: returns (ids, t, normal)."""
        L = lib()
        o = B.make_opts(L, **opts)
        od = np.ascontiguousarray(origin_dir, dtype=np.float64).reshape(-1, 6)
        n = od.shape[0]
        ids = np.zeros(n, dtype=np.int32)
        t = np.zeros(n, dtype=np.float64)
        nrm = np.zeros((n, 3), dtype=np.float64)
        B.check(L, L.rt_trace_rays(scene.handle, C.byref(o), n, B.ptr(od, C.c_double), B.ptr(ids, C.c_int32), B.ptr(t, C.c_double), B.ptr(nrm, C.c_double)))
        return ids, t, nrm

    @staticmethod
    def project(scene, pos, **opts):
        """Camera::project (src/camera.rs:77, :110) for a batch of normalised image positions."""
        L = lib()
        o = B.make_opts(L, **opts)
        p = np.ascontiguousarray(pos, dtype=np.float64).reshape(-1, 2)
        out = np.zeros((p.shape[0], 6), dtype=np.float64)
        B.check(L, L.rt_project(scene.handle, C.byref(o), p.shape[0], B.ptr(p, C.c_double), B.ptr(out, C.c_double)))
        return out


class _Color:
    """`color` module: the quantiser of src/color.rs:593-632."""

    @staticmethod
    def write_bgr(rgb, device=0):
        L = lib()
        a = np.ascontiguousarray(rgb, dtype=np.float64).reshape(-1, 3)
        out = np.zeros((a.shape[0], 3), dtype=np.uint8)
        B.check(L, L.rt_quantize_bgr(B.ptr(a, C.c_double), a.shape[0], B.ptr(out, C.c_uint8), device))
        return out


class _Bmp:
    """`bmp` module (src/bmp.rs)."""

    @staticmethod
    def bytewidth(width):
        return B.bmp_bytewidth(lib(), width)

    @staticmethod
    def header(width, height):
        return B.bmp_header(lib(), width, height)

    @staticmethod
    def write(path, width, height, rows):
        L = lib()
        r = np.ascontiguousarray(rows, dtype=np.uint8)
        if r.size != B.bmp_bytewidth(L, width) * height:
            raise ValueError("rows must hold bytewidth*height bytes")
        B.check(L, L.rt_bmp_write(os.fsencode(path), int(width), int(height), B.ptr(r, C.c_uint8)))


serialize = _Serialize
raytrace = _Raytrace
color = _Color
bmp = _Bmp

from . import scene  # noqa: E402,F401  (description helpers mirroring scene.rs / shapes.rs / camera.rs)
