"""ctypes wrapper around oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module.  The product (rust-raytrace_b200/) never does.

Scenes are exchanged as plain "scene description" dicts (schema in rust-raytrace_b200/scene.py and
DESIGN.md); `build_scene` feeds one into the oracle's flat builder API.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

RNG_COUNTER = 0
RNG_SEQUENTIAL = 1

MAT_KINDS = {"phong": 0, "indirect": 1, "fresnel": 2, "transparent": 3}
LIGHT_KINDS = {"point": 0, "directional": 1, "area": 2}


class OrcOpts(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("antialias", C.c_uint32), ("max_depth", C.c_uint32),
                ("min_significance", C.c_double), ("seed", C.c_uint64), ("rng_mode", C.c_int32), ("threads", C.c_int32)]


class OrcStats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("shape_tests", C.c_uint64), ("seconds", C.c_double),
                ("threads", C.c_int32), ("_pad", C.c_int32)]


def build_library(force=False):
    """Compile liboracle.so with the recipe in oracle/Makefile (building the checker is not using it)."""
    src = os.path.join(_HERE, "oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        env = dict(os.environ)
        env.pop("CXX", None)
        subprocess.check_call(["make", "-C", _HERE, "-s"], env=env)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build_library()
        L = C.CDLL(_LIB_PATH)
        L.orc_scene_new.restype = C.c_void_p
        L.orc_scene_free.argtypes = [C.c_void_p]
        dp = C.POINTER(C.c_double)
        L.orc_add_material.argtypes = [C.c_void_p, C.c_int, dp, dp, C.c_double, dp, C.c_double, C.c_uint32]
        L.orc_add_spheres.argtypes = [C.c_void_p, C.c_uint64, dp, C.POINTER(C.c_uint32)]
        L.orc_add_plane.argtypes = [C.c_void_p, dp, dp, C.c_uint32]
        L.orc_add_light.argtypes = [C.c_void_p, C.c_int, dp, dp, dp, dp]
        L.orc_set_camera_spc_new.argtypes = [C.c_void_p, dp, dp, dp, C.c_double]
        L.orc_set_camera_spc_look_at.argtypes = [C.c_void_p, dp, dp, dp, C.c_double, C.c_double]
        L.orc_set_camera_dof.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_uint32]
        L.orc_get_camera.argtypes = [C.c_void_p, dp, dp]
        L.orc_set_background_solid.argtypes = [C.c_void_p, dp]
        L.orc_set_skybox_face.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint8)]
        L.orc_set_options.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
        L.orc_bmp_bytewidth.argtypes = [C.c_uint32]
        L.orc_bmp_bytewidth.restype = C.c_uint32
        L.orc_bmp_header.argtypes = [C.c_uint32, C.c_uint32, C.POINTER(C.c_uint8)]
        L.orc_quantize_bgr.argtypes = [dp, C.c_uint64, C.POINTER(C.c_uint8)]
        L.orc_srgb_tables.argtypes = [dp, dp]
        L.orc_trace_rays.argtypes = [C.c_void_p, C.c_uint64, dp, C.POINTER(C.c_int32), dp, dp]
        L.orc_project.argtypes = [C.c_void_p, C.c_uint64, dp, C.c_uint64, dp]
        L.orc_ray_color.argtypes = [C.c_void_p, C.c_uint64, dp, C.c_double, C.c_uint32, C.POINTER(C.c_uint64), C.c_uint32, C.c_double, C.c_int32,
                                    C.POINTER(C.c_uint32), dp, C.POINTER(C.c_uint64)]
        L.orc_ray_color.restype = None
        L.orc_project_sequential.argtypes = [C.c_void_p, C.c_uint64, dp, C.POINTER(C.c_uint32), dp]
        L.orc_project_sequential.restype = None
        L.orc_render.argtypes = [C.c_void_p, C.POINTER(OrcOpts), C.POINTER(C.c_uint8), C.POINTER(C.c_int32), dp, C.POINTER(OrcStats)]
        L.orc_render_pixels.argtypes = [C.c_void_p, C.POINTER(OrcOpts), C.c_uint64, C.POINTER(C.c_uint32), C.POINTER(C.c_uint8),
                                        C.POINTER(C.c_int32), dp, C.POINTER(OrcStats)]
        for f in ("orc_key_sample", "orc_child_key", "orc_draw_bits"):
            getattr(L, f).restype = C.c_uint64
        L.orc_key_sample.argtypes = [C.c_uint64] * 3
        L.orc_child_key.argtypes = [C.c_uint64] * 2
        L.orc_draw_bits.argtypes = [C.c_uint64] * 2
        L.orc_xorshift.argtypes = [C.c_uint32] * 5 + [C.POINTER(C.c_uint32)]
        _lib = L
    return _lib


def _d3(v):
    a = (C.c_double * 3)(*[float(x) for x in v])
    return a


def _ptr(a, ty):
    return a.ctypes.data_as(C.POINTER(ty)) if a is not None else None


class OracleScene:
    """Owns an oracle Scene built from a scene-description dict."""

    def __init__(self, desc):
        L = lib()
        self._h = C.c_void_p(L.orc_scene_new())
        self.desc = desc
        for m in desc["materials"]:
            L.orc_add_material(self._h, MAT_KINDS[m["kind"]], _d3(m.get("diffuse", (0, 0, 0))), _d3(m.get("specular", (0, 0, 0))),
                               float(m.get("exponent", 0.0)), _d3(m.get("ambient", (0, 0, 0))), float(m.get("ior", 1.0)),
                               int(m.get("samples", 0)))
        self.n_objects = 0
        for o in desc["objects"]:
            if o["shape"] == "sphere":
                data = np.array([[*o["center"], o["radius"]]], dtype=np.float64)
                mat = np.array([o["material"]], dtype=np.uint32)
                rc = L.orc_add_spheres(self._h, 1, _ptr(data, C.c_double), _ptr(mat, C.c_uint32))
                self.n_objects += 1
            elif o["shape"] == "spheres":
                data = np.ascontiguousarray(o["data"], dtype=np.float64)
                mat = np.ascontiguousarray(o["material"], dtype=np.uint32)
                rc = L.orc_add_spheres(self._h, data.shape[0], _ptr(data, C.c_double), _ptr(mat, C.c_uint32))
                self.n_objects += data.shape[0]
            elif o["shape"] == "plane":
                rc = L.orc_add_plane(self._h, _d3(o["point"]), _d3(o["normal"]), int(o["material"]))
                self.n_objects += 1
            else:
                raise ValueError(o["shape"])
            if rc != 0:
                raise ValueError("oracle rejected object (bad material index?)")
        for l in desc["lights"]:
            k = l["model"]
            if k == "point":
                L.orc_add_light(self._h, 0, _d3(l["location"]), None, None, _d3(l["color"]))
            elif k == "directional":
                L.orc_add_light(self._h, 1, _d3(l["direction"]), None, None, _d3(l["color"]))
            else:
                L.orc_add_light(self._h, 2, _d3(l["origin"]), _d3(l["side1"]), _d3(l["side2"]), _d3(l["color"]))
        cam = desc["camera"]
        if cam["ctor"] == "new":
            L.orc_set_camera_spc_new(self._h, _d3(cam["position"]), _d3(cam["look"]), _d3(cam["up"]), float(cam["im_dist"]))
        else:
            L.orc_set_camera_spc_look_at(self._h, _d3(cam["focus"]), _d3(cam["look"]), _d3(cam["up"]), float(cam["pov"]), float(cam["h"]))
        if cam.get("dof"):
            d = cam["dof"]
            L.orc_set_camera_dof(self._h, float(d["focus"]), float(d["aperture"]), int(d["samples"]))
        bg = desc["background"]
        if bg["type"] == "solid":
            L.orc_set_background_solid(self._h, _d3(bg["color"]))
        else:
            for i, f in enumerate(bg["faces"]):
                f = np.ascontiguousarray(f, dtype=np.uint8)
                assert f.ndim == 3 and f.shape[2] == 3
                if L.orc_set_skybox_face(self._h, i, f.shape[1], f.shape[0], _ptr(f, C.c_uint8)) != 0:
                    raise ValueError("bad skybox face")
        o = desc["options"]
        L.orc_set_options(self._h, int(o["width"]), int(o["height"]), int(o["antialias"]))
        self.width, self.height, self.antialias = int(o["width"]), int(o["height"]), int(o["antialias"])

    def __del__(self):
        try:
            if self._h:
                lib().orc_scene_free(self._h)
                self._h = None
        except Exception:
            pass

    # ---------------------------------------------------------------------------------------------
    def _opts(self, width=None, height=None, antialias=None, max_depth=4, min_significance=1.0 / 512.0, seed=1,
              rng_mode=RNG_COUNTER, threads=0):
        return OrcOpts(width or self.width, height or self.height, self.antialias if antialias is None else antialias, max_depth,
                       min_significance, seed, rng_mode, threads)

    def camera(self):
        pos = np.zeros(3)
        m = np.zeros(9)
        lib().orc_get_camera(self._h, _ptr(pos, C.c_double), _ptr(m, C.c_double))
        return pos, m.reshape(3, 3)

    def render(self, want_ids=True, want_accum=True, **kw):
        """Full frame.  Returns dict(bgr=(h, bytewidth) u8, ids, accum, stats)."""
        o = self._opts(**kw)
        bw = bytewidth(o.width)
        rows = np.zeros((o.height, bw), dtype=np.uint8)
        ids = np.full((o.height, o.width), -2, dtype=np.int32) if want_ids else None
        acc = np.zeros((o.height, o.width, 3), dtype=np.float64) if want_accum else None
        st = OrcStats()
        rc = lib().orc_render(self._h, C.byref(o), _ptr(rows, C.c_uint8), _ptr(ids, C.c_int32), _ptr(acc, C.c_double), C.byref(st))
        if rc != 0:
            raise RuntimeError("orc_render failed")
        return dict(bgr=rows, ids=ids, accum=acc, stats=_stats(st))

    def render_pixels(self, pixels, **kw):
        """Sparse pixel set (pixel = y*width + x).  Returns dict(bgr=(n,3), ids=(n,), accum=(n,3), stats)."""
        o = self._opts(**kw)
        px = np.ascontiguousarray(pixels, dtype=np.uint32)
        n = px.shape[0]
        bgr = np.zeros((n, 3), dtype=np.uint8)
        ids = np.full(n, -2, dtype=np.int32)
        acc = np.zeros((n, 3), dtype=np.float64)
        st = OrcStats()
        rc = lib().orc_render_pixels(self._h, C.byref(o), n, _ptr(px, C.c_uint32), _ptr(bgr, C.c_uint8), _ptr(ids, C.c_int32),
                                     _ptr(acc, C.c_double), C.byref(st))
        if rc != 0:
            raise RuntimeError("orc_render_pixels failed")
        return dict(bgr=bgr, ids=ids, accum=acc, stats=_stats(st))

    def trace_rays(self, origin_dir):
        od = np.ascontiguousarray(origin_dir, dtype=np.float64).reshape(-1, 6)
        n = od.shape[0]
        ids = np.zeros(n, dtype=np.int32)
        t = np.zeros(n, dtype=np.float64)
        nrm = np.zeros((n, 3), dtype=np.float64)
        lib().orc_trace_rays(self._h, n, _ptr(od, C.c_double), _ptr(ids, C.c_int32), _ptr(t, C.c_double), _ptr(nrm, C.c_double))
        return ids, t, nrm

    def project(self, pos, seed=1):
        p = np.ascontiguousarray(pos, dtype=np.float64).reshape(-1, 2)
        out = np.zeros((p.shape[0], 6), dtype=np.float64)
        lib().orc_project(self._h, p.shape[0], _ptr(p, C.c_double), seed, _ptr(out, C.c_double))
        return out


def _ray_color(self, origin_dir, significance=1.0, depth=0, keys=None, max_depth=4, min_significance=1.0 / 512.0, xorshift_state=None):
    """ray_color (src/raytrace.rs:261-267) of each ray; keys: counter-RNG key per ray, or xorshift_state (x, y, z, w): ONE
    sequential xorshift128 stream threaded through all rays in order, as upstream threads its rng.  Returns (rgb, rays)."""
    od = np.ascontiguousarray(origin_dir, dtype=np.float64).reshape(-1, 6)
    n = od.shape[0]
    out = np.zeros((n, 3), dtype=np.float64)
    rays = C.c_uint64(0)
    k = None if keys is None else np.ascontiguousarray(keys, dtype=np.uint64)
    st = None if xorshift_state is None else (C.c_uint32 * 4)(*[int(v) & 0xFFFFFFFF for v in xorshift_state])
    lib().orc_ray_color(self._h, n, _ptr(od, C.c_double), float(significance), int(depth), _ptr(k, C.c_uint64), int(max_depth), float(min_significance),
                        RNG_SEQUENTIAL if xorshift_state is not None else RNG_COUNTER, st, _ptr(out, C.c_double), C.byref(rays))
    return out, int(rays.value)


def _project_sequential(self, pos, xorshift_state):
    """Camera::project with the reference's sequential rng (src/camera.rs:110-121)."""
    p = np.ascontiguousarray(pos, dtype=np.float64).reshape(-1, 2)
    out = np.zeros((p.shape[0], 6), dtype=np.float64)
    st = (C.c_uint32 * 4)(*[int(v) & 0xFFFFFFFF for v in xorshift_state])
    lib().orc_project_sequential(self._h, p.shape[0], _ptr(p, C.c_double), st, _ptr(out, C.c_double))
    return out


OracleScene.ray_color = _ray_color
OracleScene.project_sequential = _project_sequential


def _stats(st):
    return dict(rays=int(st.rays), shadow_rays=int(st.shadow_rays), shape_tests=int(st.shape_tests), seconds=float(st.seconds),
                threads=int(st.threads))


def bytewidth(width):
    return int(lib().orc_bmp_bytewidth(int(width)))


def bmp_header(width, height):
    buf = np.zeros(122, dtype=np.uint8)
    lib().orc_bmp_header(int(width), int(height), _ptr(buf, C.c_uint8))
    return buf.tobytes()


def quantize_bgr(rgb):
    a = np.ascontiguousarray(rgb, dtype=np.float64).reshape(-1, 3)
    out = np.zeros((a.shape[0], 3), dtype=np.uint8)
    lib().orc_quantize_bgr(_ptr(a, C.c_double), a.shape[0], _ptr(out, C.c_uint8))
    return out


def srgb_tables():
    v = np.zeros(256)
    a = np.zeros(255)
    lib().orc_srgb_tables(_ptr(v, C.c_double), _ptr(a, C.c_double))
    return v, a


def write_bmp(path, width, height, rows):
    with open(path, "wb") as f:
        f.write(bmp_header(width, height))
        f.write(np.ascontiguousarray(rows, dtype=np.uint8).tobytes())
