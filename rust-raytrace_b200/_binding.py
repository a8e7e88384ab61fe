"""ctypes declarations for the C ABI in include/raytrace_b200.h and the Scene wrapper built on them.

This is the Python stand-in for the Rust shim a maintainer would write: it only marshals
arguments.  All tracing happens inside libraytrace_b200.so on the GPU.
"""
import ctypes as C

import numpy as np

RT_OK, RT_ERR_INVALID, RT_ERR_SYNTAX, RT_ERR_IO, RT_ERR_CUDA, RT_ERR_UNSUPPORTED = 0, -1, -2, -3, -4, -5
ACCEL_AUTO, ACCEL_LINEAR, ACCEL_BVH = 0, 1, 2
FP_STRICT, FP_FAST = 0, 1
GATHER_AUTO, GATHER_DIRECT, GATHER_PEER, GATHER_NCCL = 0, 1, 2, 3
ANTIALIAS_FROM_SCENE = 0xFFFFFFFF
TUNE_NODES_F32, TUNE_NODES_Q16, TUNE_MORTON_AXIS, TUNE_BINARY_SHADOW, TUNE_HOST_LEVELS, TUNE_SERIAL_SHADOW, TUNE_FRAME_GRAPH = 1, 2, 4, 8, 16, 32, 64
NOALLOC_MAX_SPHERES, NOALLOC_SCRATCH_BYTES = 65536, 65536 * 32 + 256   # rt_intersect_spheres_noalloc's static scratch


class RenderOpts(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("antialias", C.c_uint32), ("max_depth", C.c_uint32),
                ("min_significance", C.c_double), ("seed", C.c_uint64), ("accel", C.c_int32), ("fp_mode", C.c_int32),
                ("device", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32), ("band_rows", C.c_uint32),
                ("max_rays_in_flight", C.c_uint64), ("tuning", C.c_uint32), ("_reserved", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("rays_total", C.c_uint64), ("rays_primary", C.c_uint64), ("rays_secondary", C.c_uint64), ("rays_shadow", C.c_uint64),
                ("paths", C.c_uint64), ("kernel_launches", C.c_uint64), ("render_ms", C.c_double), ("bvh_build_ms", C.c_double),
                ("n_bvh_nodes", C.c_uint64), ("local_rows", C.c_uint32), ("bytewidth", C.c_uint32), ("accel_used", C.c_int32),
                ("fp_mode_used", C.c_int32), ("bvh_nodes_visited", C.c_uint64 * 2), ("bvh_prims_tested", C.c_uint64 * 2),
                ("class_ms", C.c_double * 5), ("class_launches", C.c_uint64 * 5), ("heavy_rays", C.c_uint64 * 2), ("bvh_node_bytes", C.c_uint64),
                ("host_levels", C.c_uint32), ("device_driven_overflows", C.c_uint32), ("scan_sphere_tests", C.c_uint64 * 2)]

    CLASSES = ("primary", "closest", "shade", "shadow", "resolve")

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_ if not k.startswith(("class_", "bvh_nodes_", "bvh_prims_", "heavy_", "scan_"))}
        d["scan_sphere_tests"] = {"closest": int(self.scan_sphere_tests[0]), "shadow": int(self.scan_sphere_tests[1])}
        d["heavy_rays"] = {"closest": int(self.heavy_rays[0]), "shadow": int(self.heavy_rays[1])}
        d["bvh_nodes_visited"] = {"closest": int(self.bvh_nodes_visited[0]), "shadow": int(self.bvh_nodes_visited[1])}
        d["bvh_prims_tested"] = {"closest": int(self.bvh_prims_tested[0]), "shadow": int(self.bvh_prims_tested[1])}
        d["class_ms"] = dict(zip(self.CLASSES, [float(x) for x in self.class_ms]))
        d["class_launches"] = dict(zip(self.CLASSES, [int(x) for x in self.class_launches]))
        return d


class Peaks(C.Structure):
    _fields_ = [("fp64_tflops", C.c_double), ("fp32_tflops", C.c_double), ("l2_read_gbs", C.c_double), ("l2_buffer_bytes", C.c_uint64),
                ("sm_count", C.c_int32), ("_pad", C.c_int32)]


class SceneInfo(C.Structure):
    _fields_ = [("n_objects", C.c_uint64), ("n_spheres", C.c_uint64), ("n_planes", C.c_uint64), ("n_materials", C.c_uint64),
                ("n_lights", C.c_uint64), ("camera_kind", C.c_int32), ("background_kind", C.c_int32)]


class RtError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("rt error %d: %s" % (code, message))
        self.code, self.message = code, message


class SyntaxError_(RtError):
    """serialize::SyntaxError -- `message` is upstream's `row:col: text`."""


_dp = C.POINTER(C.c_double)
_u8p = C.POINTER(C.c_uint8)
_u32p = C.POINTER(C.c_uint32)
_i32p = C.POINTER(C.c_int32)

# every symbol include/raytrace_b200.h declares: name -> (restype, argtypes)
API = {
    "rt_last_error": (C.c_char_p, []),
    "rt_version": (C.c_char_p, []),
    "rt_device_count": (C.c_int, []),
    "rt_release_cached_memory": (None, []),
    "rt_scene_new": (C.c_void_p, []),
    "rt_scene_free": (None, [C.c_void_p]),
    "rt_scene_parse": (C.c_int, [C.c_char_p, C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p, C.c_size_t]),
    "rt_scene_parse_file": (C.c_int, [C.c_char_p, C.POINTER(C.c_void_p), C.c_char_p, C.c_size_t]),
    "rt_scene_warnings": (C.c_char_p, [C.c_void_p]),
    "rt_scene_add_material_phong": (C.c_int, [C.c_void_p, _dp, _dp, C.c_double, _dp]),
    "rt_scene_add_material_indirect": (C.c_int, [C.c_void_p, _dp, _dp, C.c_double, _dp, C.c_uint32]),
    "rt_scene_add_material_fresnel": (C.c_int, [C.c_void_p, _dp, _dp, C.c_double, _dp, C.c_double]),
    "rt_scene_add_material_transparent": (C.c_int, [C.c_void_p, _dp, C.c_double, C.c_double]),
    "rt_scene_add_spheres": (C.c_int, [C.c_void_p, C.c_uint64, _dp, _u32p]),
    "rt_scene_add_plane": (C.c_int, [C.c_void_p, _dp, _dp, C.c_uint32]),
    "rt_scene_update_spheres": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint64, _dp]),
    "rt_scene_add_light_point": (C.c_int, [C.c_void_p, _dp, _dp]),
    "rt_scene_add_light_directional": (C.c_int, [C.c_void_p, _dp, _dp]),
    "rt_scene_add_light_area": (C.c_int, [C.c_void_p, _dp, _dp, _dp, _dp]),
    "rt_scene_set_camera_spc_new": (C.c_int, [C.c_void_p, _dp, _dp, _dp, C.c_double]),
    "rt_scene_set_camera_spc_look_at": (C.c_int, [C.c_void_p, _dp, _dp, _dp, C.c_double, C.c_double]),
    "rt_scene_set_camera_dof": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_uint32]),
    "rt_scene_get_camera": (C.c_int, [C.c_void_p, _dp, _dp]),
    "rt_scene_set_background_solid": (C.c_int, [C.c_void_p, _dp]),
    "rt_scene_set_skybox_face": (C.c_int, [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, _u8p]),
    "rt_scene_load_skybox_face": (C.c_int, [C.c_void_p, C.c_int, C.c_char_p]),
    "rt_scene_set_options": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]),
    "rt_scene_get_options": (C.c_int, [C.c_void_p, _u32p, _u32p, _u32p]),
    "rt_scene_get_info": (C.c_int, [C.c_void_p, C.POINTER(SceneInfo)]),
    "rt_scene_dump": (C.c_size_t, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "rt_render_opts_default": (None, [C.POINTER(RenderOpts)]),
    "rt_render": (C.c_int, [C.c_void_p, C.POINTER(RenderOpts), _u8p, _i32p, _dp, C.POINTER(Stats)]),
    "rt_context_create": (C.c_int, [C.c_void_p, C.POINTER(RenderOpts), C.POINTER(C.c_void_p)]),
    "rt_context_destroy": (None, [C.c_void_p]),
    "rt_context_render": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "rt_context_render_begin": (C.c_int, [C.c_void_p]),
    "rt_context_render_end": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "rt_context_stats": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "rt_context_framebuffer": (C.c_void_p, [C.c_void_p]),
    "rt_context_framebuffer_bytes": (C.c_size_t, [C.c_void_p]),
    "rt_context_download": (C.c_int, [C.c_void_p, _u8p, _i32p, _dp]),
    "rt_context_untile": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "rt_context_synchronize": (C.c_int, [C.c_void_p]),
    "rt_context_frame_create": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "rt_context_frame_export": (C.c_int, [C.c_void_p, _u8p]),
    "rt_context_frame_import": (C.c_int, [C.c_void_p, _u8p]),
    "rt_context_frame_set": (C.c_int, [C.c_void_p, C.c_void_p]),
    "rt_context_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "rt_context_set_profiling": (C.c_int, [C.c_void_p, C.c_int]),
    "rt_multi_create": (C.c_int, [C.c_void_p, C.POINTER(RenderOpts), C.c_int, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]),
    "rt_multi_destroy": (None, [C.c_void_p]),
    "rt_multi_render": (C.c_int, [C.c_void_p, _u8p, _i32p, C.POINTER(Stats)]),
    "rt_multi_device_stats": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(Stats)]),
    "rt_multi_frame_device": (C.c_void_p, [C.c_void_p, C.c_int]),
    "rt_multi_gather_used": (C.c_int, [C.c_void_p]),
    "rt_render_multi": (C.c_int, [C.c_void_p, C.POINTER(RenderOpts), C.c_int, C.POINTER(C.c_int), C.c_int, _u8p, _i32p, C.POINTER(Stats)]),
    "rt_trace_rays": (C.c_int, [C.c_void_p, C.POINTER(RenderOpts), C.c_uint64, _dp, _i32p, _dp, _dp]),
    "rt_intersect_spheres_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    "rt_project": (C.c_int, [C.c_void_p, C.POINTER(RenderOpts), C.c_uint64, _dp, _dp]),
    "rt_quantize_bgr": (C.c_int, [_dp, C.c_uint64, _u8p, C.c_int]),
    "rt_measure_peaks": (C.c_int, [C.c_int, C.POINTER(Peaks)]),
    "rt_bmp_bytewidth": (C.c_uint32, [C.c_uint32]),
    "rt_bmp_header": (None, [C.c_uint32, C.c_uint32, _u8p]),
    "rt_bmp_write": (C.c_int, [C.c_char_p, C.c_uint32, C.c_uint32, _u8p]),
}

# the scene-construction subset (also exported by the test-only emulation library)
SCENE_API = [k for k in API if k.startswith("rt_scene_") or k in ("rt_last_error", "rt_render_opts_default", "rt_bmp_bytewidth", "rt_bmp_header",
                                                                   "rt_bmp_write", "rt_version")]


def declare(lib, names=None):
    for name in (names or API):
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = API[name]
    return lib


def d3(v):
    return (C.c_double * 3)(*[float(x) for x in v])


def ptr(a, ty):
    return a.ctypes.data_as(C.POINTER(ty)) if a is not None else None


def check(lib, rc):
    if rc < 0:
        msg = (lib.rt_last_error() or b"").decode("utf-8", "replace")
        raise (SyntaxError_ if rc == RT_ERR_SYNTAX else RtError)(rc, msg)
    return rc


def make_opts(lib, **kw):
    o = RenderOpts()
    lib.rt_render_opts_default(C.byref(o))
    for k, v in kw.items():
        if v is None:
            continue
        if not hasattr(o, k):
            raise TypeError("unknown render option %r" % k)
        setattr(o, k, v)
    return o


class Scene:
    """Owns an `rt_scene*`.  Construct from a scene-description dict (scene.py) or with Scene.parse(text)."""

    def __init__(self, lib, desc=None, handle=None):
        self._lib = lib
        self._h = None
        if handle is not None:
            self._h = C.c_void_p(handle)
            return
        self._h = C.c_void_p(lib.rt_scene_new())
        if desc is not None:
            self._fill(desc)

    @classmethod
    def parse(cls, lib, text):
        """serialize::deserialize (src/serialize.rs:427)."""
        raw = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        out = C.c_void_p()
        err = C.create_string_buffer(1024)
        rc = lib.rt_scene_parse(raw, len(raw), C.byref(out), err, len(err))
        if rc != RT_OK:
            raise (SyntaxError_ if rc == RT_ERR_SYNTAX else RtError)(rc, err.value.decode("utf-8", "replace"))
        return cls(lib, handle=out.value)

    def close(self):
        if self._h:
            self._lib.rt_scene_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    def _fill(self, desc):
        L, h = self._lib, self._h
        for m in desc["materials"]:
            k = m["kind"]
            dif, spc, amb = d3(m.get("diffuse", (0, 0, 0))), d3(m.get("specular", (0, 0, 0))), d3(m.get("ambient", (0, 0, 0)))
            e = float(m.get("exponent", 0.0))
            if k == "phong":
                check(L, L.rt_scene_add_material_phong(h, dif, spc, e, amb))
            elif k == "indirect":
                check(L, L.rt_scene_add_material_indirect(h, dif, spc, e, amb, int(m["samples"])))
            elif k == "fresnel":
                check(L, L.rt_scene_add_material_fresnel(h, dif, spc, e, amb, float(m["ior"])))
            elif k == "transparent":
                check(L, L.rt_scene_add_material_transparent(h, spc, e, float(m["ior"])))
            else:
                raise ValueError("unknown material kind %r" % k)
        for o in desc["objects"]:
            sh = o["shape"]
            if sh == "sphere":
                data = np.array([[*o["center"], o["radius"]]], dtype=np.float64)
                mat = np.array([o["material"]], dtype=np.uint32)
                check(L, L.rt_scene_add_spheres(h, 1, ptr(data, C.c_double), ptr(mat, C.c_uint32)))
            elif sh == "spheres":
                data = np.ascontiguousarray(o["data"], dtype=np.float64)
                mat = np.ascontiguousarray(o["material"], dtype=np.uint32)
                assert data.ndim == 2 and data.shape[1] == 4 and mat.shape == (data.shape[0],)
                check(L, L.rt_scene_add_spheres(h, data.shape[0], ptr(data, C.c_double), ptr(mat, C.c_uint32)))
            elif sh == "plane":
                check(L, L.rt_scene_add_plane(h, d3(o["point"]), d3(o["normal"]), int(o["material"])))
            else:
                raise ValueError("unknown shape %r" % sh)
        for l in desc["lights"]:
            k = l["model"]
            if k == "point":
                check(L, L.rt_scene_add_light_point(h, d3(l["location"]), d3(l["color"])))
            elif k == "directional":
                check(L, L.rt_scene_add_light_directional(h, d3(l["direction"]), d3(l["color"])))
            elif k == "area":
                check(L, L.rt_scene_add_light_area(h, d3(l["origin"]), d3(l["side1"]), d3(l["side2"]), d3(l["color"])))
            else:
                raise ValueError("unknown light model %r" % k)
        cam = desc.get("camera")
        if cam is not None:
            if cam["ctor"] == "new":
                check(L, L.rt_scene_set_camera_spc_new(h, d3(cam["position"]), d3(cam["look"]), d3(cam["up"]), float(cam["im_dist"])))
            else:
                check(L, L.rt_scene_set_camera_spc_look_at(h, d3(cam["focus"]), d3(cam["look"]), d3(cam["up"]), float(cam["pov"]), float(cam["h"])))
            if cam.get("dof"):
                d = cam["dof"]
                check(L, L.rt_scene_set_camera_dof(h, float(d["focus"]), float(d["aperture"]), int(d["samples"])))
        bg = desc.get("background")
        if bg is not None:
            if bg["type"] == "solid":
                check(L, L.rt_scene_set_background_solid(h, d3(bg["color"])))
            else:
                for i, f in enumerate(bg["faces"]):
                    f = np.ascontiguousarray(f, dtype=np.uint8)
                    assert f.ndim == 3 and f.shape[2] == 3
                    check(L, L.rt_scene_set_skybox_face(h, i, f.shape[1], f.shape[0], ptr(f, C.c_uint8)))
        o = desc.get("options")
        if o is not None:
            check(L, L.rt_scene_set_options(h, int(o["width"]), int(o["height"]), int(o["antialias"])))

    def update_spheres(self, first_object, center_xyz_radius):
        """Moves / resizes consecutive sphere objects (an animation step): rt_scene_update_spheres."""
        a = np.ascontiguousarray(center_xyz_radius, dtype=np.float64).reshape(-1, 4)
        check(self._lib, self._lib.rt_scene_update_spheres(self._h, int(first_object), a.shape[0], ptr(a, C.c_double)))

    # -- introspection ---------------------------------------------------------------------------------
    def options(self):
        w, h, a = C.c_uint32(), C.c_uint32(), C.c_uint32()
        check(self._lib, self._lib.rt_scene_get_options(self._h, C.byref(w), C.byref(h), C.byref(a)))
        return w.value, h.value, a.value

    def info(self):
        i = SceneInfo()
        check(self._lib, self._lib.rt_scene_get_info(self._h, C.byref(i)))
        return {k: getattr(i, k) for k, _ in i._fields_}

    def camera(self):
        pos, m = np.zeros(3), np.zeros(9)
        check(self._lib, self._lib.rt_scene_get_camera(self._h, ptr(pos, C.c_double), ptr(m, C.c_double)))
        return pos, m.reshape(3, 3)

    def warnings(self):
        return (self._lib.rt_scene_warnings(self._h) or b"").decode("utf-8", "replace")

    def dump(self):
        n = self._lib.rt_scene_dump(self._h, None, 0)
        buf = C.create_string_buffer(n + 1)
        self._lib.rt_scene_dump(self._h, buf, n + 1)
        return buf.value.decode("utf-8", "replace")


def bmp_bytewidth(lib, width):
    return int(lib.rt_bmp_bytewidth(int(width)))


def bmp_header(lib, width, height):
    buf = np.zeros(122, dtype=np.uint8)
    lib.rt_bmp_header(int(width), int(height), ptr(buf, C.c_uint8))
    return buf.tobytes()
