"""Per-class device time and BVH work per ray for a workload (development aid)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
which = sys.argv[1] if len(sys.argv) > 1 else "10k"
d, kw = {"10k": (scenes.spheres_10k, dict(max_depth=8, seed=3)), "1m": (lambda: scenes.spheres_1m(3840, 2160), dict(max_depth=8, seed=4)),
         "100k": (lambda: scenes.skybox_100k(3840, 2160, antialias=1, face_size=256), dict(max_depth=4, seed=5))}[which]
tuning = int(sys.argv[2]) if len(sys.argv) > 2 else 0   # RT_TUNE_* bits (8 = binary shadow walk, 16 = host-driven levels)
ctx = rt.Context(rt.Scene(d()), accel=rt.ACCEL_BVH, tuning=tuning, **kw)
ctx.set_profiling(True)
ctx.render()
st = ctx.render()
rc, rs = st["rays_primary"] + st["rays_secondary"], st["rays_shadow"]
print(which, "tuning", tuning, "host_levels", st["host_levels"], "node bytes", st["bvh_node_bytes"], "render %.2f ms" % st["render_ms"], {k: round(v, 2) for k, v in st["class_ms"].items()}, "launches", st["class_launches"])
print("closest: %d rays, %.1f nodes/ray, %.2f prims/ray, %.0f Mrays/s" % (rc, st["bvh_nodes_visited"]["closest"] / rc, st["bvh_prims_tested"]["closest"] / rc, rc / st["class_ms"]["closest"] / 1e3))
print("shadow : %d rays, %.1f nodes/ray, %.2f prims/ray, %.0f Mrays/s" % (rs, st["bvh_nodes_visited"]["shadow"] / rs, st["bvh_prims_tested"]["shadow"] / rs, rs / st["class_ms"]["shadow"] / 1e3))
print("heavy rays:", st["heavy_rays"])
