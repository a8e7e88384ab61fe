"""Small renders that touch every kernel (BVH both node formats, linear scan, skybox, multi-level queues) for compute-sanitizer."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
for name, desc, kw in [
    ("bvh", scenes.random_spheres(3000, 5, 30.0, 160, 96), dict(max_depth=5, seed=1, accel=rt.ACCEL_BVH)),
    ("linear", scenes.random_spheres(300, 6, 12.0, 96, 64), dict(max_depth=4, seed=2, accel=rt.ACCEL_LINEAR)),
    ("skybox", scenes.skybox_100k(96, 64, antialias=2, n=2000, face_size=64), dict(max_depth=3, seed=3, accel=rt.ACCEL_BVH)),
    ("showcase", scenes.materials_showcase(96, 64, 2), dict(max_depth=4, seed=4)),
]:
    ctx = rt.Context(rt.Scene(desc), **kw)
    st = ctx.render()
    ctx.download()
    print(name, "rays", st["rays_total"], "node bytes", st["bvh_node_bytes"], flush=True)
    ctx.close()
