"""Small renders that touch every kernel, for compute-sanitizer (memcheck / racecheck / initcheck / synccheck):
BVH with grid and f32 nodes, wide and binary shadow walks, the linear scan (TMA-staged tiles), skybox, every material,
device-driven and host-driven frames (queue slicing), heavy-ray hand-off with split scans, sharded frames resolved into
a whole-frame buffer, the LBVH build."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
for name, desc, kw in [
    ("bvh", scenes.random_spheres(3000, 5, 30.0, 160, 96), dict(max_depth=5, seed=1, accel=rt.ACCEL_BVH)),
    ("bvh-f32-binary", scenes.random_spheres(3000, 5, 30.0, 96, 64), dict(max_depth=4, seed=1, accel=rt.ACCEL_BVH, tuning=rt.TUNE_NODES_F32 | rt.TUNE_BINARY_SHADOW)),
    ("bvh-q16-binary-host", scenes.random_spheres(3000, 5, 30.0, 96, 64), dict(max_depth=4, seed=1, accel=rt.ACCEL_BVH, tuning=rt.TUNE_BINARY_SHADOW | rt.TUNE_HOST_LEVELS)),
    ("linear", scenes.random_spheres(1500, 6, 12.0, 96, 64), dict(max_depth=4, seed=2, accel=rt.ACCEL_LINEAR)),
    ("linear-host", scenes.random_spheres(300, 6, 12.0, 64, 48), dict(max_depth=4, seed=2, accel=rt.ACCEL_LINEAR, tuning=rt.TUNE_HOST_LEVELS)),
    ("skybox", scenes.skybox_100k(96, 64, antialias=2, n=2000, face_size=64), dict(max_depth=3, seed=3, accel=rt.ACCEL_BVH)),
    ("showcase", scenes.materials_showcase(96, 64, 2), dict(max_depth=4, seed=4)),
    ("sliced", scenes.materials_showcase(96, 64, 2), dict(max_depth=4, seed=4, max_rays_in_flight=1024)),
]:
    ctx = rt.Context(rt.Scene(desc), **kw)
    st = ctx.render()
    ctx.download()
    print(name, "rays", st["rays_total"], "node bytes", st["bvh_node_bytes"], "host_levels", st["host_levels"], flush=True)
    ctx.close()
# rays from 1e8 units out: runaway walks handed to k_heavy_scan (a few rays -> split scans merged by the last CTA)
desc = scenes.random_spheres(4000, 3, 30.0, 8, 8)
rng = np.random.default_rng(4)
tgt = np.stack([rng.uniform(-30, 30, 64), rng.uniform(0, 6, 64), rng.uniform(-30, 30, 64)], 1)
d = rng.normal(size=(64, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
o = tgt - d * 1e8
o[:, 1] = np.abs(o[:, 1])
ids, t, nrm = rt.raytrace.intersect(rt.Scene(desc), np.concatenate([o, tgt - o], 1), accel=rt.ACCEL_BVH)
print("heavy", int((ids >= 0).sum()), "hits", flush=True)
# shards of one frame resolved into one whole-frame buffer / copied band by band / gathered (one rank)
sc = rt.Scene(scenes.random_spheres(500, 9, 15.0, 35, 37))
for g in (rt.GATHER_DIRECT, rt.GATHER_PEER):
    fr = rt.raytrace.render_multi(sc, [0, 0, 0], gather=g, max_depth=3, seed=6)
    print("multi", g, "rays", fr.stats["rays_total"], flush=True)
fr = rt.raytrace.render_multi(sc, [0], gather=rt.GATHER_NCCL, max_depth=3, seed=6)
print("multi nccl(1) rays", fr.stats["rays_total"], flush=True)
print("peaks", rt.measure_peaks(0) if os.environ.get("SANITIZE_PEAKS") else "skipped")
