"""Per-class device time of ONE rank's shard of a frame (world = argv[2]) -- where strong scaling goes (development aid)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
which = sys.argv[1] if len(sys.argv) > 1 else "10k"
world = int(sys.argv[2]) if len(sys.argv) > 2 else 8
d, kw = {"10k": (scenes.spheres_10k, dict(max_depth=8, seed=3)), "1m8k": (lambda: scenes.spheres_1m(7680, 4320), dict(max_depth=8, seed=4))}[which]
ctx = rt.Context(rt.Scene(d()), accel=rt.ACCEL_BVH, rank=0, world=world, **kw)
ctx.render()
plain = min(ctx.render()["render_ms"] for _ in range(3))
ctx.set_profiling(True)
st = ctx.render()
print(which, "world", world, "render %.3f ms (%.3f with per-launch events)" % (plain, st["render_ms"]), {k: round(v, 3) for k, v in st["class_ms"].items()},
      "sum %.3f" % sum(st["class_ms"].values()), "launches", st["kernel_launches"], "rays", st["rays_total"])
