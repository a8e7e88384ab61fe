"""What each wavefront level costs (development aid): the frame rendered with max_depth = 0, 1, ... -- the difference between two
rows is one level's rays and device time per kernel class.  usage: gpu_levels.py [10k|1m] [world]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
which = sys.argv[1] if len(sys.argv) > 1 else "10k"
world = int(sys.argv[2]) if len(sys.argv) > 2 else 1
d, seed = {"10k": (scenes.spheres_10k, 3), "1m": (lambda: scenes.spheres_1m(3840, 2160), 4)}[which]
sc = rt.Scene(d())
prev = None
print("lib", os.path.basename(rt.LIB_PATH), which, "world", world)
for md in range(0, 9):
    ctx = rt.Context(sc, accel=rt.ACCEL_BVH, max_depth=md, seed=seed, rank=0, world=world)
    ctx.render()
    plain = min(ctx.render()["render_ms"] for _ in range(3))
    ctx.set_profiling(True)
    st = ctx.render()
    ctx.close()
    cur = dict(c=st["rays_primary"] + st["rays_secondary"], s=st["rays_shadow"], ms=plain, **{k: st["class_ms"][k] for k in ("closest", "shade", "shadow")})
    if prev:
        print("depth<=%d: +%8d closest rays +%8d shadow | +%.3f ms frame | closest +%.3f shade +%.3f shadow +%.3f" % (
            md, cur["c"] - prev["c"], cur["s"] - prev["s"], cur["ms"] - prev["ms"], cur["closest"] - prev["closest"], cur["shade"] - prev["shade"], cur["shadow"] - prev["shadow"]))
    else:
        print("depth<=0:  %8d closest rays  %8d shadow |  %.3f ms frame | closest  %.3f shade  %.3f shadow  %.3f" % (cur["c"], cur["s"], cur["ms"], cur["closest"], cur["shade"], cur["shadow"]))
    prev = cur
