"""Device time per frame with the shadow segments beside the level loop (default) and with RT_TUNE_SERIAL_SHADOW, on a
whole frame and on rank 0's shard of 8 (development aid).  usage: python tools/gpu_overlap_ab.py [10k|1m|1m8k ...]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
W = {"10k": (scenes.spheres_10k, dict(max_depth=8, seed=3)), "1m": (lambda: scenes.spheres_1m(3840, 2160), dict(max_depth=8, seed=4)),
     "1m8k": (lambda: scenes.spheres_1m(7680, 4320), dict(max_depth=8, seed=4)),
     "100k": (lambda: scenes.skybox_100k(3840, 2160, antialias=4, face_size=256), dict(max_depth=4, seed=5))}
for which in (sys.argv[1:] or ["10k", "1m"]):
    d, kw = W[which]
    sc = rt.Scene(d())
    for world in (1, 8):
        res = {}
        for tuning in (0, rt.TUNE_SERIAL_SHADOW, 0, rt.TUNE_SERIAL_SHADOW):
            ctx = rt.Context(sc, accel=rt.ACCEL_BVH, tuning=tuning, rank=0, world=world, **kw)
            for _ in range(3):
                ctx.render()
            ms = sorted(ctx.render()["render_ms"] for _ in range(9))
            res.setdefault(tuning, []).append(ms[len(ms) // 2])
            st = ctx.render()
            del ctx
        print("%-5s world %d  overlapped %s ms   serial %s ms   launches %d" % (which, world, ["%.3f" % x for x in res[0]], ["%.3f" % x for x in res[rt.TUNE_SERIAL_SHADOW]], st["kernel_launches"]), flush=True)
