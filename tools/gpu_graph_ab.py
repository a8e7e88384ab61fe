"""Device ms per frame with the frame's CUDA graph replayed (RT_TUNE_FRAME_GRAPH) and with plain launches (default), whole frame and rank 0's shard
of 8; and the host time of rt_context_render_begin (development aid)."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
W = {"10k": (scenes.spheres_10k, dict(max_depth=8, seed=3)), "1m": (lambda: scenes.spheres_1m(3840, 2160), dict(max_depth=8, seed=4))}
for which in (sys.argv[1:] or ["10k", "1m"]):
    d, kw = W[which]
    sc = rt.Scene(d())
    for world in (1, 8):
        res, host = {}, {}
        for tuning in (rt.TUNE_FRAME_GRAPH, 0, rt.TUNE_FRAME_GRAPH, 0):
            ctx = rt.Context(sc, accel=rt.ACCEL_BVH, tuning=tuning, rank=0, world=world, **kw)
            for _ in range(4):
                ctx.render()
            ms, hb = [], []
            for _ in range(9):
                t0 = time.perf_counter(); ctx.render_begin(); hb.append((time.perf_counter() - t0) * 1e3)
                ms.append(ctx.render_end()["render_ms"])
            res.setdefault(tuning, []).append(sorted(ms)[4]); host.setdefault(tuning, []).append(sorted(hb)[4])
            del ctx
        print("%-4s world %d  graph %s ms (begin %s ms on the host)   plain %s ms (begin %s ms)" % (which, world, ["%.3f" % x for x in res[rt.TUNE_FRAME_GRAPH]], ["%.3f" % x for x in host[rt.TUNE_FRAME_GRAPH]],
              ["%.3f" % x for x in res[0]], ["%.3f" % x for x in host[0]]), flush=True)
