"""Device ms per frame of rank r's shard of 8 for band heights 4 / 8 / 16 / 32 rows (development aid): taller bands keep more
of a shard's rays coherent, shorter ones balance the ranks better.  Prints the slowest rank (what a frame costs)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
W = {"10k": (scenes.spheres_10k, dict(max_depth=8, seed=3)), "1m8k": (lambda: scenes.spheres_1m(7680, 4320), dict(max_depth=8, seed=4))}
for which in (sys.argv[1:] or ["10k"]):
    d, kw = W[which]
    sc = rt.Scene(d())
    for band in (4, 8, 16, 32, 64):
        per_rank = []
        for rank in range(8):
            ctx = rt.Context(sc, accel=rt.ACCEL_BVH, rank=rank, world=8, band_rows=band, **kw)
            for _ in range(3):
                ctx.render()
            ms = sorted(ctx.render()["render_ms"] for _ in range(7))
            per_rank.append(ms[len(ms) // 2])
            del ctx
        print("%-5s band %2d rows: slowest rank %.3f ms, fastest %.3f, mean %.3f" % (which, band, max(per_rank), min(per_rank), sum(per_rank) / 8), flush=True)
