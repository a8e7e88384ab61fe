"""Would splitting one GPU's share of the frame into k independent wavefront chains on k streams pay?  k contexts on one
device (ranks 0..k-1 of world k*W, i.e. together the share of rank 0 of world W up to band placement), rendered from k host
threads at once, against one context of world W.  Wall clock over frames (development aid)."""
import importlib, os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
W = {"10k": (scenes.spheres_10k, dict(max_depth=8, seed=3)), "1m": (lambda: scenes.spheres_1m(3840, 2160), dict(max_depth=8, seed=4)),
     "1m8k": (lambda: scenes.spheres_1m(7680, 4320), dict(max_depth=8, seed=4))}
FRAMES = 20
for which in (sys.argv[1:] or ["10k", "1m"]):
    d, kw = W[which]
    sc = rt.Scene(d())
    for world in (1, 8):
        out = []
        for k in (1, 2, 3):
            ctxs = [rt.Context(sc, accel=rt.ACCEL_BVH, rank=r, world=world * k, **kw) for r in range(k)]
            for c in ctxs:
                c.render(); c.render()
            bar = threading.Barrier(k + 1)
            def run(c):
                bar.wait()
                for _ in range(FRAMES):
                    c.render()
                bar.wait()
            th = [threading.Thread(target=run, args=(c,)) for c in ctxs]
            for t in th: t.start()
            bar.wait(); t0 = time.perf_counter(); bar.wait(); t1 = time.perf_counter()
            for t in th: t.join()
            out.append("%d chain(s) %.3f ms" % (k, (t1 - t0) * 1e3 / FRAMES))
            del ctxs
        print("%-5s world %d  " % (which, world) + "   ".join(out), flush=True)
