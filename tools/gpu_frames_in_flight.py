"""Would keeping k frames in flight on one GPU hide the latency-bound deep levels of a shard?  k contexts (shards 0..k-1 of 8: equal
work each) render FRAMES frames each from k host threads at once; frames per second against one context (development aid)."""
import importlib, os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
sc = rt.Scene(scenes.spheres_10k())
FRAMES = 30
for world in (8, 1):
    for k in (1, 2, 3, 4):
        ctxs = [rt.Context(sc, accel=rt.ACCEL_BVH, rank=r % world, world=world, max_depth=8, seed=3) for r in range(k)]
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
        print("world %d: %d frame(s) in flight: %.3f ms per frame" % (world, k, (t1 - t0) * 1e3 / (FRAMES * k)), flush=True)
        del ctxs
