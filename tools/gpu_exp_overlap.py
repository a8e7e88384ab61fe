import importlib, os, sys, time
sys.path.insert(0, "/root/repo")
import torch
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
sc = rt.Scene(scenes.spheres_10k())
kw = dict(max_depth=8, seed=3, accel=rt.ACCEL_BVH)
w, h, _ = sc.options(); bw = rt.bmp.bytewidth(w)
host = torch.empty(h * bw, dtype=torch.uint8).pin_memory()
for devs in ([0], [0, 0], [0, 0, 0], [0, 0, 0, 0]):
    mc = rt.MultiContext(sc, devs, gather=rt.GATHER_PEER, **kw)
    mc.render_to(None); mc.render_to(None)
    ts = []
    for _ in range(5):
        t0 = time.perf_counter(); st = mc.render_to(None); ts.append((time.perf_counter() - t0) * 1e3)
    print(len(devs), "shards on one GPU: wall %.2f ms (min %.2f), max shard render_ms %.2f" % (sum(ts) / 5, min(ts), st["render_ms"]), flush=True)
    mc.close()
