"""Where the host-to-host time of rt_render_multi goes (development aid): create / render / destroy of an rt_multi, per gather mode."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
n = int(sys.argv[1]) if len(sys.argv) > 1 else rt.device_count()
which = sys.argv[2] if len(sys.argv) > 2 else "10k"
d, kw = {"10k": (scenes.spheres_10k, dict(max_depth=8, seed=3)), "1m8k": (lambda: scenes.spheres_1m(7680, 4320), dict(max_depth=8, seed=4))}[which]
desc = d()
sc = rt.Scene(desc)
w, h, _ = sc.options()
bw = rt.bmp.bytewidth(w)
host = torch.empty(h * bw, dtype=torch.uint8).pin_memory()
pageable = np.zeros(h * bw, dtype=np.uint8)
def T(f, reps=5):
    f(); ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); f(); ts.append((time.perf_counter() - t0) * 1e3)
    return min(ts), sum(ts) / len(ts)
print("devices", n, which)
print("rt_render (1 GPU) -> pinned   : %.2f / %.2f ms" % T(lambda: rt.raytrace.render_into(sc, host.data_ptr(), accel=rt.ACCEL_BVH, **kw)))
for name, g in (("direct", rt.GATHER_DIRECT), ("peer", rt.GATHER_PEER), ("nccl", rt.GATHER_NCCL)):
    print("rt_render_multi %-6s -> pinned : %.2f / %.2f ms" % ((name,) + T(lambda: rt.raytrace.render_into(sc, host.data_ptr(), devices=n, gather=g, accel=rt.ACCEL_BVH, **kw))))
    t0 = time.perf_counter(); mc = rt.MultiContext(sc, n, gather=g, accel=rt.ACCEL_BVH, **kw); t1 = time.perf_counter()
    r = T(lambda: mc.render_to(host.data_ptr()))
    st = mc.render_to(host.data_ptr())
    r2 = T(lambda: mc.render_to(pageable.ctypes.data))
    r3 = T(lambda: mc.render_to(None)) if g != rt.GATHER_DIRECT else (0, 0)
    t2 = time.perf_counter(); mc.close(); t3 = time.perf_counter()
    print("   create %.2f ms  render->pinned %.2f / %.2f  ->pageable %.2f / %.2f  ->device only %.2f / %.2f  (device render_ms max %.2f)  destroy %.2f ms" % (
        (t1 - t0) * 1e3, r[0], r[1], r2[0], r2[1], r3[0], r3[1], st["render_ms"], (t3 - t2) * 1e3))
