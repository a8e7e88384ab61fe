"""Wall time of the phases of one end-to-end frame on one GPU (context create / render / download / close), warm and cold
(spheres rewritten before the frame: the create then uploads the scene and rebuilds the BVH).  Development aid."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
which = sys.argv[1] if len(sys.argv) > 1 else "10k"
desc = scenes.spheres_10k() if which == "10k" else scenes.spheres_1m(3840, 2160)
scene = rt.Scene(desc)
first = 0
for ob in desc["objects"]:
    if ob["shape"] == "spheres":
        data = np.ascontiguousarray(ob["data"], dtype=np.float64)
        break
    first += 1
kw = dict(max_depth=8, seed=3 if which == "10k" else 4, accel=rt.ACCEL_BVH)
keep = rt.Context(scene, **kw)
keep.render()
host = torch.empty(keep.framebuffer_ptr()[1], dtype=torch.uint8).pin_memory()
del keep
for cold in (False, True):
    rows = []
    for it in range(7):
        t = [time.perf_counter()]
        if cold:
            scene.update_spheres(first, data)
        t.append(time.perf_counter())
        ctx = rt.Context(scene, **kw); t.append(time.perf_counter())
        st = ctx.render(); t.append(time.perf_counter())
        ctx.download_rows_to(host.data_ptr()); torch.cuda.synchronize(); t.append(time.perf_counter())
        ctx.close(); t.append(time.perf_counter())
        rows.append([1e3 * (b - a) for a, b in zip(t, t[1:])] + [st["render_ms"], ctx.create_stats["bvh_build_ms"] if hasattr(ctx, "create_stats") else 0.0])
    m = np.median(np.array(rows[2:]), axis=0)
    print("%s %s: update %.3f  create %.3f (bvh_build_ms %.3f)  render %.3f (device %.3f)  download %.3f  close %.3f  total %.3f ms" % (
        which, "cold" if cold else "warm", m[0], m[1], m[6], m[2], m[5], m[3], m[4], sum(m[:5])), flush=True)
