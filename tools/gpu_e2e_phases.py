"""Wall time of the phases of one end-to-end frame (context create / render / download / close), development aid."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
which = sys.argv[1] if len(sys.argv) > 1 else "10k"
scene = rt.Scene(scenes.spheres_10k() if which == "10k" else scenes.spheres_1m(3840, 2160))
kw = dict(max_depth=8, seed=3 if which == "10k" else 4, accel=rt.ACCEL_BVH)
keep = rt.Context(scene, **kw)   # bench.py keeps its device-resident context alive while the e2e leg runs
keep.render()
host = torch.empty(keep.framebuffer_ptr()[1], dtype=torch.uint8).pin_memory()
for it in range(6):
    t = [time.perf_counter()]
    ctx = rt.Context(scene, **kw); t.append(time.perf_counter())
    st = ctx.render(); t.append(time.perf_counter())
    ctx.download_rows_to(host.data_ptr()); torch.cuda.synchronize(); t.append(time.perf_counter())
    ctx.close(); t.append(time.perf_counter())
    d = [1e3 * (b - a) for a, b in zip(t, t[1:])]
    print("create %.2f render %.2f (device %.2f, build %.2f) download %.2f close %.2f total %.2f ms" % (d[0], d[1], st["render_ms"], st["bvh_build_ms"], d[2], d[3], sum(d)), flush=True)
