"""Context creation on the 1M-sphere scene, phase by phase: wall time of rt_context_create (first = flatten + page-lock,
later = cached), the device BVH build inside it, and -- under `ncu --metrics gpu__time_duration.sum` -- the build kernels."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
which = sys.argv[1] if len(sys.argv) > 1 else "1m"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 6
t0 = time.perf_counter()
desc = scenes.spheres_10k() if which == "10k" else scenes.spheres_1m(3840, 2160)
t1 = time.perf_counter()
scene = rt.Scene(desc)
t2 = time.perf_counter()
print("describe %.1f ms, rt_scene build %.1f ms" % (1e3 * (t1 - t0), 1e3 * (t2 - t1)), flush=True)
kw = dict(max_depth=8, seed=3 if which == "10k" else 4, accel=rt.ACCEL_BVH)
for it in range(n):
    a = time.perf_counter()
    ctx = rt.Context(scene, **kw)
    b = time.perf_counter()
    st = ctx.create_stats
    ctx.close()
    c = time.perf_counter()
    print("create %.2f ms (bvh build %.2f ms, %d nodes of %d B)  close %.2f ms" % (1e3 * (b - a), st["bvh_build_ms"], st["n_bvh_nodes"], st["bvh_node_bytes"],
                                                                                 1e3 * (c - b)), flush=True)
