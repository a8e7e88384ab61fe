"""Shadow-kernel throughput with one light at a time vs all three (is mixing lights in a warp costly?) -- development aid."""
import importlib, os, sys, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
base = scenes.spheres_10k()
for pick in ([0, 1, 2], [0], [1], [2]):
    d = dict(base); d["lights"] = [base["lights"][i] for i in pick]
    ctx = rt.Context(rt.Scene(d), accel=rt.ACCEL_BVH, max_depth=8, seed=3)
    ctx.set_profiling(True); ctx.render(); st = ctx.render()
    rs = st["rays_shadow"]
    print("lights", pick, "shadow rays %d, %.2f ms, %.1f nodes/ray, %.0f Mrays/s, %.2f ps per ray-node" % (rs, st["class_ms"]["shadow"], st["bvh_nodes_visited"]["shadow"] / rs,
          rs / st["class_ms"]["shadow"] / 1e3, 1e9 * st["class_ms"]["shadow"] / st["bvh_nodes_visited"]["shadow"]), flush=True)
    ctx.close()
