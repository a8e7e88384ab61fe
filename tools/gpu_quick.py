"""Quick device timings of the named workloads (development aid; bench.py is the measured contract)."""
import importlib
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")


def run(name, desc, reps=3, **kw):
    t0 = time.time()
    sc = rt.Scene(desc)
    t1 = time.time()
    ctx = rt.Context(sc, **kw)
    t2 = time.time()
    best = None
    for _ in range(reps):
        st = ctx.render()
        if best is None or st["render_ms"] < best["render_ms"]:
            best = st
    print("%-28s scene %.2fs ctx %.2fs | %.3f ms  %.1f Mrays/s  rays %d (p %d s %d sh %d) launches %d bvh %.2f ms nodes %d" % (
        name, t1 - t0, t2 - t1, best["render_ms"], best["rays_total"] / best["render_ms"] / 1e3, best["rays_total"], best["rays_primary"],
        best["rays_secondary"], best["rays_shadow"], best["kernel_launches"], best["bvh_build_ms"], best["n_bvh_nodes"]), flush=True)
    ctx.close()
    return best


if __name__ == "__main__":
    which = sys.argv[1:] or ["10k"]
    if "10k" in which:
        d = scenes.spheres_10k()
        run("10k/4K/d8 bvh strict", d, max_depth=8, seed=3, accel=rt.ACCEL_BVH)
        run("10k/4K/d8 bvh fast", d, max_depth=8, seed=3, accel=rt.ACCEL_BVH, fp_mode=rt.FP_FAST)
        run("10k/4K/d8 linear strict", d, reps=1, max_depth=8, seed=3, accel=rt.ACCEL_LINEAR)
        run("10k/4K/d8 linear fast", d, reps=1, max_depth=8, seed=3, accel=rt.ACCEL_LINEAR, fp_mode=rt.FP_FAST)
    if "1m" in which:
        d = scenes.spheres_1m(3840, 2160)
        run("1M/4K/d8 bvh strict", d, max_depth=8, seed=4, accel=rt.ACCEL_BVH)
        run("1M/4K/d8 bvh fast", d, max_depth=8, seed=4, accel=rt.ACCEL_BVH, fp_mode=rt.FP_FAST)
    if "cornell" in which:
        sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
        REF_TEXT_SCENE = scenes.TEST_SCENE_TEXT
        sc = rt.serialize.deserialize(REF_TEXT_SCENE)
        for aa in (1, 16):
            ctx = rt.Context(sc, width=3840, height=2160, antialias=aa, seed=1)
            st = ctx.render(); st = ctx.render()
            print("cornell 4K aa%d: %.3f ms %.1f Mrays/s rays %d launches %d" % (aa, st["render_ms"], st["rays_total"] / st["render_ms"] / 1e3, st["rays_total"], st["kernel_launches"]), flush=True)
            ctx.close()
