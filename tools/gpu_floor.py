"""Launch floor of a frame (development aid): the 10k scene at 64x36 pixels, depth 8 -- every kernel of the level chain runs on
next to nothing, so the device time is what the ~60 launches and their dependencies cost by themselves."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
sc = rt.Scene(scenes.spheres_10k())
for w, h in ((64, 36), (256, 144), (960, 540)):
    for tuning in (0, rt.TUNE_SERIAL_SHADOW, rt.TUNE_HOST_LEVELS):
        ctx = rt.Context(sc, accel=rt.ACCEL_BVH, width=w, height=h, max_depth=8, seed=3, tuning=tuning)
        for _ in range(3):
            ctx.render()
        ms = sorted(ctx.render()["render_ms"] for _ in range(9))
        st = ctx.render()
        print("%4dx%-4d tuning %2d: %.3f ms device, %d launches, %d rays" % (w, h, tuning, ms[4], st["kernel_launches"], st["rays_total"]), flush=True)
        ctx.set_profiling(True)
        st = ctx.render()
        print("      per class (profiling, serial):", {k: round(v, 3) for k, v in st["class_ms"].items()}, st["class_launches"], flush=True)
        del ctx
