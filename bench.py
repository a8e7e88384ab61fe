#!/usr/bin/env python3
"""bench.py -- measures the tracing hot path on N B200s of one node (contract: task prompt, section 4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference] [--fp strict|fast]

A step = one frame of the workload: primary-ray generation -> wavefront (closest-hit, shade, shadow kernels) ->
resolve, on a device-resident scene (inputs already in HBM).  With N > 1 the frame is sharded by interleaved row
bands (one process per GPU), every rank traces its bands, the quantised BGR bands are all-gathered over NCCL and
de-interleaved into BMP row order: strong scaling of one frame.

metric  Mrays/s, primary + secondary + shadow: 1 ray = one Scene::intersect call upstream (src/scene.rs:247),
        counted exactly from the wavefront queue lengths.
value   whole-job rays of K frames / device time of K frames (max over ranks).
e2e     the same metric through the user-facing call with HOST buffers: host scene -> upload -> device BVH build ->
        trace -> framebuffer copied back to host memory, every step.

--impl reference times the CPU restatement of the reference (oracle/, kind "port": the upstream crate is Rust and
cannot be built in this image) on the box's host cores on a bounded pixel sample of the same workload.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    # BASELINE.json configs[2]: the metric's 4K single-GPU throughput configuration (default)
    "spheres-10k-4k-d8": dict(gen="spheres_10k", gen_kw=dict(width=3840, height=2160), opts=dict(max_depth=8, seed=3), accel="bvh"),
    # configs[0]: the reference's own CPU-runnable case, test_scene.txt at its own options (800x800, 1024 spp, depth 4)
    "test-scene-800-1024spp": dict(gen="cornell", gen_kw={}, opts=dict(max_depth=4, seed=1), accel="linear"),
    # configs[1]: pixel-match configuration
    "cornell-4k-1spp": dict(gen="cornell", gen_kw={}, opts=dict(width=3840, height=2160, antialias=1, max_depth=4, seed=1), accel="linear"),
    # configs[3] and its 4K variant (north_star target statement)
    "spheres-1m-8k-d8": dict(gen="spheres_1m", gen_kw=dict(width=7680, height=4320), opts=dict(max_depth=8, seed=4), accel="bvh"),
    "spheres-1m-4k-d8": dict(gen="spheres_1m", gen_kw=dict(width=3840, height=2160), opts=dict(max_depth=8, seed=4), accel="bvh"),
    # configs[4]
    "skybox-100k-4k-16spp": dict(gen="skybox_100k", gen_kw=dict(width=3840, height=2160, antialias=16), opts=dict(max_depth=4, seed=5), accel="bvh"),
    # brute-force scan of configs[2] (FP64-pipe-bound oracle pass for the BVH)
    "spheres-10k-4k-d8-linear": dict(gen="spheres_10k", gen_kw=dict(width=3840, height=2160), opts=dict(max_depth=8, seed=3), accel="linear"),
}
DEFAULT_WORKLOAD = "spheres-10k-4k-d8"


def make_desc(name):
    w = WORKLOADS[name]
    if w["gen"] == "cornell":
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from test_gpu_parity import REF_TEXT_SCENE
        return None, REF_TEXT_SCENE
    scenes = importlib.import_module("rust-raytrace_b200.scenes")
    return getattr(scenes, w["gen"])(**w["gen_kw"]), None


def frame_size(name, desc):
    o = WORKLOADS[name]["opts"]
    if desc is not None:
        return o.get("width") or desc["options"]["width"], o.get("height") or desc["options"]["height"], o.get("antialias", desc["options"]["antialias"])
    return o.get("width", 800), o.get("height", 800), o.get("antialias", 1024)   # test_scene.txt's own options block


class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device_index):
        self.idx, self.proc, self.lines = device_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------------------------------
# CPU side: the oracle on a bounded pixel sample
# ---------------------------------------------------------------------------------------------------------------------
def oracle_scene(name):
    from oracle import oracle as orc, scene_text
    desc, text = make_desc(name)
    if desc is None:
        desc = scene_text.deserialize(text)
    return orc.OracleScene(desc), desc


def sample_pixels(width, height, n, seed):
    rng = np.random.default_rng(seed)
    return rng.choice(width * height, size=min(n, width * height), replace=False).astype(np.uint32)


def oracle_sample_rate(osc, name, desc, n_pixels, seed, threads=0):
    w, h, aa = frame_size(name, desc)
    o = WORKLOADS[name]["opts"]
    px = sample_pixels(w, h, n_pixels, seed)
    r = osc.render_pixels(px, width=w, height=h, antialias=aa, max_depth=o["max_depth"], seed=o["seed"], threads=threads)
    st = r["stats"]
    return st["rays"], st["seconds"], st["threads"], len(px)


def cpu_baseline(name, budget_s=12.0):
    """Oracle ('port') on all host cores over a pixel sample sized to ~budget_s of CPU work."""
    osc, desc = oracle_scene(name)
    rays, sec, threads, npx = oracle_sample_rate(osc, name, desc, 256, 1234)
    per_px = max(sec, 1e-4) / npx
    n = int(max(256, min(2_000_000, budget_s / per_px)))
    rays, sec, threads, npx = oracle_sample_rate(osc, name, desc, n, 4321)
    return {"value": rays / sec / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "port",
            "sample": "%d pseudo-random pixels of the %s frame (%d rays, %.1f s), counter RNG, OpenMP over pixels" % (npx, name, rays, sec)}


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (oracle port) on the host cores; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = args.workload
    osc, desc = oracle_scene(name)
    rays, sec, threads, npx = oracle_sample_rate(osc, name, desc, 256, 99)
    per_px = max(sec, 1e-4) / npx
    total_budget = 90.0
    n = int(max(128, min(1_000_000, total_budget / (args.steps + args.warmup) / per_px)))
    for i in range(args.warmup):
        oracle_sample_rate(osc, name, desc, n, 1000 + i)
    tot_rays, tot_sec = 0, 0.0
    for i in range(args.steps):
        rays, sec, threads, npx = oracle_sample_rate(osc, name, desc, n, 2000 + i)
        tot_rays += rays
        tot_sec += sec
    value = tot_rays / tot_sec / 1e6
    w, h, aa = frame_size(name, desc)
    sample = "%d pseudo-random pixels of the frame per step (%d rays per step), counter RNG, OpenMP over pixels" % (n, tot_rays // max(1, args.steps))
    line = {"impl": "reference", "metric": "Mrays/s (primary+secondary+shadow) at 4K", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_sec / max(1, args.steps) * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(name, w, h, aa, desc),
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(name, w, h, aa, desc):
    scene_mod = importlib.import_module("rust-raytrace_b200.scene")
    return {"workload": name, "width": w, "height": h, "antialias": aa, "max_depth": WORKLOADS[name]["opts"]["max_depth"],
            "objects": scene_mod.count_objects(desc), "lights": len(desc["lights"]), "accel": WORKLOADS[name]["accel"],
            "l2_policy": "inputs larger than L2: per-frame ray queues + accumulators exceed 126 MB and are rewritten every step"}


# ---------------------------------------------------------------------------------------------------------------------
# GPU side
# ---------------------------------------------------------------------------------------------------------------------
def scene_h2d_bytes(desc):
    """bytes the library uploads for this scene (flattened records; see DESIGN.md 'Data layout')"""
    scene_mod = importlib.import_module("rust-raytrace_b200.scene")
    n = scene_mod.count_objects(desc)
    b = n * (32 + 4 + 4 + 4) + len(desc["materials"]) * 96 + len(desc["lights"]) * 104 + 511 * 8
    if desc["background"]["type"] == "skybox":
        b += sum(int(f.size) for f in desc["background"]["faces"])
    return int(b)


def roofline_for(stats_sum, steps, peak_gbs, peak_src, accel):
    """Dominant kernel class by measured device time; algorithmic bytes per SURVEY.md 8(d) / DESIGN.md."""
    cls = max(stats_sum["class_ms"], key=lambda k: stats_sum["class_ms"][k])
    ms = stats_sum["class_ms"][cls]
    launches = max(1, stats_sum["class_launches"][cls])
    R_CLOSEST, H_CLOSEST = 48, 12          # origin+direction read, id+t written
    R_SHADOW = 84                          # origin, direction, range, pending colour, pixel read
    R_SHADE_IN, R_CHILD, R_SHADOW_OUT = 108, 96, 84
    rays_c = stats_sum["rays_primary"] + stats_sum["rays_secondary"]
    if cls == "closest":
        queue = rays_c * (R_CLOSEST + H_CLOSEST)
        tree = stats_sum["bvh_nodes"]["closest"] * stats_sum.get("node_bytes", 64) + stats_sum["bvh_prims"]["closest"] * 32
    elif cls == "shadow":
        queue = stats_sum["rays_shadow"] * R_SHADOW
        tree = stats_sum["bvh_nodes"]["shadow"] * stats_sum.get("node_bytes", 64) + stats_sum["bvh_prims"]["shadow"] * 32
    elif cls == "shade":
        queue = rays_c * R_SHADE_IN + stats_sum["rays_secondary"] * R_CHILD + stats_sum["rays_shadow"] * R_SHADOW_OUT
        tree = 0
    else:
        queue, tree = 0, 0
    achieved = queue / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
    return {"bound": "hbm", "kernel": {"closest": "k_closest_bvh" if accel == "bvh" else "k_closest_scan",
                                       "shadow": "k_shadow_bvh" if accel == "bvh" else "k_shadow_scan", "shade": "k_shade",
                                       "primary": "k_primary", "resolve": "k_resolve"}[cls],
            "achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs, "traffic": None,
            "peak_source": peak_src, "algorithmic_bytes_per_launch": queue / launches, "avg_launch_ms": ms / launches,
            "share_of_step": ms / max(1e-9, sum(stats_sum["class_ms"].values())),
            "class_ms_per_step": {k: v / steps for k, v in stats_sum["class_ms"].items()},
            "tree_bytes_per_launch_L2": tree / launches, "bvh_node_bytes": stats_sum.get("node_bytes", 0),
            "tree_gbs_L2": tree / (ms * 1e-3) / 1e9 if ms > 0 else 0.0,
            "note": "achieved = compulsory queue-record bytes of the dominant kernel / its device time; BVH node+leaf fetches "
                    "(tree_*) are served by L1/L2 and reported beside it"}


def add_ncu_traffic(roof, workload, world):
    """roofline.traffic = DRAM bytes (read + write) of ONE launch of the dominant kernel from the committed ncu --set full
    capture of this workload (profiles/ncu_traffic.json, written from profiles/*_ncu_<kernel>.txt); null when there is no
    capture for this workload / kernel / world size.  bench.py never runs under a profiler itself."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            t = json.load(f)
        k = t["kernels"].get(roof["kernel"])
        if world == 1 and t.get("workload") == workload and k:
            roof["traffic"] = k["dram_bytes_read"] + k["dram_bytes_write"]
            roof["traffic_source"] = k["source"]
            # what the same capture says binds the kernel when it is not the HBM roofline
            roof["ncu"] = {m: k[m] for m in ("issue_slots_pct", "lanes_per_instruction", "l1_lsu_wavefronts_pct", "warps_active_pct") if m in k}
    except (OSError, ValueError, KeyError):
        pass
    return roof


def run_ours(args):
    import torch
    import torch.distributed as dist

    rt = importlib.import_module("rust-raytrace_b200")
    name = args.workload
    wl = WORKLOADS[name]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch N>1 with torchrun (one rank per GPU)" % (args.gpus, world))
    if rt.device_count() < 1:
        raise SystemExit("no CUDA device visible: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    desc, text = make_desc(name)
    if desc is None:
        from oracle import scene_text  # only to describe the workload in `config`; the product parses the text itself
        scene = rt.serialize.deserialize(text)
        desc = scene_text.deserialize(text)
    else:
        scene = rt.Scene(desc)
    w, h, aa = frame_size(name, desc)
    accel = rt.ACCEL_BVH if wl["accel"] == "bvh" else rt.ACCEL_LINEAR
    fp_mode = rt.FP_FAST if args.fp == "fast" else rt.FP_STRICT
    opts = dict(wl["opts"], accel=accel, fp_mode=fp_mode, device=local_rank, rank=rank, world=world)
    ctx = rt.Context(scene, **opts)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    ctx.set_profiling(True)

    fb_ptr, fb_bytes = ctx.framebuffer_ptr()
    bytewidth = ctx.bytewidth
    final = torch.empty(h * bytewidth, dtype=torch.uint8, device="cuda")
    gathered = torch.empty(world * fb_bytes, dtype=torch.uint8, device="cuda") if world > 1 else None
    # zero-copy view of the context's device framebuffer for the NCCL gather
    local_fb = None
    if world > 1:
        class _Holder:
            pass
        hold = _Holder()
        hold.__cuda_array_interface__ = {"shape": (fb_bytes,), "typestr": "|u1", "data": (fb_ptr, False), "version": 2}
        local_fb = torch.as_tensor(hold, device="cuda")

    # N > 1: the gather of frame k runs on its own stream beside the trace of frame k+1 (the shard's bytes are first copied
    # out of the context's framebuffer, which the next frame's resolve overwrites); frame k is assembled (untile) right
    # after trace k+1.  drain() assembles the last frame: it is called before the timed region closes.
    gather_stream = torch.cuda.Stream() if world > 1 else None
    staging = torch.empty_like(local_fb) if world > 1 else None
    ev_traced, ev_gathered = torch.cuda.Event(), torch.cuda.Event()
    pending = [False]

    def drain():
        if pending[0]:
            stream.wait_event(ev_gathered)
            ctx.untile(gathered.data_ptr(), final.data_ptr())
            pending[0] = False

    def step():
        st = ctx.render()
        if world > 1:
            drain()
            staging.copy_(local_fb)
            ev_traced.record(stream)
            gather_stream.wait_event(ev_traced)
            with torch.cuda.stream(gather_stream):
                dist.all_gather_into_tensor(gathered, staging)
                ev_gathered.record(gather_stream)
            pending[0] = True
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(3, args.warmup)):
        step()
    drain()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    acc = {"rays_total": 0, "rays_primary": 0, "rays_secondary": 0, "rays_shadow": 0, "kernel_launches": 0, "render_ms": 0.0,
           "class_ms": {}, "class_launches": {}, "bvh_nodes": {"closest": 0, "shadow": 0}, "bvh_prims": {"closest": 0, "shadow": 0}}
    barrier()
    t0 = time.perf_counter()
    ev0.record(stream)
    for _ in range(args.steps):
        st = step()
        for k in ("rays_total", "rays_primary", "rays_secondary", "rays_shadow", "kernel_launches", "render_ms"):
            acc[k] += st[k]
        for k, v in st["class_ms"].items():
            acc["class_ms"][k] = acc["class_ms"].get(k, 0.0) + v
            acc["class_launches"][k] = acc["class_launches"].get(k, 0) + st["class_launches"][k]
        acc["node_bytes"] = st["bvh_node_bytes"]
        for k in ("closest", "shadow"):
            acc["bvh_nodes"][k] += st["bvh_nodes_visited"][k]
            acc["bvh_prims"][k] += st["bvh_prims_tested"][k]
    drain()
    ev1.record(stream)
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    dev_ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        acc["kernel_launches"] += args.steps  # k_untile, once per frame
    return dict(rt=rt, torch=torch, dist=dist, scene=scene, desc=desc, ctx=ctx, acc=acc, dev_ms=dev_ms, wall_ms=wall_ms, clocks=clocks, world=world,
                rank=rank, local_rank=local_rank, opts=opts, w=w, h=h, aa=aa, final=final, gathered=gathered, local_fb=local_fb, step=step,
                barrier=barrier, bytewidth=bytewidth)


def run_e2e(R, steps):
    """End to end through the user-facing objects with HOST buffers: every step builds the device context from the host
    scene (upload + device BVH build + queue allocation), traces, gathers (N > 1) and copies the frame to host memory."""
    rt, torch, dist = R["rt"], R["torch"], R["dist"]
    world, rank = R["world"], R["rank"]
    h, bytewidth = R["h"], R["bytewidth"]
    host = torch.empty(max(h * bytewidth, R["ctx"].framebuffer_ptr()[1]), dtype=torch.uint8).pin_memory()
    rays = 0

    def one():
        ctx = rt.Context(R["scene"], **R["opts"])
        st = ctx.render()
        if world > 1:
            ptr, nbytes = ctx.framebuffer_ptr()

            class _H:
                pass
            hh = _H()
            hh.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
            dist.all_gather_into_tensor(R["gathered"], torch.as_tensor(hh, device="cuda"))
            torch.cuda.synchronize()
            ctx.untile(R["gathered"].data_ptr(), R["final"].data_ptr())
            ctx.synchronize()
            if rank == 0:
                host.copy_(R["final"])
        else:
            ctx.download_rows_to(host.data_ptr())  # pinned host memory; local_rows == padded height at world 1
        torch.cuda.synchronize()
        ctx.close()
        return st["rays_total"]

    one()
    R["barrier"]()
    t0 = time.perf_counter()
    for _ in range(steps):
        rays += one()
    R["barrier"]()
    sec = time.perf_counter() - t0
    t = torch.tensor([sec, float(rays)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        sec, rays = float(tmax[0]), float(tsum[1])
    return rays / sec / 1e6, sec / steps * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--fp", default="strict", choices=["strict", "fast"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
        return
    R = run_ours(args)
    torch, dist, world, rank, acc = R["torch"], R["dist"], R["world"], R["rank"], R["acc"]
    t = torch.tensor([R["dev_ms"], R["wall_ms"]], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(acc["rays_total"]), float(acc["kernel_launches"])], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    dev_ms, wall_ms = float(t[0]), float(t[1])
    rays_all, launches_all = float(tot[0]), int(tot[1])
    e2e_val, e2e_ms = (None, None)
    if not args.no_e2e:
        e2e_val, e2e_ms = run_e2e(R, min(args.steps, 5))
    if rank == 0:
        peak, peak_src = load_peaks()
        name = args.workload
        line = {"metric": "Mrays/s (primary+secondary+shadow) at 4K", "value": rays_all / (dev_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": dev_ms / args.steps,
                "wall_ms_per_step": wall_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64", "fp_mode": args.fp, "data": "synthetic", "config": workload_config(name, R["w"], R["h"], R["aa"], R["desc"]),
                "rays_per_step": rays_all / args.steps, "gpu_launches": launches_all, "clocks": R["clocks"],
                "roofline": add_ncu_traffic(roofline_for(acc, args.steps, peak, peak_src, WORKLOADS[name]["accel"]), name, world),
                "bvh_build_ms": R["ctx"].create_stats["bvh_build_ms"]}
        if world > 1:
            line["config"]["multi_gpu"] = ("image row bands (4 rows) interleaved over ranks, scene replicated; one NCCL all_gather of the "
                                           "BGR8 shards per frame, pipelined one frame deep with the next frame's trace, then untile")
        if e2e_val is not None:
            line["e2e"] = {"value": e2e_val, "unit": "Mrays/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": scene_h2d_bytes(R["desc"]),
                           "d2h_bytes_per_step": R["h"] * R["bytewidth"],
                           "what": "host scene -> rt_context_create (upload, device BVH build, queues) -> render -> gather -> frame in host memory"}
        if args.gpus == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(name)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
