#!/usr/bin/env python3
"""bench.py -- measures the tracing hot path on N B200s of one node (contract: task prompt, section 4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference] [--fp strict|fast]
                    [--gather peer|nccl] [--frames-in-flight F] [--tuning BITS] [--no-also] [--no-e2e] [--no-cpu-baseline]

A step = one frame of the workload: primary-ray generation -> wavefront (closest-hit, shade, shadow kernels) -> resolve,
on a device-resident scene (inputs already in HBM).  With N > 1 the frame is sharded by interleaved row bands, one process
per GPU (torchrun); every rank traces its bands and its resolve kernel stores the quantised BGR rows straight into rank 0's
whole-frame buffer over NVLink (CUDA IPC peer memory: the gather is fused into the kernel that produces the bytes).
`--gather nccl` assembles the frame with an NCCL all-gather + de-interleave instead; at N > 1 the default line reports that
variant beside the headline (`also.gather_nccl`).  Strong scaling of one frame.

metric  Mrays/s, primary + secondary + shadow: 1 ray = one Scene::intersect call upstream (src/scene.rs:247), counted
        exactly from the wavefront queue lengths.
value   whole-job rays of K frames / device time of K frames (CUDA events on the launching stream, max over ranks); no
        per-launch profiling events inside the timed region (kernel-class times come from a separate pass).  The frames
        are independent (upstream's frame loop run for an animation), so --frames-in-flight F (default 2 for the BVH workloads) of them are queued
        at once per GPU, one rt_context each (rt_context_render_begin / _end): on a shard the deep wavefront levels are bound
        by the latency of their longest walks and the next frame fills those gaps.  `frame_ms_alone` is one frame by itself.
e2e     the same metric through the user-facing call with HOST buffers, every step: the caller moves its spheres
        (rt_scene_update_spheres from host memory) -> rt_render / rt_render_multi (flatten, upload, device BVH build on
        every GPU, trace) -> frame in pinned host memory.  e2e_warm: the same call on an unchanged scene (the device copy
        kept with the rt_scene is reused: nothing but the frame crosses PCIe).
also    BASELINE config 4 (1M spheres, 7680x4320, the north_star scaling configuration) measured in the same run.

--impl reference times the CPU restatement of the reference (oracle/, kind "port": the upstream crate is Rust and cannot
be built in this image) on all host cores on a bounded pixel sample of the same workload.
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
ALSO_WORKLOAD = "spheres-1m-8k-d8"   # north_star's scaling configuration, carried by the default line
METRIC = "Mrays/s (primary+secondary+shadow) at 4K"


_JSON_OUT = None


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries print there too (NCCL's version banner at N > 1): keep a private
    handle on the real stdout for that line and point file descriptor 1 at stderr for everything else."""
    global _JSON_OUT
    if _JSON_OUT is None:
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def host_threads():
    """Host cores this process may use.  Never omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def pkg():
    return importlib.import_module("rust-raytrace_b200")


def make_desc(name):
    """(description dict | None, scene text | None) of a workload; the generators live in the product package"""
    w = WORKLOADS[name]
    scenes = importlib.import_module("rust-raytrace_b200.scenes")
    if w["gen"] == "cornell":
        return None, scenes.TEST_SCENE_TEXT
    return getattr(scenes, w["gen"])(**w["gen_kw"]), None


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
# CPU side: the oracle on a bounded pixel sample (test infrastructure; only these two legs may touch oracle/)
# ---------------------------------------------------------------------------------------------------------------------
def oracle_scene(name):
    from oracle import oracle as orc, scene_text
    desc, text = make_desc(name)
    if desc is None:
        desc = scene_text.deserialize(text)
    return orc.OracleScene(desc), desc


def oracle_frame_size(name, desc):
    o = WORKLOADS[name]["opts"]
    return (o.get("width") or desc["options"]["width"], o.get("height") or desc["options"]["height"], o.get("antialias", desc["options"]["antialias"]))


def sample_pixels(width, height, n, seed):
    rng = np.random.default_rng(seed)
    return rng.choice(width * height, size=min(n, width * height), replace=False).astype(np.uint32)


def oracle_sample_rate(osc, name, desc, n_pixels, seed, threads):
    w, h, aa = oracle_frame_size(name, desc)
    o = WORKLOADS[name]["opts"]
    px = sample_pixels(w, h, n_pixels, seed)
    r = osc.render_pixels(px, width=w, height=h, antialias=aa, max_depth=o["max_depth"], seed=o["seed"], threads=threads)
    st = r["stats"]
    return st["rays"], st["seconds"], st["threads"], len(px)


def cpu_baseline(name, budget_s=12.0):
    """Oracle ('port') on all host cores over a pixel sample sized to ~budget_s of CPU work."""
    threads = host_threads()
    osc, desc = oracle_scene(name)
    rays, sec, used, npx = oracle_sample_rate(osc, name, desc, 64 * threads, 1234, threads)
    per_px = max(sec, 1e-4) / npx
    n = int(max(64 * threads, min(2_000_000, budget_s / per_px)))
    rays, sec, used, npx = oracle_sample_rate(osc, name, desc, n, 4321, threads)
    return {"value": rays / sec / 1e6, "unit": "Mrays/s", "cores": used, "kind": "port",
            "sample": "%d pseudo-random pixels of the %s frame (%d rays, %.1f s), counter RNG, OpenMP over pixels, linear scan as upstream" % (npx, name, rays, sec)}


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (oracle port) on all host cores; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    name = args.workload
    osc, desc = oracle_scene(name)
    rays, sec, used, npx = oracle_sample_rate(osc, name, desc, 64 * threads, 99, threads)
    per_px = max(sec, 1e-4) / npx
    total_budget = 90.0
    n = int(max(32 * threads, min(1_000_000, total_budget / (args.steps + args.warmup) / per_px)))
    for i in range(args.warmup):
        oracle_sample_rate(osc, name, desc, n, 1000 + i, threads)
    tot_rays, tot_sec = 0, 0.0
    for i in range(args.steps):
        rays, sec, used, npx = oracle_sample_rate(osc, name, desc, n, 2000 + i, threads)
        tot_rays += rays
        tot_sec += sec
    value = tot_rays / tot_sec / 1e6
    w, h, aa = oracle_frame_size(name, desc)
    scene_mod = importlib.import_module("rust-raytrace_b200.scene")
    sample = "%d pseudo-random pixels of the frame per step (%d rays per step), counter RNG, OpenMP over pixels, linear scan as upstream" % (
        n, tot_rays // max(1, args.steps))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_sec / max(1, args.steps) * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(name, w, h, aa, scene_mod.count_objects(desc), len(desc["lights"])),
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": used, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def workload_config(name, w, h, aa, n_objects, n_lights):
    return {"workload": name, "width": w, "height": h, "antialias": aa, "max_depth": WORKLOADS[name]["opts"]["max_depth"],
            "objects": int(n_objects), "lights": int(n_lights), "accel": WORKLOADS[name]["accel"],
            "l2_policy": "inputs larger than L2: per-frame ray queues + accumulators exceed 126 MB and are rewritten every step"}


# ---------------------------------------------------------------------------------------------------------------------
# GPU side
# ---------------------------------------------------------------------------------------------------------------------
def scene_h2d_bytes(info, desc):
    """bytes the library uploads per device for this scene (flattened records; see DESIGN.md 'Data layout')"""
    b = info["n_objects"] * (4 + 4) + info["n_spheres"] * (32 + 4) + info["n_planes"] * 56 + info["n_materials"] * 96 + info["n_lights"] * 104 + 511 * 8
    if desc is not None and desc["background"]["type"] == "skybox":
        b += sum(int(f.size) for f in desc["background"]["faces"])
    return int(b)


def roofline_for(prof, peak_gbs, peak_src, accel, hw):
    """Dominant kernel class by measured device time (profiling pass); algorithmic bytes / flops per SURVEY.md 8(d) and
    DESIGN.md.  BVH and shading kernels: compulsory queue-record bytes against the measured HBM copy peak, the L1/L2-served
    tree bytes beside it against the measured L2 read bandwidth.  Linear-scan kernels: 25 flop per exact sphere test
    (src/shapes.rs:60-65; device counter) + 14 per plane test against the measured FP64 FMA peak."""
    steps = prof["steps"]
    cls = max(prof["class_ms"], key=lambda k: prof["class_ms"][k])
    ms = prof["class_ms"][cls]
    launches = max(1, prof["class_launches"][cls])
    R_CLOSEST, H_CLOSEST = 48, 12          # origin+direction read, id+t written
    R_SHADOW = 84                          # origin, direction, range, pending colour, pixel read
    R_SHADE_IN, R_CHILD, R_SHADOW_OUT = 108, 96, 84
    rays_c = prof["rays_primary"] + prof["rays_secondary"]
    tree = 0
    if cls == "closest":
        queue = rays_c * (R_CLOSEST + H_CLOSEST)
        tree = prof["bvh_nodes"]["closest"] * prof["node_bytes"] + prof["bvh_prims"]["closest"] * 32
    elif cls == "shadow":
        queue = prof["rays_shadow"] * R_SHADOW
        tree = prof["bvh_nodes"]["shadow"] * prof["shadow_node_bytes"] + prof["bvh_prims"]["shadow"] * 32
    elif cls == "shade":
        queue = rays_c * R_SHADE_IN + prof["rays_secondary"] * R_CHILD + prof["rays_shadow"] * R_SHADOW_OUT
    else:
        queue = 0
    kernel = {"closest": "k_closest_bvh" if accel == "bvh" else "k_closest_scan", "shadow": "k_shadow_bvh" if accel == "bvh" else "k_shadow_scan",
              "shade": "k_shade", "primary": "k_primary", "resolve": "k_resolve"}[cls]
    sec = ms * 1e-3
    out = {"kernel": kernel, "peak_source": peak_src, "avg_launch_ms": ms / launches, "launches_per_step": launches / steps,
           "share_of_step": ms / max(1e-9, sum(prof["class_ms"].values())),
           "class_ms_per_step": {k: v / steps for k, v in prof["class_ms"].items()},
           "hbm_peak_gbs": peak_gbs, "fp64_peak_tflops": hw["fp64_tflops"], "fp32_peak_tflops": hw["fp32_tflops"], "l2_peak_gbs": hw["l2_read_gbs"],
           "traffic": None}
    if accel == "linear" and cls in ("closest", "shadow"):
        n_planes = prof["n_planes"]
        rays = rays_c if cls == "closest" else prof["rays_shadow"]
        flops = 25.0 * prof["scan_tests"][cls] + 14.0 * n_planes * rays
        achieved = flops / sec / 1e12 if sec > 0 else 0.0
        out.update({"bound": "fp64", "achieved": achieved, "peak": hw["fp64_tflops"], "unit": "TFLOP/s", "frac": achieved / hw["fp64_tflops"],
                    "frac_of_fp32_peak": achieved / hw["fp32_tflops"], "algorithmic_flops_per_launch": flops / launches,
                    "sphere_tests_per_ray": prof["scan_tests"][cls] / max(1, rays),
                    "note": "achieved = (25 flop x exact sphere tests counted on the device + 14 flop x plane tests) / device time of the scan "
                            "kernel, against the FP64 FMA throughput measured in this run (rt_measure_peaks); tensor cores unused by design"})
    else:
        achieved = queue / sec / 1e9 if sec > 0 else 0.0
        tree_gbs = tree / sec / 1e9 if sec > 0 else 0.0
        out.update({"bound": "hbm", "achieved": achieved, "peak": peak_gbs, "unit": "GB/s", "frac": achieved / peak_gbs,
                    "algorithmic_bytes_per_launch": queue / launches, "tree_bytes_per_launch_L2": tree / launches,
                    "bvh_node_bytes": prof["shadow_node_bytes"] if cls == "shadow" else prof["node_bytes"], "tree_gbs_L2": tree_gbs,
                    "tree_frac_of_l2_peak": tree_gbs / hw["l2_read_gbs"] if hw["l2_read_gbs"] else None,
                    "note": "achieved = compulsory queue-record bytes of the dominant kernel / its device time against the measured HBM copy "
                            "peak; BVH node + leaf fetches (tree_*) are served by L1/L2 and quoted against the L2 read bandwidth measured in "
                            "this run (rt_measure_peaks).  The traversal is bound by instruction issue, not by either (profiles/)"})
    return out


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
            roof["ncu"] = {m: k[m] for m in ("issue_slots_pct", "lanes_per_instruction", "l1_lsu_wavefronts_pct", "warps_active_pct") if m in k}
    except (OSError, ValueError, KeyError):
        pass
    return roof


class Dist:
    """torch.distributed plumbing: one process per GPU, NCCL; trivial at world 1"""

    def __init__(self, args):
        import torch
        self.torch = torch
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus:
            raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch N>1 with torchrun (one rank per GPU)" % (args.gpus, self.world))
        torch.cuda.set_device(self.local_rank)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
            self.dist = dist
            self.host_group = dist.new_group(backend="gloo")   # barriers that keep the waiting ranks' GPUs idle

    def barrier(self):
        if self.dist:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def host_barrier(self):
        """Host-side barrier (gloo): an NCCL barrier parks a spinning kernel on every waiting rank's GPU, which would time-slice
        with the work rank 0 runs on those GPUs during the end-to-end leg."""
        self.torch.cuda.synchronize()
        if self.dist:
            self.dist.barrier(group=self.host_group)

    def reduce(self, values, op):
        t = self.torch.tensor([float(v) for v in values], dtype=self.torch.float64, device="cuda")
        if self.dist:
            self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op))
        return [float(x) for x in t]

    def broadcast_bytes(self, arr, n):
        t = self.torch.zeros(n, dtype=self.torch.uint8, device="cuda")
        if self.rank == 0:
            t.copy_(self.torch.from_numpy(np.ascontiguousarray(arr, dtype=np.uint8)))
        if self.dist:
            self.dist.broadcast(t, src=0)
        return t.cpu().numpy()

    def close(self):
        if self.dist:
            self.dist.barrier()
            self.dist.destroy_process_group()


def device_view(torch, ptr, nbytes):
    class _H:
        pass
    h = _H()
    h.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(h, device="cuda")


def measure(name, args, D, steps, warmup, hw, want_e2e=True, want_cpu=True, gather="peer", want_nccl_variant=False):
    """One workload on this process's GPU (shard `rank` of `world`): returns the JSON fields of a bench line."""
    rt, torch = pkg(), D.torch
    wl = WORKLOADS[name]
    world, rank = D.world, D.rank
    desc, text = make_desc(name)
    scene = rt.serialize.deserialize(text) if desc is None else rt.Scene(desc)
    info = scene.info()
    sw, sh, saa = scene.options()
    o = wl["opts"]
    w, h, aa = o.get("width") or sw, o.get("height") or sh, o.get("antialias", saa)
    accel = rt.ACCEL_BVH if wl["accel"] == "bvh" else rt.ACCEL_LINEAR
    fp_mode = rt.FP_FAST if args.fp == "fast" else rt.FP_STRICT
    base_opts = dict(wl["opts"], accel=accel, fp_mode=fp_mode)
    if getattr(args, "tuning", 0):
        base_opts["tuning"] = int(args.tuning)   # RT_TUNE_* bits (ncu captures use 32: one frame-wide shadow launch, as the class timing pass has it)
    opts = dict(base_opts, device=D.local_rank, rank=rank, world=world)
    ctx = rt.Context(scene, **opts)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    # frames in flight (peer gather / one GPU): one context per frame slot, each on its own stream; frame k is queued on slot
    # k % F before frame k - F + 1 is collected.  The nccl gather pipelines its own way (below) with one context.
    # Default: two for the BVH workloads (their deep levels are latency-bound: DESIGN.md 5b), one for the linear-scan workloads
    # (every launch fills the machine; a second frame only competes: 427.8 against 408.9 ms on test-scene-800-1024spp).
    want_f = args.frames_in_flight if getattr(args, "frames_in_flight", None) else (2 if wl["accel"] == "bvh" else 1)
    F = max(1, int(want_f)) if (world == 1 or gather == "peer") else 1
    ctxs = [ctx] + [rt.Context(scene, **opts) for _ in range(F - 1)]
    bytewidth = ctx.bytewidth
    frame_bytes = h * bytewidth
    fb_ptr, fb_bytes = ctx.framebuffer_ptr()

    # ---- how the shards become one frame (N > 1)
    frame0 = None         # rank 0: device view of the assembled frame
    nccl = None
    if world > 1 and gather == "peer":
        for i, c in enumerate(ctxs):   # one whole-frame buffer on rank 0 per frame slot
            handle = None
            if rank == 0:
                p = c.frame_create()
                handle = c.frame_export()
                if i == 0:
                    frame0 = device_view(torch, p, frame_bytes)
            handle = D.broadcast_bytes(handle, 64)
            if rank != 0:
                c.frame_import(handle)
    elif world > 1:
        nccl = dict(final=torch.empty(frame_bytes, dtype=torch.uint8, device="cuda"), gathered=torch.empty(world * fb_bytes, dtype=torch.uint8, device="cuda"),
                    local=device_view(torch, fb_ptr, fb_bytes), stream=torch.cuda.Stream(), ev_traced=torch.cuda.Event(), ev_gathered=torch.cuda.Event(),
                    pending=False)
        nccl["staging"] = torch.empty_like(nccl["local"])
        frame0 = nccl["final"]

    # nccl: the gather of frame k runs on its own stream beside the trace of frame k+1 (the shard's bytes are first copied out
    # of the context's framebuffer, which the next resolve overwrites); frame k is assembled right after trace k+1
    def drain():
        if nccl and nccl["pending"]:
            stream.wait_event(nccl["ev_gathered"])
            ctx.untile(nccl["gathered"].data_ptr(), nccl["final"].data_ptr())
            nccl["pending"] = False

    def step():
        st = ctx.render()
        if nccl:
            drain()
            nccl["staging"].copy_(nccl["local"])
            nccl["ev_traced"].record(stream)
            nccl["stream"].wait_event(nccl["ev_traced"])
            with torch.cuda.stream(nccl["stream"]):
                D.dist.all_gather_into_tensor(nccl["gathered"], nccl["staging"])
                nccl["ev_gathered"].record(nccl["stream"])
            nccl["pending"] = True
        return st

    for _ in range(max(3, warmup)):
        step()
    for c in ctxs[1:]:
        c.render()
    drain()
    D.barrier()
    # one frame by itself (no other frame in flight): what a caller waiting for this frame sees
    alone = sorted(ctx.render()["render_ms"] for _ in range(5))[2]
    D.barrier()
    sampler = ClockSampler(D.local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rays = launches = 0
    host_levels = 0
    D.barrier()
    t0 = time.perf_counter()
    ev0.record(stream)
    def account(st):
        nonlocal rays, launches, host_levels
        rays += st["rays_total"]
        launches += st["kernel_launches"] + (1 if nccl else 0)
        host_levels |= st["host_levels"]

    if F == 1:
        for _ in range(steps):
            account(step())
    else:
        queued = []
        for k in range(steps):
            c = ctxs[k % F]
            if len(queued) == F:
                account(queued.pop(0).render_end())
            c.render_begin()
            queued.append(c)
        for c in queued:
            account(c.render_end())
        torch.cuda.synchronize()   # every slot's stream (render_end has already waited for each frame's last event)
    drain()
    ev1.record(stream)
    D.barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    dev_ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    dev_ms, wall_ms = D.reduce([dev_ms, wall_ms], "MAX")
    rays_all, launches_all = D.reduce([rays, launches], "SUM")

    # ---- kernel-class times: a separate pass with one event pair per launch (not part of the headline timing)
    ctx.set_profiling(True)
    P = 3
    prof = {"steps": P, "class_ms": {}, "class_launches": {}, "rays_primary": 0, "rays_secondary": 0, "rays_shadow": 0,
            "bvh_nodes": {"closest": 0, "shadow": 0}, "bvh_prims": {"closest": 0, "shadow": 0}, "scan_tests": {"closest": 0, "shadow": 0},
            "n_planes": info["n_planes"]}
    for _ in range(P):
        st = ctx.render()
        for k in ("rays_primary", "rays_secondary", "rays_shadow"):
            prof[k] += st[k]
        for k, v in st["class_ms"].items():
            prof["class_ms"][k] = prof["class_ms"].get(k, 0.0) + v
            prof["class_launches"][k] = prof["class_launches"].get(k, 0) + st["class_launches"][k]
        for k in ("closest", "shadow"):
            prof["bvh_nodes"][k] += st["bvh_nodes_visited"][k]
            prof["bvh_prims"][k] += st["bvh_prims_tested"][k]
            prof["scan_tests"][k] += st["scan_sphere_tests"][k]
        prof["node_bytes"] = st["bvh_node_bytes"]
        prof["shadow_node_bytes"] = 64 if st["bvh_node_bytes"] == 32 else st["bvh_node_bytes"]   # four-wide nodes over the grid form
    ctx.set_profiling(False)
    D.barrier()

    # ---- correctness of the assembled multi-GPU frame: rank 0 renders the whole frame on its own GPU and compares
    frame_check = None
    if world > 1:
        ctx.render()
        if nccl:
            step(); drain()
        D.barrier()
        if rank == 0:
            got = frame0.cpu().numpy().reshape(h, bytewidth)
            one = rt.Context(scene, **dict(base_opts, device=D.local_rank))
            one.render()
            ref_rows, _, _ = one.download()
            one.close()
            ref_rows = ref_rows[:h]
            d = np.abs(got.astype(np.int16) - ref_rows.astype(np.int16))
            frame_check = {"max_lsb": int(d.max()), "identical": float((d == 0).mean()), "bytes": int(d.size),
                           "against": "the same frame rendered by one GPU (rank 0, world 1) in this run; differences come from the order of "
                                      "f64 atomic adds into a pixel only"}
        D.barrier()

    (alone,) = D.reduce([alone], "MAX")
    line = {"value": rays_all / (dev_ms * 1e-3) / 1e6, "ms_per_step": dev_ms / steps, "wall_ms_per_step": wall_ms / steps,
            "frames_in_flight": F, "frame_ms_alone": alone,
            "rays_per_step": rays_all / steps, "gpu_launches": int(launches_all), "clocks": clocks, "steps": steps, "warmup": max(3, warmup),
            "config": workload_config(name, w, h, aa, info["n_objects"], info["n_lights"]),
            "bvh_build_ms": ctx.create_stats["bvh_build_ms"], "wavefront": "host-driven levels" if host_levels else
            ("device-driven (no count read-backs); each level's shadow queries traced on a second stream beside the next level"
             if WORKLOADS[name]["accel"] == "bvh" else "device-driven (no count read-backs)")}
    if world > 1:
        line["config"]["multi_gpu"] = ("image row bands (4 rows) interleaved over ranks, scene replicated; " +
                                       ("every rank's resolve kernel stores its rows into rank 0's frame over NVLink (CUDA IPC peer memory)"
                                        if gather == "peer" else "one NCCL all_gather of the BGR8 shards per frame, pipelined one frame deep with "
                                        "the next frame's trace, then de-interleave"))
        line["frame_check"] = frame_check
    if rank == 0:
        peak, peak_src = load_peaks()
        line["roofline"] = add_ncu_traffic(roofline_for(prof, peak, peak_src, wl["accel"], hw), name, world)

    # ---- NCCL variant of the gather, for the record (N > 1, default line only)
    if want_nccl_variant and world > 1 and gather == "peer":
        ctx.frame_set(0)
        sub = measure_nccl_variant(ctx, D, steps, frame_bytes, fb_ptr, fb_bytes)
        if rank == 0:
            line["gather_nccl"] = sub

    for c in ctxs:
        c.close()
    del ctx, ctxs

    # ---- end to end through the user-facing call with host buffers
    if want_e2e:
        e2e = run_e2e(rt, D, scene, desc, info, base_opts, h, bytewidth, max(4, min(steps, 6)), text=text, in_flight=max(1, int(args.frames_in_flight or 2)))
        if rank == 0:
            line.update(e2e)
    if want_cpu and world == 1 and rank == 0:
        line["cpu_baseline"] = cpu_baseline(name)
    return line


def measure_nccl_variant(ctx, D, steps, frame_bytes, fb_ptr, fb_bytes):
    torch = D.torch
    stream = torch.cuda.current_stream()
    final = torch.empty(frame_bytes, dtype=torch.uint8, device="cuda")
    gathered = torch.empty(D.world * fb_bytes, dtype=torch.uint8, device="cuda")
    local = device_view(torch, fb_ptr, fb_bytes)
    staging = torch.empty_like(local)
    side = torch.cuda.Stream()
    ev_t, ev_g = torch.cuda.Event(), torch.cuda.Event()
    pending = [False]

    def drain():
        if pending[0]:
            stream.wait_event(ev_g)
            ctx.untile(gathered.data_ptr(), final.data_ptr())
            pending[0] = False

    def step():
        st = ctx.render()
        drain()
        staging.copy_(local)
        ev_t.record(stream)
        side.wait_event(ev_t)
        with torch.cuda.stream(side):
            D.dist.all_gather_into_tensor(gathered, staging)
            ev_g.record(side)
        pending[0] = True
        return st["rays_total"]

    for _ in range(3):
        step()
    drain()
    D.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rays = 0
    ev0.record(stream)
    for _ in range(steps):
        rays += step()
    drain()
    ev1.record(stream)
    D.barrier()
    ms, = D.reduce([ev0.elapsed_time(ev1)], "MAX")
    rays_all, = D.reduce([rays], "SUM")
    return {"value": rays_all / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": ms / steps,
            "what": "the same frames gathered with one NCCL all_gather_into_tensor per frame (side stream, one frame deep) + k_untile"}


def run_e2e(rt, D, scene, desc, info, base_opts, h, bytewidth, steps, text=None, in_flight=2):
    """End to end through rt_render / rt_render_multi with HOST buffers.  N > 1: rank 0 drives all N GPUs through the library's
    own multi-GPU call (one host thread per device inside the library), the other ranks idle at the barrier.
    cold: the spheres are rewritten from host memory before every frame (flatten, upload, device BVH build per GPU, trace,
          frame to pinned host memory);  warm: unchanged scene (the device copy kept with the rt_scene is reused).
    in_flight frames at once: that many caller threads, each with its own rt_scene and pinned frame, each making the same
    blocking call for every other frame (an animation renders frame k+1 while frame k is on its way to the host); the figure
    for one call at a time is reported beside it (`one_at_a_time_ms`)."""
    torch = D.torch
    world, rank = D.world, D.rank
    out = {}
    rt.release_cached_memory()   # the bench context's queues: rank 0 is about to allocate on every GPU
    D.host_barrier()
    if rank == 0:
        spheres = None
        if desc is not None:
            first = 0
            for ob in desc["objects"]:
                if ob["shape"] == "spheres":
                    spheres = (first, np.ascontiguousarray(ob["data"], dtype=np.float64))
                    break
                first += 1
        devices = None if world == 1 else list(range(world))
        slots = [(scene, torch.empty(h * bytewidth, dtype=torch.uint8).pin_memory())]
        for _ in range(max(1, in_flight) - 1):
            slots.append((rt.serialize.deserialize(text) if desc is None else rt.Scene(desc), torch.empty(h * bytewidth, dtype=torch.uint8).pin_memory()))

        def one(slot, cold):
            sc, host = slot
            if cold and spheres is not None:
                sc.update_spheres(spheres[0], spheres[1])
            st = rt.raytrace.render_into(sc, host.data_ptr(), devices=devices, **dict(base_opts, device=D.local_rank))
            return st["rays_total"]

        def timed(cold, n_slots, frames):
            """frames calls in all, shared out over n_slots caller threads; -> (rays, seconds)"""
            per = [frames // n_slots + (1 if i < frames % n_slots else 0) for i in range(n_slots)]
            rays = [0] * n_slots
            gate = threading.Barrier(n_slots + 1)

            def run(i):
                gate.wait()
                for _ in range(per[i]):
                    rays[i] += one(slots[i], cold)
                gate.wait()

            th = [threading.Thread(target=run, args=(i,)) for i in range(n_slots)]
            for t in th:
                t.start()
            gate.wait()
            t0 = time.perf_counter()
            gate.wait()
            sec = time.perf_counter() - t0
            for t in th:
                t.join()
            return sum(rays), sec

        for label, cold in (("e2e", True), ("e2e_warm", False)):
            for sl in slots:
                one(sl, cold); one(sl, cold)
            rays1, sec1 = timed(cold, 1, steps)
            raysF, secF = timed(cold, len(slots), steps) if len(slots) > 1 else (rays1, sec1)
            # the caller picks what serves it best: with many GPUs and a big scene to re-upload, two callers' host work (flatten,
            # one upload thread per device each) can get in each other's way -- both figures are reported, the faster is the line's
            used = len(slots) if secF < sec1 else 1
            rays, sec = (raysF, secF) if used > 1 else (rays1, sec1)
            per_dev = scene_h2d_bytes(info, desc) if (cold and spheres is not None) else 0
            out[label] = {"value": rays / sec / 1e6, "unit": "Mrays/s", "ms_per_step": sec / steps * 1e3,
                          "h2d_bytes_per_step": per_dev * world, "d2h_bytes_per_step": h * bytewidth,
                          "frames_in_flight": used, "one_at_a_time_ms": sec1 / steps * 1e3, "two_in_flight_ms": secF / steps * 1e3 if len(slots) > 1 else None,
                          "what": ("caller rewrites the spheres from host memory (rt_scene_update_spheres) -> " if cold and spheres is not None else "unchanged scene -> ") +
                                  ("rt_render" if world == 1 else "rt_render_multi (%d GPUs driven by one process, one thread per device, each device DMAs "
                                   "its bands to the host frame)" % world) +
                                  (": flatten, upload, device BVH build, trace, frame to pinned host memory" if cold and spheres is not None
                                   else ": device copy of the scene reused, trace, frame to pinned host memory") +
                                  ("; %d caller threads, each with its own rt_scene and host frame, keep %d frames in flight" % (used, used)
                                   if used > 1 else "; one call at a time")}
        del slots
        rt.release_cached_memory()
    D.host_barrier()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--fp", default="strict", choices=["strict", "fast"])
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"])
    ap.add_argument("--frames-in-flight", type=int, default=None,
                    help="frames queued at once per GPU (one context each); default: 2 for BVH workloads, 1 for linear-scan workloads")
    ap.add_argument("--tuning", type=int, default=0, help="RT_TUNE_* bits for every context (0 = the defaults)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-also", action="store_true")
    args = ap.parse_args()
    claim_stdout()
    if args.impl == "reference":
        run_reference(args)
        return
    rt = pkg()
    if rt.device_count() < 1:
        raise SystemExit("no CUDA device visible: the product has no CPU path")
    D = Dist(args)
    hw = rt.measure_peaks(D.local_rank)
    m = measure(args.workload, args, D, args.steps, args.warmup, hw, want_e2e=not args.no_e2e, want_cpu=not args.no_cpu_baseline,
                gather=args.gather, want_nccl_variant=(args.workload == DEFAULT_WORKLOAD))
    also = None
    if not args.no_also and args.workload == DEFAULT_WORKLOAD:
        a = measure(ALSO_WORKLOAD, args, D, min(args.steps, 5), 3, hw, want_e2e=not args.no_e2e, want_cpu=not args.no_cpu_baseline, gather=args.gather)
        if D.rank == 0:
            a.update({"metric": "Mrays/s (primary+secondary+shadow) at 8K", "unit": "Mrays/s", "n_gpus": args.gpus})
            also = {ALSO_WORKLOAD: a}
    if D.rank == 0:
        line = {"metric": METRIC, "value": m.pop("value"), "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": max(3, args.warmup),
                "ms_per_step": m.pop("ms_per_step"), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "fp_mode": args.fp, "data": "synthetic"}
        m.pop("steps", None); m.pop("warmup", None)
        line.update(m)
        line["host_threads"] = host_threads()
        if also:
            line["also"] = also
        emit(line)
    D.close()


if __name__ == "__main__":
    main()
