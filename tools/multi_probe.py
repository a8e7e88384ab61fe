"""Where an N-GPU step goes: per rank, device ms of render / all_gather / untile, plus the bare all_gather at the same size
(run under torchrun; development aid)."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
rt = importlib.import_module("rust-raytrace_b200")
scenes = importlib.import_module("rust-raytrace_b200.scenes")
which = sys.argv[1] if len(sys.argv) > 1 else "1m8k"
d, kw = {"10k": (scenes.spheres_10k, dict(max_depth=8, seed=3)), "1m8k": (lambda: scenes.spheres_1m(7680, 4320), dict(max_depth=8, seed=4))}[which]
ctx = rt.Context(rt.Scene(d()), accel=rt.ACCEL_BVH, device=lr, rank=rank, world=world, **kw)
stream = torch.cuda.current_stream()
ctx.set_stream(stream.cuda_stream)
fb_ptr, fb_bytes = ctx.framebuffer_ptr()
class H: pass
h = H(); h.__cuda_array_interface__ = {"shape": (fb_bytes,), "typestr": "|u1", "data": (fb_ptr, False), "version": 2}
local_fb = torch.as_tensor(h, device="cuda")
gathered = torch.empty(world * fb_bytes, dtype=torch.uint8, device="cuda")
final = torch.empty(ctx.height * ctx.bytewidth, dtype=torch.uint8, device="cuda")
ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
for it in range(6):
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    ev[0].record(); st = ctx.render(); ev[1].record()
    dist.all_gather_into_tensor(gathered, local_fb); ev[2].record()
    ctx.untile(gathered.data_ptr(), final.data_ptr()); ev[3].record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) * 1e3
    if it >= 3:
        print("rank %d it %d: render %.2f (stats %.2f) gather %.2f untile %.2f wall %.2f ms, fb %d B" % (rank, it, ev[0].elapsed_time(ev[1]), st["render_ms"], ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3]), wall, fb_bytes), flush=True)
for _ in range(3):
    dist.all_gather_into_tensor(gathered, local_fb)
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
ev[0].record()
for _ in range(10):
    dist.all_gather_into_tensor(gathered, local_fb)
ev[1].record(); torch.cuda.synchronize()
if rank == 0:
    print("bare all_gather of %d B x %d ranks: %.3f ms" % (fb_bytes, world, ev[0].elapsed_time(ev[1]) / 10), flush=True)
dist.destroy_process_group()
