"""all_gather_into_tensor timing at the frame sizes the path gathers (4K: 24.9 MB total, 8K: 99.5 MB total); run under torchrun."""
import os, time, torch, torch.distributed as dist
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
for total in (24883200, 99532800):
    per = total // world
    src = torch.full((per,), rank, dtype=torch.uint8, device="cuda")
    dst = torch.empty(per * world, dtype=torch.uint8, device="cuda")
    for _ in range(5):
        dist.all_gather_into_tensor(dst, src)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        dist.all_gather_into_tensor(dst, src)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    if rank == 0:
        print("all_gather total %.1f MB world %d: %.3f ms  (%.1f GB/s received per rank)" % (total / 1e6, world, ms, total * (world - 1) / world / ms / 1e6), flush=True)
    # one-shot with skew: rank r sleeps r*0.2 ms first
dist.destroy_process_group()
