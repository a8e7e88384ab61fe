"""N > 1 path on CPU: world_size 2 and 3 over gloo (the GPU run uses the same plumbing over NCCL).

Each rank renders its interleaved row bands (shard mapping of rt_device.cuh slot_to_pixel, run through the CPU emulation
harness), the quantised BGR bands are all-gathered rank-major -- the only collective of the path, one per frame --
and the de-interleave (untile_source_word, the arithmetic of k_untile) must reproduce the single-rank frame exactly,
padding rows and row padding bytes included.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, width, height, band_rows, out_path):
    for p in (ROOT, HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    import importlib
    import emu
    scenes = importlib.import_module("rust-raytrace_b200.scenes")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        desc = scenes.random_spheres(300, 9, 12.0, width, height)
        sc = emu.Scene(desc)
        kw = dict(max_depth=4, seed=6, band_rows=band_rows)
        shard = emu.render_shard(sc, rank=rank, world=world, **kw)
        local = torch.from_numpy(shard["rows"].reshape(-1).copy())
        gathered = torch.empty(world * local.numel(), dtype=torch.uint8)
        dist.all_gather_into_tensor(gathered, local)
        rays = torch.tensor([shard["stats"]["rays_total"]], dtype=torch.int64)
        dist.all_reduce(rays)
        if rank == 0:
            final = emu.untile(gathered.numpy(), height, shard["bytewidth"], band_rows or 4, world, shard["local_rows"])
            whole = emu.render(sc, **kw)
            np.savez(out_path, final=final, whole=whole["bgr"], rays=int(rays[0]), whole_rays=whole["stats"]["rays_total"],
                     local_rows=shard["local_rows"])
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,width,height,band_rows", [(2, 96, 70, 0), (3, 50, 37, 8), (2, 33, 8, 4)])
def test_row_band_shards_gather_and_untile(tmp_path, world, width, height, band_rows):
    out = str(tmp_path / "out.npz")
    mp.spawn(_worker, args=(world, _free_port(), width, height, band_rows, out), nprocs=world, join=True)
    r = np.load(out)
    assert r["final"].shape == r["whole"].shape == (height, (3 * width + 3) // 4 * 4)
    assert np.array_equal(r["final"], r["whole"])          # every pixel is rendered by exactly one rank, identically
    assert int(r["rays"]) == int(r["whole_rays"])           # shards partition the work: ray counts add up
    band = band_rows or 4
    n_bands = (height + band - 1) // band
    assert int(r["local_rows"]) == (n_bands + world - 1) // world * band


def test_untile_is_a_permutation():
    """untile_source_word maps every final word to a distinct gathered word (size-independent property, 8K sizes)"""
    for p in (ROOT, HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    import emu
    for world, height, band in [(8, 4320, 4), (4, 2160, 4), (3, 37, 8), (1, 10, 4)]:
        bw = 64
        n_bands = (height + band - 1) // band
        local_rows = (n_bands + world - 1) // world * band
        gathered = np.arange(world * local_rows * bw // 4, dtype=np.uint32).view(np.uint8)
        out = emu.untile(gathered, height, bw, band, world, local_rows).view(np.uint32).reshape(-1)
        assert len(np.unique(out)) == out.size
        # row y of the image lives on rank (y // band) % world at local row (y // band // world) * band + y % band
        y = np.arange(height)
        src_row = ((y // band) % world) * local_rows + (y // band // world) * band + y % band
        assert np.array_equal(out.reshape(height, bw // 4)[:, 0], src_row * (bw // 4))
