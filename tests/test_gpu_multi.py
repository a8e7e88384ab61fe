"""Multi-GPU frames behind the C ABI (rt_multi_*, rt_render_multi): the frame sharded by row bands over the
devices of the node, gathered three ways, must be the frame one GPU renders.

On a one-GPU machine the DIRECT and PEER modes still run with the same ordinal listed several times (the shards share the
GPU), so the sharding, the strided band copies and the resolve-into-the-whole-frame path are exercised by the round-end GPU
run; the tests marked `multi` need two or more GPUs (real NVLink peer stores, real ncclAllGather).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _same_frame(a_rows, b_rows, width):
    # same rays, same answers; only the order of the f64 atomic adds into a pixel may differ between two renders
    assert a_rows.shape == b_rows.shape
    assert (a_rows != b_rows).mean() < 1e-5
    d = np.abs(a_rows.astype(np.int16) - b_rows.astype(np.int16))
    assert d.max() <= 1
    assert np.array_equal(a_rows[:, 3 * width:], b_rows[:, 3 * width:])  # row padding bytes


@pytest.mark.parametrize("gather", ["direct", "peer"])
@pytest.mark.parametrize("shards,width,height", [(1, 200, 150), (2, 200, 150), (3, 35, 37), (5, 321, 98)])
def test_sharded_frame_equals_one_gpu_frame(gpu, scenes, gather, shards, width, height):
    """heights that are not a multiple of band*world, widths whose rows are not a multiple of 16 bytes (k_resolve's short
    last piece), more shards than some bands"""
    desc = scenes.random_spheres(500, 9, 15.0, width, height)
    sc = gpu.Scene(desc)
    whole = gpu.raytrace.render(sc, max_depth=4, seed=6)
    mode = gpu.GATHER_DIRECT if gather == "direct" else gpu.GATHER_PEER
    fr = gpu.raytrace.render_multi(sc, [0] * shards, gather=mode, max_depth=4, seed=6)
    _same_frame(fr.rows, whole.rows, width)
    assert np.array_equal(fr.ids, whole.ids)
    assert fr.stats["rays_total"] == whole.stats["rays_total"] and fr.stats["paths"] == whole.stats["paths"]


def test_multi_context_renders_repeatedly_and_keeps_the_frame_on_the_device(gpu, scenes):
    import torch
    desc = scenes.materials_showcase(160, 100, 2)
    sc = gpu.Scene(desc)
    whole = gpu.raytrace.render(sc, max_depth=5, seed=4)
    mc = gpu.MultiContext(sc, [0, 0, 0], gather=gpu.GATHER_PEER, max_depth=5, seed=4)
    assert mc.gather == gpu.GATHER_PEER
    a = mc.render(want_ids=True)
    b = mc.render()
    _same_frame(a.rows, whole.rows, 160)
    _same_frame(b.rows, whole.rows, 160)
    assert np.array_equal(a.ids, whole.ids)
    st = mc.render_to(None)   # frame stays in device 0's buffer
    assert st["rays_total"] == whole.stats["rays_total"]
    ptr = mc.frame_device_ptr(0)
    assert ptr

    class H:
        __cuda_array_interface__ = {"shape": (whole.rows.size,), "typestr": "|u1", "data": (ptr, False), "version": 2}
    dev = torch.as_tensor(H(), device="cuda").cpu().numpy().reshape(whole.rows.shape)
    _same_frame(dev, whole.rows, 160)
    assert sum(mc.device_stats(i)["local_rows"] for i in range(3)) >= 100
    mc.close()


def test_nccl_gather_on_one_device(gpu, scenes):
    """RT_GATHER_NCCL with a single rank: no collective, the frame is copied into the per-device frame buffer"""
    desc = scenes.random_spheres(300, 4, 12.0, 96, 64)
    sc = gpu.Scene(desc)
    whole = gpu.raytrace.render(sc, max_depth=3, seed=2)
    fr = gpu.raytrace.render_multi(sc, [0], gather=gpu.GATHER_NCCL, max_depth=3, seed=2)
    _same_frame(fr.rows, whole.rows, 96)
    with pytest.raises(gpu.RtError):
        gpu.raytrace.render_multi(sc, [0, 0], gather=gpu.GATHER_NCCL, max_depth=3, seed=2)   # one rank per GPU
    with pytest.raises(gpu.RtError):
        gpu.raytrace.render_multi(sc, [gpu.device_count()], max_depth=3, seed=2)              # no such device


@pytest.mark.parametrize("gather", ["direct", "peer", "nccl"])
def test_real_multi_gpu_frame(gpu, scenes, gather):
    """two or more GPUs: NVLink peer stores / ncclAllGather + k_untile / per-device strided DMA against the one-GPU frame"""
    n = gpu.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    desc = scenes.random_spheres(20000, 0x5EED0003, 60.0, 1280, 722)
    sc = gpu.Scene(desc)
    whole = gpu.raytrace.render(sc, max_depth=8, seed=3, accel=gpu.ACCEL_BVH)
    mode = {"direct": gpu.GATHER_DIRECT, "peer": gpu.GATHER_PEER, "nccl": gpu.GATHER_NCCL}[gather]
    for devs in sorted({2, n}):
        mc = gpu.MultiContext(sc, devs, gather=mode, max_depth=8, seed=3, accel=gpu.ACCEL_BVH)
        for _ in range(2):
            fr = mc.render(want_ids=True)
            _same_frame(fr.rows, whole.rows, 1280)
            assert np.array_equal(fr.ids, whole.ids)
            assert fr.stats["rays_total"] == whole.stats["rays_total"]
        mc.close()


def test_render_multi_keeps_its_state_between_calls_and_follows_scene_edits(gpu, scenes):
    """rt_render_multi is one-shot for the caller; the library keeps the rt_multi it made (threads, contexts, streams) for the
    next call with the same scene handle / options / devices and re-binds it to the scene's current revision: moved spheres are
    re-uploaded, a changed structure (new light) makes a new one, a freed scene drops it."""
    import ctypes as C
    desc = scenes.random_spheres(2000, 17, 25.0, 160, 100)
    sc = gpu.Scene(desc)
    kw = dict(max_depth=4, seed=7, accel=gpu.ACCEL_BVH)
    one = gpu.raytrace.render(sc, **kw)
    a = gpu.raytrace.render_multi(sc, [0, 0, 0], gather=gpu.GATHER_DIRECT, **kw)
    b = gpu.raytrace.render_multi(sc, [0, 0, 0], gather=gpu.GATHER_DIRECT, **kw)      # kept state
    for fr in (a, b):
        _same_frame(fr.rows, one.rows, 160)
        assert np.array_equal(fr.ids, one.ids) and fr.stats["rays_total"] == one.stats["rays_total"]
    assert a.stats["bvh_build_ms"] == 0 and b.stats["bvh_build_ms"] == 0             # the scene's device copy came from `one`
    # the spheres move: same handle, new revision
    data = np.array(desc["objects"][1]["data"])
    data[:, 1] += 0.75
    sc.update_spheres(1, data)
    moved = gpu.raytrace.render_multi(sc, [0, 0, 0], gather=gpu.GATHER_DIRECT, **kw)
    ref = gpu.raytrace.render(sc, **kw)
    _same_frame(moved.rows, ref.rows, 160)
    assert np.array_equal(moved.ids, ref.ids) and (moved.ids != one.ids).mean() > 0.01
    assert moved.stats["bvh_build_ms"] > 0                                            # re-uploaded and rebuilt by the re-bound contexts
    # a new light changes what the queues were sized for
    L = gpu.lib()
    assert L.rt_scene_add_light_point(sc.handle, (C.c_double * 3)(5.0, 30.0, 5.0), (C.c_double * 3)(0.5, 0.5, 0.5)) == 0
    lit = gpu.raytrace.render_multi(sc, [0, 0, 0], gather=gpu.GATHER_DIRECT, **kw)
    ref2 = gpu.raytrace.render(sc, **kw)
    _same_frame(lit.rows, ref2.rows, 160)
    assert lit.stats["rays_shadow"] == ref2.stats["rays_shadow"] > moved.stats["rays_shadow"]
    # another scene (the first one freed: whatever was kept for it is dropped with it)
    del sc
    sc2 = gpu.Scene(scenes.random_spheres(500, 3, 12.0, 160, 100))
    x = gpu.raytrace.render_multi(sc2, [0, 0, 0], gather=gpu.GATHER_DIRECT, **kw)
    y = gpu.raytrace.render(sc2, **kw)
    _same_frame(x.rows, y.rows, 160)
    gpu.release_cached_memory()
