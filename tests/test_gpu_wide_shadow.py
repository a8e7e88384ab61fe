"""GPU parity tests of the four-wide shadow walk (rt_types.h BvhWNode, rt_device.cuh trav_descend_wide; the default for shadow
queries wherever the grid nodes are in use) against the binary walk it is derived from (RT_TUNE_BINARY_SHADOW) and the oracle.
First hardware run and timings: profiles/r02_v1_wide_ab.txt."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,extent", [(1, 3.0), (3, 3.0), (65, 5.0), (3000, 20.0), (60000, 90.0)])
def test_wide_shadow_walk_renders_the_binary_walks_frame(gpu, scenes, n, extent):
    desc = scenes.random_spheres(n, 100 + n, extent, 320, 180)
    a = gpu.raytrace.render(gpu.Scene(desc), max_depth=6, seed=9, accel=gpu.ACCEL_BVH, tuning=gpu.TUNE_BINARY_SHADOW)
    b = gpu.raytrace.render(gpu.Scene(desc), max_depth=6, seed=9, accel=gpu.ACCEL_BVH)
    assert np.array_equal(a.ids, b.ids)
    assert a.stats["rays_total"] == b.stats["rays_total"] and a.stats["rays_shadow"] == b.stats["rays_shadow"]
    # same rays, same occlusion answers; only the order of the f64 atomic adds into a pixel may differ
    assert (a.rows != b.rows).mean() < 1e-5
    assert a.stats["bvh_nodes_visited"]["closest"] == b.stats["bvh_nodes_visited"]["closest"]
    if n >= 3000:
        assert b.stats["bvh_node_bytes"] == 32
        va, vb = a.stats["bvh_nodes_visited"]["shadow"], b.stats["bvh_nodes_visited"]["shadow"]
        assert 0.35 * va < vb < 0.7 * va, (va, vb)


def test_wide_shadow_walk_against_the_oracle(gpu, orc, scenes):
    desc = scenes.random_spheres(1500, 0x5EED0003, 25.0, 160, 90)
    ref = orc.OracleScene(desc).render(max_depth=8, seed=3)
    fr = gpu.raytrace.render(gpu.Scene(desc), max_depth=8, seed=3, accel=gpu.ACCEL_BVH)
    assert fr.stats["bvh_node_bytes"] == 32
    assert np.array_equal(fr.ids, ref["ids"])
    diff = np.abs(fr.rows.astype(np.int16) - ref["bgr"].astype(np.int16))
    assert diff.max() <= 1 and (diff == 0).mean() >= 0.999
    assert fr.stats["rays_total"] == ref["stats"]["rays"]
