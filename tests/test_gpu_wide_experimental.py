"""Opt-in GPU parity tests of the experimental four-wide shadow walk (RAYTRACE_B200_BVH_WIDE=1).

The path is off by default: its arithmetic is covered on the CPU by tests/test_emu_parity.py::test_wide_*, and on a B200
only the 65- and 3000-sphere cases below have run so far, in mode 1 and with the first (branchy) formulation of the
walk (both green, with the round's last GPU seconds); the branch-free walks and mode 2 have only run in the emulation, and
nothing has been timed.  These tests only run with RAYTRACE_B200_TEST_WIDE=1, so that a fault in an unmeasured kernel cannot poison the
CUDA context of the regular `-m gpu` run:

    RAYTRACE_B200_TEST_WIDE=1 python -m pytest tests/test_gpu_wide_experimental.py -m gpu -q
"""
import os

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("RAYTRACE_B200_TEST_WIDE") != "1", reason="opt-in: RAYTRACE_B200_TEST_WIDE=1")]


@pytest.mark.parametrize("mode", ["1", "2"])  # 1: shadow queries over the wide nodes, 2: closest-hit queries too (never run on hardware yet)
@pytest.mark.parametrize("n,extent", [(1, 3.0), (3, 3.0), (65, 5.0), (3000, 20.0), (60000, 90.0)])
def test_wide_shadow_walk_renders_the_binary_walks_frame(gpu, scenes, monkeypatch, n, extent, mode):
    desc = scenes.random_spheres(n, 100 + n, extent, 320, 180)
    monkeypatch.delenv("RAYTRACE_B200_BVH_WIDE", raising=False)
    a = gpu.raytrace.render(gpu.Scene(desc), max_depth=6, seed=9, accel=gpu.ACCEL_BVH)
    monkeypatch.setenv("RAYTRACE_B200_BVH_WIDE", mode)
    b = gpu.raytrace.render(gpu.Scene(desc), max_depth=6, seed=9, accel=gpu.ACCEL_BVH)
    assert np.array_equal(a.ids, b.ids)
    assert a.stats["rays_total"] == b.stats["rays_total"] and a.stats["rays_shadow"] == b.stats["rays_shadow"]
    # same rays, same occlusion answers; only the order of the f64 atomic adds into a pixel may differ
    assert (a.rows != b.rows).mean() < 1e-5
    if n >= 3000:
        va, vb = a.stats["bvh_nodes_visited"]["shadow"], b.stats["bvh_nodes_visited"]["shadow"]
        assert 0.35 * va < vb < 0.7 * va, (va, vb)
        ca, cb = a.stats["bvh_nodes_visited"]["closest"], b.stats["bvh_nodes_visited"]["closest"]
        assert (0.35 * ca < cb < 0.75 * ca) if mode == "2" else ca == cb, (ca, cb)


@pytest.mark.parametrize("mode", ["1", "2"])
def test_wide_shadow_walk_against_the_oracle(gpu, orc, scenes, monkeypatch, mode):
    desc = scenes.random_spheres(1500, 0x5EED0003, 25.0, 160, 90)
    ref = orc.OracleScene(desc).render(max_depth=8, seed=3)
    monkeypatch.setenv("RAYTRACE_B200_BVH_WIDE", mode)
    fr = gpu.raytrace.render(gpu.Scene(desc), max_depth=8, seed=3, accel=gpu.ACCEL_BVH)
    assert np.array_equal(fr.ids, ref["ids"])
    diff = np.abs(fr.rows.astype(np.int16) - ref["bgr"].astype(np.int16))
    assert diff.max() <= 1 and (diff == 0).mean() >= 0.999
    assert fr.stats["rays_total"] == ref["stats"]["rays"]
