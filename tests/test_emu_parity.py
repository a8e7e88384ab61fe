"""CPU-side parity of the product's per-ray arithmetic (rust-raytrace_b200/csrc/rt_device.cuh) against the oracle.

tests/emu compiles the product header with g++ and drives it with a serial loop shaped like the GPU wavefront
(tests/emu/README.md).  It is a test of the header, not a fallback: what is parallel (kernels, atomics, TMA staging,
the device BVH build) is only covered by `pytest -m gpu`.

Bars as on the GPU: primary hit ids bit-exact, bytes within +-1 LSB and >= 99.9 % identical, ray counts equal; in
strict arithmetic the intersection results (t, normal) are bit-identical.
"""
import numpy as np
import pytest

import emu
from test_gpu_parity import REF_TEXT_SCENE


def check(got, ref, what):
    assert np.array_equal(got["ids"], ref["ids"]), "%s: primary ids differ in %d pixels" % (what, int((got["ids"] != ref["ids"]).sum()))
    diff = np.abs(got["bgr"].astype(np.int16) - ref["bgr"].astype(np.int16))
    assert diff.max() <= 1, "%s: %d LSB" % (what, int(diff.max()))
    same = float((diff == 0).mean())
    assert same >= 0.999, "%s: %.5f identical" % (what, same)
    assert got["stats"]["rays_total"] == ref["stats"]["rays"], (what, got["stats"]["rays_total"], ref["stats"]["rays"])
    assert got["stats"]["rays_shadow"] == ref["stats"]["shadow_rays"]
    return same


def _rays(rng, n, extent, ymax):
    o = np.stack([rng.uniform(-extent, extent, n), rng.uniform(0.1, ymax, n), rng.uniform(-extent, extent, n)], axis=1)
    d = rng.normal(size=(n, 3))
    d[: n // 2] /= np.linalg.norm(d[: n // 2], axis=1, keepdims=True)
    return np.concatenate([o, d], axis=1)


def test_quantiser(orc):
    v, a = orc.srgb_tables()
    vals = np.concatenate([a, np.nextafter(a, -np.inf), np.nextafter(a, np.inf), v, np.random.default_rng(1).uniform(-0.1, 1.2, 20001),
                           [0.0, -1.0, 1.0, 5.0, np.nan, np.inf, -np.inf, -0.0]])
    vals = np.resize(vals, (vals.size // 3) * 3).reshape(-1, 3)
    assert np.array_equal(emu.quantize_bgr(vals), orc.quantize_bgr(vals))


@pytest.mark.parametrize("accel", ["linear", "bvh"])
def test_intersect_is_bit_identical(rt, orc, scenes, accel):
    desc = scenes.random_spheres(2000, 77, 25.0, 8, 8)
    rays = _rays(np.random.default_rng(5), 30000, 30.0, 9.0)
    data = desc["objects"][1]["data"]
    inside = np.concatenate([data[:1000, :3], np.random.default_rng(6).normal(size=(1000, 3))], axis=1)  # rays starting inside spheres
    surf = data[1000:2000, :3] + data[1000:2000, 3:4] * np.array([0.0, 1.0, 0.0])                         # rays leaving sphere tops
    leaving = np.concatenate([surf, np.random.default_rng(8).normal(size=(1000, 3))], axis=1)
    rays = np.concatenate([rays, inside, leaving])
    ids, t, nrm = emu.trace_rays(emu.Scene(desc), rays, accel=rt.ACCEL_BVH if accel == "bvh" else rt.ACCEL_LINEAR)
    rid, rt_, rn = orc.OracleScene(desc).trace_rays(rays)
    assert np.array_equal(ids, rid) and np.array_equal(t, rt_) and np.array_equal(nrm, rn)
    assert (ids >= 0).mean() > 0.3


def test_bvh_equals_linear_scan_on_degenerate_input(rt, orc):
    """ties, duplicates, nested and huge/tiny spheres, NaN/inf/zero rays: the BVH must return the scan's object index"""
    S = rt.scene
    b = S.SceneBuilder()
    m = b.add_material(S.PhongMaterial((0.5, 0.5, 0.5), (0, 0, 0), 1.0, (0, 0, 0)))
    b.add_object(S.Plane((0, -1, 0), (0, 1, 0)), m)
    rng = np.random.default_rng(11)
    c = rng.uniform(-5, 5, size=(60, 3))
    r = rng.uniform(0.05, 1.5, size=60)
    data = np.concatenate([np.c_[c, r], np.c_[c, r],                        # every sphere twice (coincident: lower index wins)
                           np.c_[c[:20], r[:20] * 0.5],                      # nested
                           [[0, 0, 0, 40.0], [0, 0, 0, 1e-6], [1e6, 0, 0, 1.0]]])
    b.add_spheres(data, np.zeros(len(data), dtype=np.uint32))
    b.camera = S.SimplePerspectiveCamera.new((0, 0, 20), (0, 0, -1), (0, 1, 0), 2.0)
    b.background = S.SolidColorBackground((0, 0, 0))
    b.options = S.Options(8, 8, 1)
    desc = b.build()
    rays = _rays(rng, 20000, 8.0, 6.0)
    rays[:5000, :3] = c[rng.integers(0, 60, 5000)]                           # from sphere centres
    special = np.array([[0, 0, 0, 0, 0, 0], [np.nan, 0, 0, 0, 0, 1], [0, 0, 0, np.nan, 0, 1], [0, 0, 0, np.inf, 0, 0],
                        [0, 0, 100, 0, 0, -1e-300], [0, 0, 100, 0, 0, -1e300], [1e300, 0, 0, -1, 0, 0], [0, -1, 5, 1, 0, 0]])
    rays = np.concatenate([rays, special])
    a = emu.trace_rays(emu.Scene(desc), rays, accel=rt.ACCEL_LINEAR)
    bb = emu.trace_rays(emu.Scene(desc), rays, accel=rt.ACCEL_BVH)
    ref = orc.OracleScene(desc).trace_rays(rays)
    for got in (a, bb):
        assert np.array_equal(got[0], ref[0])
        assert np.array_equal(got[1], ref[1], equal_nan=True)
        assert np.array_equal(got[2], ref[2], equal_nan=True)
    n_nodes, depth = emu.bvh_stats(emu.Scene(desc))
    assert n_nodes == len(data) - 1 and depth < 96


@pytest.mark.parametrize("accel", ["linear", "bvh"])
def test_shadow_any_hit_equals_closest_hit_rule(rt, orc, scenes, accel):
    """src/raytrace.rs:43-50: occluded iff the CLOSEST hit exists and (no range or t*t < r2); the kernels use any-hit"""
    desc = scenes.random_spheres(800, 5, 12.0, 8, 8)
    rng = np.random.default_rng(2)
    rays = _rays(rng, 20000, 14.0, 8.0)
    r2 = np.where(rng.uniform(size=20000) < 0.3, -1.0, rng.uniform(0.0, 30.0, 20000) ** 2)
    occ = emu.shadow_rays(emu.Scene(desc), rays, r2, accel=rt.ACCEL_BVH if accel == "bvh" else rt.ACCEL_LINEAR)
    ids, t, _ = orc.OracleScene(desc).trace_rays(rays)
    with np.errstate(invalid="ignore"):
        want = (ids >= 0) & ((r2 < 0) | (t * t < r2))
    assert np.array_equal(occ, want)
    assert 0.1 < want.mean() < 0.9


def test_render_cornell_text(rt, orc):
    from oracle import scene_text
    desc = scene_text.deserialize(REF_TEXT_SCENE)
    kw = dict(width=120, height=100, antialias=3, seed=21)
    check(emu.render(emu.Scene.parse(REF_TEXT_SCENE), **kw), orc.OracleScene(desc).render(**kw), "cornell")


@pytest.mark.parametrize("accel", ["linear", "bvh"])
def test_render_all_materials(rt, orc, scenes, accel):
    desc = scenes.materials_showcase(128, 96, 2)
    ref = orc.OracleScene(desc).render(max_depth=5, seed=4)
    got = emu.render(emu.Scene(desc), max_depth=5, seed=4, accel=rt.ACCEL_BVH if accel == "bvh" else rt.ACCEL_LINEAR)
    same = check(got, ref, "showcase/" + accel)
    # the accumulators themselves agree to re-association error (path-throughput form vs recursion)
    ok = np.isfinite(ref["accum"])
    assert np.allclose(got["accum"][ok] / 2.0, ref["accum"][ok], rtol=1e-11, atol=1e-13)  # product keeps the SUM over the 2 aa samples
    assert same > 0.9999


def test_render_spheres_depth8(rt, orc, scenes):
    desc = scenes.random_spheres(1500, 0x5EED0003, 25.0, 160, 90)
    ref = orc.OracleScene(desc).render(max_depth=8, seed=3)
    check(emu.render(emu.Scene(desc), max_depth=8, seed=3, accel=rt.ACCEL_BVH), ref, "spheres1500")
    assert ref["stats"]["rays"] > 5 * 160 * 90


def test_render_skybox_dof_area(rt, orc, scenes):
    S = rt.scene
    cam = S.DepthOfFieldCamera.new(S.SimplePerspectiveCamera.new((0, 4, 14), (0, -0.2, -1), (0, 1, 0), 2.0), 12.0, 0.15, 2)
    desc = scenes.random_spheres(120, 5, 10.0, 100, 72, antialias=2, reflective_only=True, background=S.SkyboxBackground(*scenes.skybox_faces(64, 8)), camera=cam)
    desc["lights"].append(S.AreaLight((-3.0, 9.0, -3.0), (6.0, 0.0, 0.0), (0.0, 0.0, 6.0), (0.6, 0.6, 0.5)))
    ref = orc.OracleScene(desc).render(max_depth=4, seed=8)
    got = emu.render(emu.Scene(desc), max_depth=4, seed=8)
    check(got, ref, "skybox/dof/area")  # same libm on both sides here, so this is exact on the CPU


def test_depth_and_significance_edges(rt, orc, scenes):
    desc = scenes.materials_showcase(48, 36, 1)
    for md, ms in [(0, 1 / 512), (1, 1 / 512), (12, 1 / 512), (6, 0.0), (6, 0.2)]:
        ref = orc.OracleScene(desc).render(max_depth=md, min_significance=ms, seed=1)
        check(emu.render(emu.Scene(desc), max_depth=md, min_significance=ms, seed=1), ref, "md%d ms%g" % (md, ms))


def test_antialias_zero_and_empty_scene(rt, orc):
    S = rt.scene
    b = S.SceneBuilder()
    b.camera = S.SimplePerspectiveCamera.new((0, 0, 0), (0, 0, -1), (0, 1, 0), 1.0)
    b.background = S.SolidColorBackground((0.051, 0.18, 0.5))
    b.options = S.Options(5, 3, 1)   # width 5: bytewidth 16, one padding byte per row
    desc = b.build()
    got = emu.render(emu.Scene(desc))
    ref = orc.OracleScene(desc).render()
    check(got, ref, "empty scene")
    assert (got["ids"] == -1).all() and got["bgr"][0, :15].tolist() == [188, 118, 64] * 5 and got["bgr"][0, 15] == 0
    z = emu.render(emu.Scene(desc), antialias=0)
    assert (z["bgr"][:, :15] == 255).all()


@pytest.mark.parametrize("accel", ["linear", "bvh"])
def test_far_origins_advance_to_the_scene(rt, orc, scenes, accel):
    """rays that start 1e3 .. 1e9 units away (ground-plane hits near the horizon): the BVH filter advances them to the
    sphere bounds first; results must still be the scan's, bit for bit, including the scan's own rounding artefacts"""
    desc = scenes.random_spheres(600, 21, 15.0, 8, 8)
    rng = np.random.default_rng(4)
    n = 6000
    target = np.stack([rng.uniform(-15, 15, n), rng.uniform(0.0, 6.0, n), rng.uniform(-15, 15, n)], 1)
    dirs = rng.normal(size=(n, 3))
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    dist = 10.0 ** rng.uniform(3, 9, n)
    o = target - dirs * dist[:, None]
    o[:, 1] = np.abs(o[:, 1])                       # keep the origins above the ground plane
    d = target - o
    d[: n // 2] /= np.linalg.norm(d[: n // 2], axis=1, keepdims=True)   # half unit, half un-normalised (|d| up to 1e9)
    away = np.concatenate([o[:500], -d[:500]], 1)   # pointing away from the scene: must miss everything but the plane
    rays = np.concatenate([np.concatenate([o, d], 1), away])
    ids, t, nrm = emu.trace_rays(emu.Scene(desc), rays, accel=rt.ACCEL_BVH if accel == "bvh" else rt.ACCEL_LINEAR)
    rid, rt_, rn = orc.OracleScene(desc).trace_rays(rays)
    assert np.array_equal(ids, rid) and np.array_equal(t, rt_) and np.array_equal(nrm, rn)
    assert (rid[:n] > 0).mean() > 0.2               # a good share of them does hit spheres
    r2 = np.where(rng.uniform(size=len(rays)) < 0.5, -1.0, (dist.max() * 2) ** 2)
    occ = emu.shadow_rays(emu.Scene(desc), rays, r2, accel=rt.ACCEL_BVH if accel == "bvh" else rt.ACCEL_LINEAR)
    assert np.array_equal(occ, rid >= 0)


@pytest.mark.parametrize("nodes", ["q16", "f32"])
def test_both_node_formats_equal_the_scan(rt, orc, scenes, nodes, monkeypatch):
    """32-byte grid nodes (boxes rounded outward to a 16-bit grid, padded 1.25 cells) and 64-byte f32 nodes are both only
    a conservative filter in front of the scan's exact test: ids, t and normals bit-identical with either"""
    tune = rt.TUNE_NODES_Q16 if nodes == "q16" else rt.TUNE_NODES_F32
    desc = scenes.random_spheres(3000, 31, 40.0, 8, 8)
    assert emu.bvh_node_bytes(emu.Scene(desc), tune) == (32 if nodes == "q16" else 64)
    rng = np.random.default_rng(12)
    rays = _rays(rng, 40000, 45.0, 8.0)
    rays[:2000, 3:] = np.array([1.0, 0.0, 0.0]) * rng.choice([-1.0, 1.0], 2000)[:, None]   # axis-parallel: clamped reciprocals
    rays[2000:4000, 4] = 0.0
    rays[4000:5000, 3] = 1e-35
    ids, t, nrm = emu.trace_rays(emu.Scene(desc), rays, accel=rt.ACCEL_BVH, tuning=tune)
    rid, rt_, rn = orc.OracleScene(desc).trace_rays(rays)
    assert np.array_equal(ids, rid) and np.array_equal(t, rt_) and np.array_equal(nrm, rn)
    r2 = np.where(rng.uniform(size=len(rays)) < 0.3, -1.0, rng.uniform(0.0, 60.0, len(rays)) ** 2)
    for shadow_tune in (tune, tune | rt.TUNE_BINARY_SHADOW):   # four-wide walk (grid nodes only) and binary walk
        occ = emu.shadow_rays(emu.Scene(desc), rays, r2, accel=rt.ACCEL_BVH, tuning=shadow_tune)
        with np.errstate(invalid="ignore"):
            assert np.array_equal(occ, (rid >= 0) & ((r2 < 0) | (rt_ * rt_ < r2)))


def test_node_format_follows_the_grid_resolution(rt, scenes):
    assert emu.bvh_node_bytes(emu.Scene(scenes.random_spheres(500, 3, 60.0, 8, 8))) == 32       # cell 0.002 vs radius 0.5
    assert emu.bvh_node_bytes(emu.Scene(scenes.random_spheres(500, 3, 30000.0, 8, 8))) == 64    # cell 0.9: too coarse


def test_coarse_grid_is_still_conservative(rt, orc, scenes, monkeypatch):
    """forcing the grid on a scene it is far too coarse for (cells larger than the spheres) may only cost time"""
    desc = scenes.random_spheres(800, 9, 30000.0, 8, 8)
    assert emu.bvh_node_bytes(emu.Scene(desc), rt.TUNE_NODES_Q16) == 32
    data = desc["objects"][1]["data"]
    rng = np.random.default_rng(2)
    k = rng.integers(0, len(data), 6000)
    o = np.stack([rng.uniform(-30000, 30000, 6000), rng.uniform(1, 50, 6000), rng.uniform(-30000, 30000, 6000)], 1)
    tgt = data[k, :3] + rng.normal(size=(6000, 3)) * data[k, 3:4] * 0.7
    rays = np.concatenate([o, tgt - o], 1)
    ids, t, nrm = emu.trace_rays(emu.Scene(desc), rays, accel=rt.ACCEL_BVH)
    rid, rt_, rn = orc.OracleScene(desc).trace_rays(rays)
    assert np.array_equal(ids, rid) and np.array_equal(t, rt_) and np.array_equal(nrm, rn)
    assert (rid > 0).mean() > 0.3


@pytest.mark.parametrize("n", [1, 2, 3])
def test_tiny_trees(rt, orc, scenes, n):
    """one, two, three spheres forced through the BVH (single-leaf root, one internal node, ...) in both node formats"""
    desc = scenes.random_spheres(n, 40 + n, 3.0, 8, 8)
    rng = np.random.default_rng(n)
    rays = _rays(rng, 4000, 5.0, 6.0)
    data = desc["objects"][1]["data"]
    aim = data[rng.integers(0, n, 2000), :3] + rng.normal(size=(2000, 3)) * 0.3
    rays[:2000, 3:] = aim - rays[:2000, :3]
    rid, rt_, rn = orc.OracleScene(desc).trace_rays(rays)
    assert (rid > 0).mean() > 0.05
    ids, t, nrm = emu.trace_rays(emu.Scene(desc), rays, accel=rt.ACCEL_BVH)
    assert np.array_equal(ids, rid) and np.array_equal(t, rt_) and np.array_equal(nrm, rn)
    assert emu.bvh_node_bytes(emu.Scene(desc)) == 32


@pytest.mark.parametrize("case", range(8))
def test_filter_is_conservative_on_odd_scenes(rt, orc, case):
    """CPU twin of tests/test_gpu_parity.py::test_bvh_filter_is_conservative_on_odd_scenes (same scene recipes, fewer rays):
    the header's BVH path against the oracle's scan"""
    from test_gpu_parity import _stress_scene
    rng = np.random.default_rng(100 + case)
    centre, extent, rlo, rhi = [
        ((0.0, 0.0, 0.0), (50.0, 50.0, 50.0), 0.05, 5.0),
        ((4000.0, -2500.0, 9000.0), (40.0, 5.0, 40.0), 0.1, 1.0),
        ((0.0, 0.0, 0.0), (2000.0, 0.5, 2000.0), 0.2, 0.8),
        ((0.0, 0.0, 0.0), (0.2, 300.0, 0.2), 0.01, 0.3),
        ((10.0, 10.0, 10.0), (30.0, 30.0, 30.0), 1e-3, 30.0),
        ((-1e5, 2e5, 5e4), (500.0, 500.0, 500.0), 0.5, 5.0),
        ((0.0, 0.0, 0.0), (1e-3, 1e-3, 1e-3), 1e-5, 1e-4),
        ((0.0, 0.0, 0.0), (3e4, 10.0, 3e4), 0.3, 0.6),
    ][case]
    centre, extent = np.array(centre), np.array(extent)
    desc, c, r = _stress_scene(rt.scene, rng, 1500, centre, extent, rlo, rhi)
    n = 8000
    k = rng.integers(0, len(c), n)
    target = c[k] + rng.normal(size=(n, 3)) * r[k][:, None] * rng.choice([0.3, 1.0, 1.5], n)[:, None]
    dist = np.exp(rng.uniform(np.log(np.linalg.norm(extent) * 1e-3), np.log(np.linalg.norm(extent) * 1e3), n))
    dirs = rng.normal(size=(n, 3)); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    o = target - dirs * dist[:, None]
    d = dirs * np.exp(rng.uniform(-3, 3, n))[:, None]
    d[:300, 0] = 0.0; d[300:600, 1] = 0.0; d[600:800, [0, 2]] = 0.0
    o[800:1000] = c[k[800:1000]]
    rays = np.concatenate([o, d], axis=1)
    ids, t, nrm = emu.trace_rays(emu.Scene(desc), rays, accel=rt.ACCEL_BVH)
    rid, rt_, rn = orc.OracleScene(desc).trace_rays(rays)
    assert np.array_equal(ids, rid), "%d ids differ" % int((ids != rid).sum())
    assert np.array_equal(t, rt_, equal_nan=True) and np.array_equal(nrm, rn, equal_nan=True)
    assert emu.bvh_node_bytes(emu.Scene(desc)) == (64 if case in (2, 7) else 32)   # 4000- and 60000-wide fields: grid too coarse


def test_flatten_cache_follows_scene_edits(rt, scenes):
    """rt_context_create flattens a scene once and keeps the result with the rt_scene (host_scene.h FlatCache); every
    builder call that changes objects, materials or skybox faces must drop it (it may be page-locked in place)."""
    import ctypes as C
    B = emu.B
    L = emu.lib()
    sc = emu.Scene(scenes.random_spheres(50, 3, 10.0, 32, 24))

    def probe(tuning=0):
        out = (C.c_uint64 * 6)()
        assert L.emu_flat_cache_probe(sc.handle, tuning, out) == 0
        return list(out)

    a, b = probe(), probe()
    assert a == b and a[2:5] == [50, 1, 51]          # same entry, 50 spheres + the ground plane
    ball = np.array([[0.0, 1.0, 0.0, 1.0]])
    mat = np.array([0], dtype=np.uint32)
    assert L.rt_scene_add_spheres(sc.handle, 1, B.ptr(ball, C.c_double), B.ptr(mat, C.c_uint32)) == 0
    c = probe()
    assert c[1] > a[1] and c[2:5] == [51, 1, 52]
    rgb = (C.c_double * 3)(0.5, 0.5, 0.5)
    assert L.rt_scene_add_material_indirect(sc.handle, rgb, rgb, 1.0, rgb, 5) == 51  # max_children follows materials
    d = probe()
    assert d[1] > c[1] and d[5] == 5 and c[5] == 2
    assert L.rt_scene_add_plane(sc.handle, (C.c_double * 3)(0, 9, 0), (C.c_double * 3)(0, -1, 0), 0) == 0
    e = probe()
    assert e[1] > d[1] and e[2:5] == [51, 2, 53]
    # the node-format override is part of the key (A/B runs switch it between contexts of one scene)
    f = probe(rt.TUNE_NODES_F32)
    assert f[1] == e[1] and f[0] != e[0] and probe(rt.TUNE_NODES_F32) == f
    assert probe(rt.TUNE_NODES_F32 | rt.TUNE_MORTON_AXIS) == f      # only the node bits decide the flattened form
    assert probe()[1] == e[1]


# ---------------------------------------------------------------------------------------------------------------------
# four-wide nodes for shadow queries (default where the grid nodes are in use; rt_types.h BvhWNode, rt_device.cuh
# trav_descend_wide); RT_TUNE_BINARY_SHADOW selects the binary walk they are derived from
# ---------------------------------------------------------------------------------------------------------------------
def _wide_scenes(rt, scenes):
    S = rt.scene
    out = [("random800", scenes.random_spheres(800, 5, 12.0, 8, 8))]
    for n in (1, 2, 3, 5, 65):
        out.append(("tiny%d" % n, scenes.random_spheres(n, 40 + n, 3.0, 8, 8)))
    b = S.SceneBuilder()
    m = b.add_material(S.PhongMaterial((0.5, 0.5, 0.5), (0, 0, 0), 1.0, (0, 0, 0)))
    b.add_object(S.Plane((0, -1, 0), (0, 1, 0)), m)
    rng = np.random.default_rng(11)
    c = rng.uniform(-5, 5, size=(60, 3))
    r = rng.uniform(0.05, 1.5, size=60)
    data = np.concatenate([np.c_[c, r], np.c_[c, r], np.c_[c[:20], r[:20] * 0.5], [[0, 0, 0, 40.0], [0, 0, 0, 1e-3], [30.0, 0, 0, 1.0]]])
    b.add_spheres(data, np.zeros(len(data), dtype=np.uint32))
    b.camera = S.SimplePerspectiveCamera.new((0, 0, 20), (0, 0, -1), (0, 1, 0), 2.0)
    b.background = S.SolidColorBackground((0, 0, 0))
    b.options = S.Options(8, 8, 1)
    out.append(("degenerate", b.build()))
    return out


def test_wide_shadow_walk_equals_scan_and_binary_walk(rt, orc, scenes):
    """the four-wide shadow walk answers every query like the linear scan (and like the binary walk it is derived from)"""
    rng = np.random.default_rng(7)
    for name, desc in _wide_scenes(rt, scenes):
        rays = _rays(rng, 6000, 14.0, 8.0)
        r2 = np.where(rng.uniform(size=6000) < 0.3, -1.0, rng.uniform(0.0, 30.0, 6000) ** 2)
        scan = emu.shadow_rays(emu.Scene(desc), rays, r2, accel=rt.ACCEL_LINEAR)
        binary = emu.shadow_rays(emu.Scene(desc), rays, r2, accel=rt.ACCEL_BVH, tuning=rt.TUNE_NODES_Q16 | rt.TUNE_BINARY_SHADOW)
        wide = emu.shadow_rays(emu.Scene(desc), rays, r2, accel=rt.ACCEL_BVH, tuning=rt.TUNE_NODES_Q16)
        assert np.array_equal(binary, scan), name
        assert np.array_equal(wide, scan), name


def test_wide_shadow_walk_renders_the_oracle_frame(rt, orc, scenes):
    desc = scenes.random_spheres(1500, 0x5EED0003, 25.0, 96, 54)
    ref = orc.OracleScene(desc).render(max_depth=6, seed=3)
    assert emu.bvh_node_bytes(emu.Scene(desc)) == 32      # grid nodes, hence wide shadow nodes, by default
    got = emu.render(emu.Scene(desc), max_depth=6, seed=3, accel=rt.ACCEL_BVH)
    check(got, ref, "spheres1500 wide")
    assert ref["stats"]["shadow_rays"] > 96 * 54


def test_wide_nodes_halve_the_dependent_fetches(rt, scenes):
    """what the wide nodes are for: about half as many node visits per shadow query (each one fetch of 64 bytes)"""
    desc = scenes.random_spheres(1500, 0x5EED0003, 25.0, 96, 54)
    a = emu.render(emu.Scene(desc), max_depth=3, seed=3, accel=rt.ACCEL_BVH, tuning=rt.TUNE_BINARY_SHADOW)
    b = emu.render(emu.Scene(desc), max_depth=3, seed=3, accel=rt.ACCEL_BVH)
    va, vb = a["stats"]["bvh_nodes_visited"]["shadow"], b["stats"]["bvh_nodes_visited"]["shadow"]
    assert a["stats"]["bvh_nodes_visited"]["closest"] == b["stats"]["bvh_nodes_visited"]["closest"] > 0
    assert np.array_equal(a["bgr"], b["bgr"])
    assert 0.35 * va < vb < 0.7 * va, (va, vb)


def test_any_tree_over_the_same_spheres_gives_the_scan_answer(rt, orc, scenes, monkeypatch):
    """BVH == linear scan does not depend on WHICH tree: the walks of rt_device.cuh over a full-sweep SAH tree (the emulation's
    experimental builder, EMU_BVH_BUILDER=sah) return the frame the LBVH and the oracle's scan return, bit for bit -- and the
    better tree saves only a few per cent of the node visits on these uniformly filled scenes (tree
    quality is not where the traversal kernels' time is)."""
    desc = scenes.random_spheres(1500, 0x5EED0004, 25.0, 96, 54)
    ref = orc.OracleScene(desc).render(max_depth=6, seed=3)
    lbvh = emu.render(emu.Scene(desc), max_depth=6, seed=3, accel=rt.ACCEL_BVH)
    monkeypatch.setenv("EMU_BVH_BUILDER", "sah")
    sah = emu.render(emu.Scene(desc), max_depth=6, seed=3, accel=rt.ACCEL_BVH)
    check(sah, ref, "spheres1500 sah tree")
    assert np.array_equal(sah["bgr"], lbvh["bgr"]) and np.array_equal(sah["ids"], lbvh["ids"])
    for cls in ("closest", "shadow"):
        a, b = lbvh["stats"]["bvh_nodes_visited"][cls], sah["stats"]["bvh_nodes_visited"][cls]
        assert 0.75 * a < b < 1.02 * a, (cls, a, b)
    rays = np.random.default_rng(3).normal(size=(5000, 6)) * np.array([20, 3, 20, 1, 1, 1.0]) + np.array([0, 4, 0, 0, 0, 0.0])
    ids, t, _ = emu.trace_rays(emu.Scene(desc), rays, accel=rt.ACCEL_BVH)
    rid, rt_, _ = orc.OracleScene(desc).trace_rays(rays)
    assert np.array_equal(ids, rid) and np.array_equal(t, rt_)
