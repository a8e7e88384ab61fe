"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star): primary hit ids bit-exact; every byte of the quantised frame within +-1 LSB and
>= 99.9 % identical; ray counts (Scene::intersect calls) equal.  In RT_FP_STRICT mode geometry is the same
unfused f64 arithmetic as the oracle, so hit distances are required to be bit-identical too.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

import importlib

REF_TEXT_SCENE = importlib.import_module("rust-raytrace_b200.scenes").TEST_SCENE_TEXT   # upstream test_scene.txt, values restated


def assert_parity(frame, ref, what, exact=False):
    assert np.array_equal(frame.ids, ref["ids"]), "%s: primary hit ids differ in %d pixels" % (what, int((frame.ids != ref["ids"]).sum()))
    a, b = frame.rows.astype(np.int16), ref["bgr"].astype(np.int16)
    assert a.shape == b.shape
    diff = np.abs(a - b)
    assert diff.max() <= 1, "%s: a byte differs by %d LSB" % (what, int(diff.max()))
    same = float((diff == 0).mean())
    assert same >= 0.999, "%s: only %.5f of the bytes identical" % (what, same)
    if exact:
        assert same == 1.0, "%s: %d bytes differ" % (what, int((diff != 0).sum()))
    assert frame.stats["rays_total"] == ref["stats"]["rays"], "%s: ray count %d vs oracle %d" % (what, frame.stats["rays_total"], ref["stats"]["rays"])
    assert frame.stats["rays_shadow"] == ref["stats"]["shadow_rays"]
    return same


def test_quantiser(gpu, orc):
    v, a = orc.srgb_tables()
    rng = np.random.default_rng(1)
    vals = np.concatenate([a, np.nextafter(a, -np.inf), np.nextafter(a, np.inf), v, rng.uniform(-0.1, 1.2, 50000),
                           [0.0, -1.0, 0.051, 0.18, 0.5, 1.0, 5.0, np.nan, np.inf, -np.inf, -0.0]])
    vals = np.resize(vals, (vals.size // 3) * 3).reshape(-1, 3)
    assert np.array_equal(gpu.color.write_bgr(vals), orc.quantize_bgr(vals))


def _random_rays(rng, n, extent, ymax):
    o = np.stack([rng.uniform(-extent, extent, n), rng.uniform(0.1, ymax, n), rng.uniform(-extent, extent, n)], axis=1)
    d = rng.normal(size=(n, 3))
    d[: n // 2] /= np.linalg.norm(d[: n // 2], axis=1, keepdims=True)  # half unit, half un-normalised directions
    return np.concatenate([o, d], axis=1)


@pytest.mark.parametrize("accel", ["linear", "bvh"])
@pytest.mark.parametrize("fp_mode", ["strict", "fast"])
def test_intersect_random_rays(gpu, orc, scenes, accel, fp_mode):
    desc = scenes.random_spheres(3000, 77, 30.0, 64, 64)
    sc = gpu.Scene(desc)
    rays = _random_rays(np.random.default_rng(5), 200_000, 35.0, 9.0)
    # rays leaving sphere surfaces (secondary-ray geometry) and rays from inside spheres
    data = desc["objects"][1]["data"]
    inside = np.concatenate([data[:2000, :3], np.random.default_rng(6).normal(size=(2000, 3))], axis=1)
    rays = np.concatenate([rays, inside])
    ids, t, nrm = gpu.raytrace.intersect(sc, rays, accel=gpu.ACCEL_BVH if accel == "bvh" else gpu.ACCEL_LINEAR,
                                         fp_mode=gpu.FP_FAST if fp_mode == "fast" else gpu.FP_STRICT)
    rid, rt_, rn = orc.OracleScene(desc).trace_rays(rays)
    if fp_mode == "strict":
        assert np.array_equal(ids, rid)
        assert np.array_equal(t, rt_), "hit distances must be bit-identical in strict mode"
        assert np.array_equal(nrm, rn)
    else:
        assert (ids != rid).mean() < 1e-5
        ok = ids == rid
        assert np.allclose(t[ok], rt_[ok], rtol=1e-9, atol=0)


@pytest.mark.parametrize("accel", ["linear", "bvh"])
def test_intersect_far_origins(gpu, orc, scenes, accel):
    """origins 1e3 .. 1e9 units out (horizon hits of the ground plane): the BVH advances the ray to the sphere bounds
    before the f32 filter and must still return the scan's result bit for bit"""
    desc = scenes.random_spheres(3000, 21, 30.0, 8, 8)
    rng = np.random.default_rng(4)
    n = 100_000
    target = np.stack([rng.uniform(-30, 30, n), rng.uniform(0.0, 6.0, n), rng.uniform(-30, 30, n)], 1)
    dirs = rng.normal(size=(n, 3))
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    dist = 10.0 ** rng.uniform(3, 9, n)
    o = target - dirs * dist[:, None]
    o[:, 1] = np.abs(o[:, 1])
    d = target - o
    d[: n // 2] /= np.linalg.norm(d[: n // 2], axis=1, keepdims=True)
    rays = np.concatenate([np.concatenate([o, d], 1), np.concatenate([o[:5000], -d[:5000]], 1)])
    ids, t, nrm = gpu.raytrace.intersect(gpu.Scene(desc), rays, accel=gpu.ACCEL_BVH if accel == "bvh" else gpu.ACCEL_LINEAR)
    rid, rt_, rn = orc.OracleScene(desc).trace_rays(rays)
    assert np.array_equal(ids, rid) and np.array_equal(t, rt_) and np.array_equal(nrm, rn)
    assert (rid[:n] > 0).mean() > 0.2
    if accel == "bvh":
        # a handful of rays from ~1e8 out among ordinary ones: their walks run away (the filter must pad by the exact test's own
        # rounding error, ~10 units here) and, being few, each is scanned by many CTAs in slices and merged (k_heavy_scan)
        far = dist > 3e7
        few = np.concatenate([rays[:n][far][:40], _random_rays(rng, 5000, 35.0, 9.0)])
        ids2, t2, nrm2 = gpu.raytrace.intersect(gpu.Scene(desc), few, accel=gpu.ACCEL_BVH)
        rid2, rt2, rn2 = orc.OracleScene(desc).trace_rays(few)
        assert np.array_equal(ids2, rid2) and np.array_equal(t2, rt2) and np.array_equal(nrm2, rn2)
        assert (rid2[:40] > 0).sum() >= 5


@pytest.mark.parametrize("nodes,morton", [("q16", "cubic"), ("f32", "cubic"), ("q16", "axis")])
def test_bvh_node_formats_and_morton_grids(gpu, orc, scenes, nodes, morton):
    """the 32-byte grid nodes, the 64-byte f32 nodes and either Morton scaling are only different conservative filters /
    different trees in front of the same exact test: ids, t, normals bit-identical, a whole frame byte-identical in ids"""
    tune = (gpu.TUNE_NODES_Q16 if nodes == "q16" else gpu.TUNE_NODES_F32) | (gpu.TUNE_MORTON_AXIS if morton == "axis" else 0)
    desc = scenes.random_spheres(8000, 13, 50.0, 320, 200)
    sc = gpu.Scene(desc)
    rng = np.random.default_rng(9)
    rays = _random_rays(rng, 100_000, 55.0, 9.0)
    rays[:3000, 3:] = np.array([0.0, 0.0, 1.0]) * rng.choice([-1.0, 1.0], 3000)[:, None]   # axis-parallel: clamped reciprocals
    rays[3000:6000, 4] = 0.0
    rays[6000:7000, 3] = 1e-35
    ids, t, nrm = gpu.raytrace.intersect(sc, rays, accel=gpu.ACCEL_BVH, tuning=tune)
    rid, rt_, rn = orc.OracleScene(desc).trace_rays(rays)
    assert np.array_equal(ids, rid) and np.array_equal(t, rt_) and np.array_equal(nrm, rn)
    bvh = gpu.raytrace.render(sc, max_depth=4, seed=2, accel=gpu.ACCEL_BVH, tuning=tune)
    assert bvh.stats["bvh_node_bytes"] == (32 if nodes == "q16" else 64)
    lin = gpu.raytrace.render(sc, max_depth=4, seed=2, accel=gpu.ACCEL_LINEAR)
    assert np.array_equal(bvh.ids, lin.ids)
    assert bvh.stats["rays_total"] == lin.stats["rays_total"] and bvh.stats["rays_shadow"] == lin.stats["rays_shadow"]
    diff = np.abs(bvh.bgr.astype(np.int16) - lin.bgr.astype(np.int16))
    assert diff.max() <= 1 and (diff == 0).mean() >= 0.999


def test_intersect_edge_cases(gpu, orc):
    S = gpu.scene
    b = S.SceneBuilder()
    m = b.add_material(S.PhongMaterial((0.5, 0.5, 0.5), (0, 0, 0), 1.0, (0, 0, 0)))
    b.add_object(S.Plane((0, 0, -3), (0, 0, 1)), m)       # 0
    b.add_object(S.Sphere((0, 1.5, 0), 1.5), m)           # 1
    b.add_object(S.Sphere((0, 1.5, 0), 1.5), m)           # 2 coincident with 1 -> lower index wins
    b.add_object(S.Plane((0, -1, 0), (0, 1, 0)), m)       # 3
    b.add_object(S.Sphere((5, 5, 5), 1.0), m)             # 4
    b.camera = S.SimplePerspectiveCamera.new((0, 3, 17), (0, 0, -1), (0, 1, 0), 3.6)
    b.background = S.SolidColorBackground((0, 0, 0))
    b.options = S.Options(8, 8, 1)
    desc = b.build()
    rays = np.array([
        [0, 3, 17, 0, 0, -1],        # exactly tangent to sphere 1 (disc == 0 -> miss), hits plane 0 at t = 20
        [0, 1.5, 17, 0, 0, -1],      # through the centre of spheres 1/2 -> id 1
        [0, 1.5, 0, 0, 0, 1],        # from inside the spheres -> far root
        [0, 1.5, 0, 0, 0, -2.5],     # inside, un-normalised direction
        [0, 5, 5, 1, 0, 0],          # parallel to plane 3 with positive numerator?  n.(p-o) = -6 -> -inf -> rejected
        [0, -5, 5, 1, 0, 0],         # parallel, numerator +4 -> t = +inf accepted unless something closer (sphere 4? no)
        [0, -1, 5, 1, 0, 0],         # in plane 3: 0/0 = NaN -> plane 3 wins whatever else is hit
        [5, -1, 5, 0, 0.0, 0.0],     # zero direction
        [5, 20, 5, 0, -1, 0],        # straight down through sphere 4 then plane 3
        [np.nan, 0, 0, 0, 0, -1],
        [0, 1, 30, 0, 0, -1e-3],     # tiny direction length
    ], dtype=np.float64)
    ref = orc.OracleScene(desc).trace_rays(rays)
    for accel in (gpu.ACCEL_LINEAR, gpu.ACCEL_BVH):
        ids, t, nrm = gpu.raytrace.intersect(gpu.Scene(desc), rays, accel=accel)
        assert np.array_equal(ids, ref[0]), (accel, ids, ref[0])
        assert np.array_equal(t, ref[1], equal_nan=True), (accel, t, ref[1])
        assert np.array_equal(nrm, ref[2], equal_nan=True)
    assert ref[0][0] == 0 and ref[1][0] == 20.0 and ref[0][1] == 1 and ref[0][6] == 3 and np.isnan(ref[1][6])


def test_project(gpu, orc):
    S = gpu.scene
    rng = np.random.default_rng(3)
    pos = rng.uniform(-1.5, 1.5, size=(5000, 2))
    cam = S.SimplePerspectiveCamera.look_at((0.3, 1.0, -2.0), (0.2, -0.3, -1.0), (0.0, 1.0, 0.1), np.deg2rad(40.0), 2.5)
    for dof in (None, dict(focus=7.5, aperture=0.2, samples=3)):
        b = S.SceneBuilder()
        b.camera = dict(cam, dof=dof) if dof else cam
        b.background = S.SolidColorBackground((0, 0, 0))
        b.options = S.Options(4, 4, 1)
        desc = b.build()
        got = gpu.raytrace.project(gpu.Scene(desc), pos, seed=9)
        ref = orc.OracleScene(desc).project(pos, seed=9)
        if dof is None:
            assert np.array_equal(got, ref)
        else:
            assert np.allclose(got, ref, rtol=1e-12, atol=1e-13)  # sin/cos: CUDA libm vs glibc, <= 2 ulp


def test_render_cornell_text(gpu, orc):
    from oracle import scene_text
    desc = scene_text.deserialize(REF_TEXT_SCENE)
    sc = gpu.serialize.deserialize(REF_TEXT_SCENE)
    assert sc.options() == (800, 800, 1024)
    kw = dict(width=240, height=200, antialias=3, seed=21)
    ref = orc.OracleScene(desc).render(**kw)
    fr = gpu.raytrace.render(sc, **kw)
    assert_parity(fr, ref, "cornell")
    assert fr.stats["accel_used"] == gpu.ACCEL_LINEAR


@pytest.mark.parametrize("accel", ["linear", "bvh"])
@pytest.mark.parametrize("fp_mode", ["strict", "fast"])
def test_render_all_materials(gpu, orc, scenes, accel, fp_mode):
    desc = scenes.materials_showcase(256, 192, 2)
    ref = orc.OracleScene(desc).render(max_depth=5, seed=4)
    fr = gpu.raytrace.render(gpu.Scene(desc), max_depth=5, seed=4, accel=gpu.ACCEL_BVH if accel == "bvh" else gpu.ACCEL_LINEAR,
                             fp_mode=gpu.FP_FAST if fp_mode == "fast" else gpu.FP_STRICT)
    if fp_mode == "strict":
        assert_parity(fr, ref, "showcase/%s/%s" % (accel, fp_mode))
    else:  # contraction can move a near-threshold significance test: ray counts may differ by a hair
        assert np.array_equal(fr.ids, ref["ids"])
        diff = np.abs(fr.rows.astype(np.int16) - ref["bgr"].astype(np.int16))
        assert diff.max() <= 1 and (diff == 0).mean() >= 0.999
        assert abs(fr.stats["rays_total"] - ref["stats"]["rays"]) <= 1e-4 * ref["stats"]["rays"]


def test_render_spheres_bvh_depth8_text_and_builder(gpu, orc, scenes):
    desc = scenes.random_spheres(4000, 0x5EED0003, 40.0, 320, 180)
    ref = orc.OracleScene(desc).render(max_depth=8, seed=3)
    fr_b = gpu.raytrace.render(gpu.Scene(desc), max_depth=8, seed=3, accel=gpu.ACCEL_BVH)
    assert_parity(fr_b, ref, "spheres4000/builder")
    fr_t = gpu.raytrace.render(gpu.serialize.deserialize(gpu.scene.to_text(desc)), max_depth=8, seed=3, accel=gpu.ACCEL_BVH)
    assert np.array_equal(fr_t.rows, fr_b.rows) and np.array_equal(fr_t.ids, fr_b.ids), "text and builder scenes must render identically"
    fr_l = gpu.raytrace.render(gpu.Scene(desc), max_depth=8, seed=3, accel=gpu.ACCEL_LINEAR)
    assert np.array_equal(fr_l.ids, fr_b.ids)
    assert fr_l.stats["rays_total"] == fr_b.stats["rays_total"]
    assert (fr_l.rows != fr_b.rows).mean() < 1e-5  # same rays, only the accumulation order may differ


def test_render_skybox_dof_area(gpu, orc, scenes):
    S = gpu.scene
    desc = scenes.random_spheres(200, 5, 10.0, 200, 144, antialias=2, reflective_only=True,
                                 background=S.SkyboxBackground(*scenes.skybox_faces(64, 8)),
                                 camera=S.DepthOfFieldCamera.new(S.SimplePerspectiveCamera.new((0, 4, 14), (0, -0.2, -1), (0, 1, 0), 2.0), 12.0, 0.15, 2))
    desc["lights"].append(S.AreaLight((-3.0, 9.0, -3.0), (6.0, 0.0, 0.0), (0.0, 0.0, 6.0), (0.6, 0.6, 0.5)))
    ref = orc.OracleScene(desc).render(max_depth=4, seed=8)
    fr = gpu.raytrace.render(gpu.Scene(desc), max_depth=4, seed=8)
    # depth-of-field rays go through sin/cos (libm differs by <= 2 ulp), so ids may differ on a handful of edge pixels
    assert (fr.ids != ref["ids"]).mean() < 1e-4
    diff = np.abs(fr.rows.astype(np.int16) - ref["bgr"].astype(np.int16))
    assert (diff <= 1).mean() >= 0.9999 and (diff == 0).mean() >= 0.999


def test_queue_slicing_is_invisible(gpu, scenes):
    desc = scenes.materials_showcase(160, 120, 2)
    sc = gpu.Scene(desc)
    big = gpu.raytrace.render(sc, max_depth=6, seed=2)
    small = gpu.raytrace.render(sc, max_depth=6, seed=2, max_rays_in_flight=4096)
    assert np.array_equal(big.ids, small.ids)
    assert big.stats["rays_total"] == small.stats["rays_total"]
    assert (big.rows != small.rows).mean() < 1e-5
    assert small.stats["kernel_launches"] > big.stats["kernel_launches"]


def test_device_driven_and_host_driven_frames_agree(gpu, orc, scenes):
    """The default frame queues every launch without reading a count back (context.cu frame_device_driven); RT_TUNE_HOST_LEVELS
    reads every level's count on the host.  Same rays, same frame.  A queue that overflows in the device-driven frame makes
    the context redo it level by level (sliced): results unchanged, and the stats say so."""
    desc = scenes.random_spheres(400, 21, 12.0, 160, 120, antialias=2)   # 10 % transparent spheres: levels branch a little
    sc = gpu.Scene(desc)
    ref = orc.OracleScene(desc).render(max_depth=6, seed=2)
    dd = gpu.raytrace.render(sc, max_depth=6, seed=2)
    hl = gpu.raytrace.render(sc, max_depth=6, seed=2, tuning=gpu.TUNE_HOST_LEVELS)
    # the default frame also traces each level's shadow queries on a second stream beside the next level; RT_TUNE_SERIAL_SHADOW
    # keeps them for one launch at the end of the frame
    ss = gpu.raytrace.render(sc, max_depth=6, seed=2, tuning=gpu.TUNE_SERIAL_SHADOW)
    assert dd.stats["host_levels"] == 0 and dd.stats["device_driven_overflows"] == 0 and hl.stats["host_levels"] == 1
    assert ss.stats["host_levels"] == 0 and ss.stats["kernel_launches"] < dd.stats["kernel_launches"]
    assert (dd.rows != ss.rows).mean() < 1e-5
    for fr in (dd, hl, ss):
        assert np.array_equal(fr.ids, ref["ids"])
        assert fr.stats["rays_total"] == ref["stats"]["rays"]
        for k in ("rays_primary", "rays_secondary", "rays_shadow", "paths"):
            assert fr.stats[k] == hl.stats[k]
    assert (dd.rows != hl.rows).mean() < 1e-5
    # between two facing IndirectPhong planes every hit spawns four rays that hit again: the levels grow fourfold, outgrow any
    # queue sized for the paths, the optimistic frame overflows and is redone level by level (sliced)
    S = gpu.scene
    b = S.SceneBuilder()
    m = b.add_material(S.IndirectPhongMaterial((0.5, 0.45, 0.4), (0.0, 0.0, 0.0), 1.0, (0.05, 0.05, 0.05), 4))
    b.add_object(S.Plane((0, 0, 0), (0, 1, 0)), m)
    b.add_object(S.Plane((0, 10, 0), (0, -1, 0)), m)
    b.add_object(S.Sphere((0, 3, -8), 2.0), b.add_material(S.TransparentMaterial((0.9, 0.9, 0.9), 100.0, 1.5)))
    b.add_light(S.PointLight((0.0, 5.0, 0.0), (0.7, 0.7, 0.7)))
    b.camera = S.SimplePerspectiveCamera.new((0, 5, 10), (0, -0.1, -1), (0, 1, 0), 2.0)
    b.background = S.SolidColorBackground((0.1, 0.1, 0.2))
    b.options = S.Options(64, 48, 1)
    desc = b.build()
    sc = gpu.Scene(desc)
    ref = orc.OracleScene(desc).render(max_depth=4, seed=2)
    whole = gpu.raytrace.render(sc, max_depth=4, seed=2, tuning=gpu.TUNE_HOST_LEVELS)
    ctx = gpu.Context(sc, max_depth=4, seed=2, max_rays_in_flight=2048)   # several primary batches, levels that outgrow the queue
    first = ctx.render()
    again = ctx.render()
    assert first["device_driven_overflows"] == 1 and first["host_levels"] == 1
    assert again["device_driven_overflows"] == 1 and again["host_levels"] == 1       # stays host-driven: no second wasted attempt
    assert first["rays_total"] == again["rays_total"] == ref["stats"]["rays"] == whole.stats["rays_total"]
    rows, ids, _ = ctx.download(want_ids=True)
    assert np.array_equal(ids[:48], ref["ids"]) and (rows[:48] != whole.rows).mean() < 1e-5
    ctx.close()
    # the scene remembers: the next context (rt_render builds one per frame) does not pay for the failed attempt again
    later = gpu.raytrace.render(sc, max_depth=4, seed=2, max_rays_in_flight=2048)
    assert later.stats["host_levels"] == 1 and later.stats["device_driven_overflows"] == 0
    assert later.stats["rays_total"] == ref["stats"]["rays"]
    # 16 samples per pixel in batches of 8192 paths, single-child materials: every batch device-driven, nothing overflows
    desc2 = scenes.random_spheres(300, 7, 12.0, 64, 48, antialias=16, transparent=False)
    ref2 = orc.OracleScene(desc2).render(max_depth=4, seed=5)
    fr2 = gpu.raytrace.render(gpu.Scene(desc2), max_depth=4, seed=5, max_rays_in_flight=8192)
    assert fr2.stats["host_levels"] == 0 and fr2.stats["rays_total"] == ref2["stats"]["rays"]
    assert np.array_equal(fr2.ids, ref2["ids"])
    fr3 = gpu.raytrace.render(gpu.Scene(desc2), max_depth=4, seed=5, max_rays_in_flight=8192, tuning=gpu.TUNE_SERIAL_SHADOW)
    assert fr3.stats["rays_total"] == fr2.stats["rays_total"] and (fr3.rows != fr2.rows).mean() < 1e-5
    d = np.abs(fr2.rows.astype(np.int16) - ref2["bgr"].astype(np.int16))
    assert d.max() <= 1 and (d == 0).mean() >= 0.999


def test_device_scene_stays_with_the_scene(gpu, scenes):
    """rt_render builds a context per call; the device copy of the scene (buffers + BVH) is kept with the rt_scene per device
    and reused until the scene changes (host_scene.h DeviceSceneSlot): the second call reports no BVH build"""
    desc = scenes.random_spheres(5000, 3, 40.0, 96, 64)
    sc = gpu.Scene(desc)
    a = gpu.raytrace.render(sc, max_depth=3, seed=1, accel=gpu.ACCEL_BVH)
    b = gpu.raytrace.render(sc, max_depth=3, seed=1, accel=gpu.ACCEL_BVH)
    c = gpu.raytrace.render(sc, max_depth=3, seed=1, accel=gpu.ACCEL_BVH, tuning=gpu.TUNE_NODES_F32)   # another tree: built
    d = gpu.raytrace.render(sc, max_depth=3, seed=1, accel=gpu.ACCEL_LINEAR)
    assert a.stats["bvh_build_ms"] > 0 and b.stats["bvh_build_ms"] == 0 and c.stats["bvh_build_ms"] > 0 and d.stats["bvh_build_ms"] == 0
    assert c.stats["bvh_node_bytes"] == 64 and a.stats["bvh_node_bytes"] == 32
    for fr in (b, c, d):
        assert np.array_equal(fr.ids, a.ids) and (fr.rows != a.rows).mean() < 1e-5
    # camera and background are read per context (they do not drop the device copy)
    import ctypes as C
    L = gpu.lib()
    assert L.rt_scene_set_background_solid(sc.handle, (C.c_double * 3)(1.0, 0.0, 0.0)) == 0
    e = gpu.raytrace.render(sc, max_depth=3, seed=1, accel=gpu.ACCEL_BVH)
    assert e.stats["bvh_build_ms"] == 0 and (e.rows != a.rows).mean() > 0.01
    # a new light drops it
    assert L.rt_scene_add_light_point(sc.handle, (C.c_double * 3)(0.0, 50.0, 0.0), (C.c_double * 3)(1.0, 1.0, 1.0)) == 0
    f = gpu.raytrace.render(sc, max_depth=3, seed=1, accel=gpu.ACCEL_BVH)
    assert f.stats["bvh_build_ms"] > 0 and f.stats["rays_shadow"] > e.stats["rays_shadow"]


def test_scene_cache_follows_edits(gpu, scenes):
    """rt_context_create keeps the flattened, page-locked scene with the rt_scene between contexts (host_scene.h
    FlatCache); an edit of the scene must drop it.  60k spheres with their own materials = 8 MB: the page-locked path."""
    import copy
    import ctypes as C
    from importlib import import_module
    B = import_module("rust-raytrace_b200._binding")
    desc = scenes.random_spheres(60000, 21, 150.0, 160, 96)
    sc = gpu.Scene(desc)
    first = gpu.raytrace.render(sc, max_depth=3, seed=5)
    again = gpu.raytrace.render(sc, max_depth=3, seed=5)  # second context: cached buffers
    assert np.array_equal(first.rows, again.rows) and np.array_equal(first.ids, again.ids)
    assert first.stats["rays_total"] == again.stats["rays_total"]
    # a mirror ball in front of the camera, appended through the builder API to the live scene
    L = gpu.lib()
    ball = np.array([[0.0, 12.0, 150.0, 4.0]], dtype=np.float64)
    mat = np.array([1], dtype=np.uint32)
    B.check(L, L.rt_scene_add_spheres(sc.handle, 1, B.ptr(ball, C.c_double), B.ptr(mat, C.c_uint32)))
    edited = gpu.raytrace.render(sc, max_depth=3, seed=5)
    desc2 = copy.copy(desc)
    desc2["objects"] = list(desc["objects"]) + [{"shape": "sphere", "center": (0.0, 12.0, 150.0), "radius": 4.0, "material": 1}]
    fresh = gpu.raytrace.render(gpu.Scene(desc2), max_depth=3, seed=5)
    assert np.array_equal(edited.rows, fresh.rows) and np.array_equal(edited.ids, fresh.ids)
    assert (edited.ids == 60001).any() and not (first.ids == 60001).any()
    # a new material reallocates the page-locked material vector
    rgb = (C.c_double * 3)(0.5, 0.5, 0.5)
    assert L.rt_scene_add_material_phong(sc.handle, rgb, rgb, 10.0, rgb) == len(desc["materials"])
    after = gpu.raytrace.render(sc, max_depth=3, seed=5)
    assert np.array_equal(after.rows, fresh.rows)


def test_antialias_zero_is_nan_white(gpu, scenes):
    fr = gpu.raytrace.render(gpu.Scene(scenes.materials_showcase(16, 12, 1)), antialias=0)
    assert (fr.bgr == 255).all()  # 0/0 -> NaN -> code 255 (src/main.rs:56, src/color.rs:593-600)


@pytest.mark.parametrize("world", [2, 3])
def test_row_band_shards_reassemble(gpu, scenes, world):
    import torch
    desc = scenes.random_spheres(500, 9, 15.0, 200, 150)  # 150 rows: not a multiple of band*world
    sc = gpu.Scene(desc)
    whole = gpu.raytrace.render(sc, max_depth=4, seed=6)
    parts, ctxs = [], []
    for r in range(world):
        ctx = gpu.Context(sc, max_depth=4, seed=6, rank=r, world=world)
        ctx.render()
        ptr, nbytes = ctx.framebuffer_ptr()
        rows, _, _ = ctx.download()
        parts.append(torch.from_numpy(rows.reshape(-1)).cuda())
        assert rows.size == nbytes
        ctxs.append(ctx)
    gathered = torch.cat(parts).contiguous()
    out = torch.zeros(whole.rows.size, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    ctxs[0].untile(gathered.data_ptr(), out.data_ptr())
    ctxs[0].synchronize()
    got = out.cpu().numpy().reshape(whole.rows.shape)
    assert (got != whole.rows).mean() < 1e-5
    assert np.array_equal(got[:, 600:], whole.rows[:, 600:])  # row padding bytes


def test_config2_cornell_4k_1spp(gpu, orc):
    """BASELINE config 2: test scene at 3840x2160, 1 spp, full-frame pixel match against the CPU oracle."""
    from oracle import scene_text
    desc = scene_text.deserialize(REF_TEXT_SCENE)
    kw = dict(width=3840, height=2160, antialias=1, seed=1)
    ref = orc.OracleScene(desc).render(want_accum=False, **kw)
    fr = gpu.raytrace.render(gpu.serialize.deserialize(REF_TEXT_SCENE), **kw)
    same = assert_parity(fr, ref, "config2")
    print("config2: identical bytes %.6f, rays %d, %.2f ms" % (same, fr.stats["rays_total"], fr.stats["render_ms"]))


# ---------------------------------------------------------------------------------------------------------------------
# BASELINE.json configurations at full size.  The CPU oracle is O(rays x objects), so full frames are checked against
# the GPU's own f64 brute-force scan and a seed-determined sparse pixel set against the oracle.
# ---------------------------------------------------------------------------------------------------------------------
def _sparse_check(frame, osc, n_pixels, what, **kw):
    h, w = frame.ids.shape
    px = np.random.default_rng(2026).choice(w * h, size=n_pixels, replace=False).astype(np.uint32)
    ref = osc.render_pixels(px, width=w, height=h, **kw)
    ys, xs = px // w, px % w
    assert np.array_equal(frame.ids[ys, xs], ref["ids"]), "%s: primary ids differ on the sparse pixel set" % what
    diff = np.abs(frame.bgr[ys, xs].astype(np.int16) - ref["bgr"].astype(np.int16))
    assert diff.max() <= 1, "%s: %d LSB" % (what, int(diff.max()))
    same = float((diff == 0).mean())
    assert same >= 0.999, "%s: %.5f identical" % (what, same)
    return same


def test_config3_spheres10k_4k_depth8(gpu, orc, scenes):
    desc = scenes.spheres_10k()
    sc = gpu.Scene(desc)
    kw = dict(max_depth=8, seed=3)
    bvh = gpu.raytrace.render(sc, accel=gpu.ACCEL_BVH, **kw)
    lin = gpu.raytrace.render(sc, accel=gpu.ACCEL_LINEAR, **kw)            # ~1.6 s of f64 brute force: the full-frame oracle
    assert np.array_equal(bvh.ids, lin.ids), "BVH and scan disagree on %d primary ids" % int((bvh.ids != lin.ids).sum())
    assert bvh.stats["rays_total"] == lin.stats["rays_total"] and bvh.stats["rays_shadow"] == lin.stats["rays_shadow"]
    assert (bvh.rows != lin.rows).mean() < 1e-5                            # identical rays; only the atomic accumulation order differs
    same = _sparse_check(bvh, orc.OracleScene(desc), 4096, "config3", antialias=1, **kw)
    print("config3: %d rays, bvh %.2f ms, scan %.1f ms, sparse identical %.5f" % (bvh.stats["rays_total"], bvh.stats["render_ms"], lin.stats["render_ms"], same))


def test_config4_spheres1m_8k_depth8(gpu, orc, scenes):
    """BASELINE config 4 at its own size: 1M spheres through the device BVH at 7680x4320, depth 8.
      * final pixels and primary ids of a seed-determined sparse pixel set against the CPU oracle (linear scan, recursion);
      * every primary id of one 1/8 row-band shard (4.1 M rays) against the GPU's f64 brute-force scan of the same shard;
      * the frame assembled from eight shards (rt_multi, the shards sharing this GPU) against the one-context frame."""
    desc = scenes.spheres_1m(7680, 4320)
    sc = gpu.Scene(desc)
    kw = dict(max_depth=8, seed=4)
    bvh = gpu.raytrace.render(sc, accel=gpu.ACCEL_BVH, **kw)
    assert bvh.ids.shape == (4320, 7680) and bvh.stats["rays_primary"] == 7680 * 4320
    same = _sparse_check(bvh, orc.OracleScene(desc), 1024, "config4", antialias=1, **kw)
    # (almost) primary rays only through the scan: max_depth 0 and an unreachable significance threshold leave only the
    # refraction children of transparent hits (upstream spawns those without a significance test, src/raytrace.rs:214-223)
    shard = dict(rank=3, world=8)
    a = gpu.Context(sc, accel=gpu.ACCEL_BVH, **kw, **shard)
    a.render()
    _, ids_bvh, _ = a.download(want_rows=False, want_ids=True)
    a.close()
    p = gpu.Context(sc, accel=gpu.ACCEL_LINEAR, max_depth=0, min_significance=1e30, seed=4, **shard)
    st = p.render()
    _, ids_scan, _ = p.download(want_rows=False, want_ids=True)
    p.close()
    assert st["rays_primary"] == 7680 * 4320 // 8 and st["rays_shadow"] == 0 and st["rays_secondary"] < 0.1 * st["rays_primary"]
    assert np.array_equal(ids_bvh, ids_scan), "BVH and scan disagree on %d primary ids of the shard" % int((ids_bvh != ids_scan).sum())
    # the shard's rows are the frame's bands 3, 11, 19, ...: same ids as the whole frame there
    rows_of_shard = (np.arange(ids_bvh.shape[0]) // 4 * 8 + 3) * 4 + np.arange(ids_bvh.shape[0]) % 4
    assert np.array_equal(ids_bvh, bvh.ids[rows_of_shard])
    gpu.release_cached_memory()   # the whole-frame context's 29 GB of queues: eight shard contexts are about to share this GPU
    multi = gpu.raytrace.render_multi(sc, [0] * 8, gather=gpu.GATHER_PEER, accel=gpu.ACCEL_BVH, **kw)
    assert np.array_equal(multi.ids, bvh.ids) and multi.stats["rays_total"] == bvh.stats["rays_total"]
    d = np.abs(multi.rows.astype(np.int16) - bvh.rows.astype(np.int16))
    assert d.max() <= 1 and (d != 0).mean() < 1e-5
    print("config4@8K: %d rays, bvh %.1f ms (%d nodes), shard scan %.0f ms, sparse identical %.5f" % (
        bvh.stats["rays_total"], bvh.stats["render_ms"], bvh.stats["n_bvh_nodes"], st["render_ms"], same))


def test_config5_skybox100k_4k_16spp(gpu, orc, scenes):
    """BASELINE config 5 at its own size: 100k reflective spheres under six 1024x1024 skybox faces (the only texture path
    upstream), 3840x2160 at 16 samples per pixel (133 M paths in primary batches); sparse pixel set against the CPU oracle"""
    desc = scenes.skybox_100k(3840, 2160, antialias=16, face_size=1024)
    fr = gpu.raytrace.render(gpu.Scene(desc), max_depth=4, seed=5)
    assert fr.stats["paths"] == 3840 * 2160 * 16 and fr.stats["host_levels"] == 0
    same = _sparse_check(fr, orc.OracleScene(desc), 1024, "config5", antialias=16, max_depth=4, seed=5)
    print("config5@16spp: %d rays, %.1f ms, sparse identical %.5f" % (fr.stats["rays_total"], fr.stats["render_ms"], same))


def test_cli_renders_test_scene_to_out_bmp(gpu, tmp_path):
    """the reference binary's behaviour (src/main.rs:16,34): ./test_scene.txt -> ./out.bmp; same bytes as the library call"""
    import subprocess
    cli = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "rust-raytrace_b200", "lib", "raytrace_b200")
    (tmp_path / "test_scene.txt").write_text(REF_TEXT_SCENE)
    r = subprocess.run([cli, "--width", "200", "--height", "160", "--antialias", "2", "--seed", "5", "--stats"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    assert "Mrays/s" in r.stderr
    raw = (tmp_path / "out.bmp").read_bytes()
    fr = gpu.raytrace.render(gpu.serialize.deserialize(REF_TEXT_SCENE), width=200, height=160, antialias=2, seed=5)
    assert len(raw) == 122 + fr.rows.size
    assert raw[:2] == b"BM"
    got = np.frombuffer(raw[122:], dtype=np.uint8).reshape(fr.rows.shape)
    diff = np.abs(got.astype(np.int16) - fr.rows.astype(np.int16))   # f64 atomics: the sum order differs between two runs
    assert diff.max() <= 1 and (diff == 0).mean() >= 0.999


def test_config1_test_scene_at_its_own_options_matches_out_bmp(gpu, orc):
    """BASELINE config 1: test_scene.txt unmodified (800x800, 1024 spp, depth 4; 6.6e8 paths) on the GPU, against the
    8x8-block linear means of upstream's own out.bmp (tests/golden/outbmp_block_means.npz) -- a statistical golden, the
    only rendered artefact upstream ships"""
    golden = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "outbmp_block_means.npz"))
    sc = gpu.serialize.deserialize(REF_TEXT_SCENE)
    assert sc.options() == (800, 800, 1024)
    fr = gpu.raytrace.render(sc, seed=11, want_ids=False)
    assert fr.stats["paths"] == 800 * 800 * 1024
    v, _ = orc.srgb_tables()
    lin = v[fr.bgr]                                                     # decoded 8-bit codes, as a BMP reader sees them
    mine = lin.reshape(100, 8, 100, 8, 3).mean(axis=(1, 3))
    g = golden["linear_bgr"].astype(np.float64)
    ok = (golden["saturated"].max(axis=2) < 0.02)[:, :, None] & (g > 2e-3)
    ratio = mine[ok] / g[ok]
    assert ok.sum() > 3000
    assert abs(np.median(ratio) - 1.0) < 0.01, np.median(ratio)
    # per-block noise of two independent 1024-spp renders (64 pixels per block): a few per cent
    assert np.abs(ratio - 1.0).mean() < 0.05, np.abs(ratio - 1.0).mean()
    assert np.quantile(np.abs(ratio - 1.0), 0.99) < 0.15, np.quantile(np.abs(ratio - 1.0), 0.99)
    for flipped in (mine[::-1], mine[:, ::-1]):                         # controls: orientation matters
        assert np.abs(flipped[ok] / g[ok] - 1.0).mean() > 3.0 * np.abs(ratio - 1.0).mean()


@pytest.mark.parametrize("n", [1, 2, 3, 65])
def test_tiny_trees(gpu, orc, scenes, n):
    """one, two, three spheres forced through the device BVH (single-leaf root, one internal node, ...), and the smallest
    scene ACCEL_AUTO sends there"""
    desc = scenes.random_spheres(n, 40 + n, 3.0, 64, 48)
    rng = np.random.default_rng(n)
    rays = _random_rays(rng, 4000, 5.0, 6.0)
    data = desc["objects"][1]["data"]
    aim = data[rng.integers(0, n, 2000), :3] + rng.normal(size=(2000, 3)) * 0.3
    rays[:2000, 3:] = aim - rays[:2000, :3]
    rid, rt_, rn = orc.OracleScene(desc).trace_rays(rays)
    ids, t, nrm = gpu.raytrace.intersect(gpu.Scene(desc), rays, accel=gpu.ACCEL_BVH)
    assert np.array_equal(ids, rid) and np.array_equal(t, rt_) and np.array_equal(nrm, rn)
    fr = gpu.raytrace.render(gpu.Scene(desc), max_depth=4, seed=n, accel=gpu.ACCEL_BVH)
    assert_parity(fr, orc.OracleScene(desc).render(max_depth=4, seed=n), "tiny tree n=%d" % n)
    assert fr.stats["accel_used"] == gpu.ACCEL_BVH and fr.stats["n_bvh_nodes"] == max(1, n - 1)


def _stress_scene(S, rng, n, centre, extent, rlo, rhi):
    b = S.SceneBuilder()
    m = b.add_material(S.PhongMaterial((0.5, 0.5, 0.5), (0.1, 0.1, 0.1), 10.0, (0.01, 0.01, 0.01)))
    b.add_object(S.Plane((centre[0], centre[1] - extent[1] - 1.0, centre[2]), (0.0, 1.0, 0.0)), m)
    c = centre + rng.uniform(-1.0, 1.0, size=(n, 3)) * extent
    r = np.exp(rng.uniform(np.log(rlo), np.log(rhi), n))
    b.add_spheres(np.concatenate([c, r[:, None]], axis=1), np.zeros(n, dtype=np.uint32))
    b.add_light(S.PointLight((centre[0], centre[1] + 3 * extent[1] + 10, centre[2]), (1.0, 1.0, 1.0)))
    b.camera = S.SimplePerspectiveCamera.new((centre[0], centre[1], centre[2] + 2 * extent[2] + 5), (0.0, 0.0, -1.0), (0.0, 1.0, 0.0), 2.0)
    b.background = S.SolidColorBackground((0.1, 0.1, 0.1))
    b.options = S.Options(8, 8, 1)
    return b.build(), c, r


@pytest.mark.parametrize("case", range(8))
def test_bvh_filter_is_conservative_on_odd_scenes(gpu, case):
    """the f32 / 16-bit-grid filter in front of the exact test must never lose a hit: BVH == GPU linear scan, bit for bit,
    on scenes far from the origin, flat, needle-like, with radii over five decades, and with rays that start inside,
    far outside, on axis planes and along box faces"""
    rng = np.random.default_rng(100 + case)
    centre, extent, rlo, rhi = [
        ((0.0, 0.0, 0.0), (50.0, 50.0, 50.0), 0.05, 5.0),
        ((4000.0, -2500.0, 9000.0), (40.0, 5.0, 40.0), 0.1, 1.0),        # far from the origin: large |c|, small extent
        ((0.0, 0.0, 0.0), (2000.0, 0.5, 2000.0), 0.2, 0.8),              # very flat
        ((0.0, 0.0, 0.0), (0.2, 300.0, 0.2), 0.01, 0.3),                 # a needle
        ((10.0, 10.0, 10.0), (30.0, 30.0, 30.0), 1e-3, 30.0),            # radii over 4.5 decades, heavy overlap
        ((-1e5, 2e5, 5e4), (500.0, 500.0, 500.0), 0.5, 5.0),             # 1e5 from the origin
        ((0.0, 0.0, 0.0), (1e-3, 1e-3, 1e-3), 1e-5, 1e-4),               # a tiny scene
        ((0.0, 0.0, 0.0), (3e4, 10.0, 3e4), 0.3, 0.6),                   # grid too coarse: f32 nodes are chosen
    ][case]
    centre, extent = np.array(centre), np.array(extent)
    desc, c, r = _stress_scene(gpu.scene, rng, 6000, centre, extent, rlo, rhi)
    sc = gpu.Scene(desc)
    n = 60000
    k = rng.integers(0, len(c), n)
    target = c[k] + rng.normal(size=(n, 3)) * r[k][:, None] * rng.choice([0.3, 1.0, 1.5], n)[:, None]
    dist = np.exp(rng.uniform(np.log(np.linalg.norm(extent) * 1e-3), np.log(np.linalg.norm(extent) * 1e3), n))
    dirs = rng.normal(size=(n, 3)); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    o = target - dirs * dist[:, None]
    d = dirs * np.exp(rng.uniform(-3, 3, n))[:, None]                      # un-normalised directions
    d[:2000, 0] = 0.0; d[2000:4000, 1] = 0.0; d[4000:5000, [0, 2]] = 0.0  # axis-plane and axis-parallel rays
    o[5000:6000] = c[k[5000:6000]]                                         # from sphere centres
    rays = np.concatenate([o, d], axis=1)
    a = gpu.raytrace.intersect(sc, rays, accel=gpu.ACCEL_LINEAR)
    b = gpu.raytrace.intersect(sc, rays, accel=gpu.ACCEL_BVH)
    assert np.array_equal(a[0], b[0]), "%d ids differ" % int((a[0] != b[0]).sum())
    assert np.array_equal(a[1], b[1], equal_nan=True) and np.array_equal(a[2], b[2], equal_nan=True)
    assert (a[0] > 0).mean() > 0.15
    # the same through whole frames (shadow queries and reflections included)
    fb = gpu.raytrace.render(sc, width=96, height=64, antialias=2, max_depth=3, seed=case, accel=gpu.ACCEL_BVH)
    fl = gpu.raytrace.render(sc, width=96, height=64, antialias=2, max_depth=3, seed=case, accel=gpu.ACCEL_LINEAR)
    assert np.array_equal(fb.ids, fl.ids)
    assert fb.stats["rays_total"] == fl.stats["rays_total"] and fb.stats["rays_shadow"] == fl.stats["rays_shadow"]
    diff = np.abs(fb.rows.astype(np.int16) - fl.rows.astype(np.int16))
    assert diff.max() <= 1 and (diff == 0).mean() >= 0.999


def test_two_frames_in_flight_equal_one_at_a_time(gpu, orc, scenes):
    """rt_context_render_begin / _end: two contexts queue their frames back to back and are collected afterwards -- same rays,
    same ids, same bytes as one blocking rt_context_render each; one _end per _begin."""
    desc = scenes.random_spheres(600, 31, 14.0, 192, 108, antialias=2)
    sc = gpu.Scene(desc)
    a = gpu.Context(sc, max_depth=6, seed=2)
    b = gpu.Context(sc, max_depth=6, seed=9)
    one = {}
    for name, c in (("a", a), ("b", b)):
        st = c.render()
        rows, ids, _ = c.download(want_ids=True)
        one[name] = (st, rows.copy(), ids.copy())
    ref = orc.OracleScene(desc).render(max_depth=6, seed=2)
    assert np.array_equal(one["a"][2][:108], ref["ids"]) and one["a"][0]["rays_total"] == ref["stats"]["rays"]
    for _ in range(3):
        a.render_begin()
        b.render_begin()
        with pytest.raises(gpu.RtError):
            a.render_begin()                       # the frame in flight has not been collected
        sa, sb = a.render_end(), b.render_end()
        for name, c, st in (("a", a, sa), ("b", b, sb)):
            rows, ids, _ = c.download(want_ids=True)
            assert st["rays_total"] == one[name][0]["rays_total"] and st["host_levels"] == 0
            assert np.array_equal(ids, one[name][2])
            assert (rows != one[name][1]).mean() < 1e-5     # order of f64 atomic adds into a pixel only
    with pytest.raises(gpu.RtError):
        a.render_end()                             # nothing in flight
    assert not np.array_equal(one["a"][1], one["b"][1])     # two different frames were in flight
    a.close(); b.close()


def test_replayed_frame_graph_equals_plain_launches(gpu, orc, scenes):
    """With RT_TUNE_FRAME_GRAPH a context that renders the same frame again replays a CUDA graph of it (captured when the second
    frame is queued from the same arguments); by default every launch is queued every time.  Same counts, ids and bytes --
    also after the scene moved (new device pointers: plain launches again, then a new graph) and with several primary batches."""
    desc = scenes.random_spheres(500, 41, 13.0, 160, 90, antialias=2)
    sc = gpu.Scene(desc)
    ref = orc.OracleScene(desc).render(max_depth=6, seed=4)
    plain = gpu.Context(sc, max_depth=6, seed=4)
    st0 = plain.render()
    rows0, ids0, _ = plain.download(want_ids=True)
    assert np.array_equal(ids0[:90], ref["ids"]) and st0["rays_total"] == ref["stats"]["rays"]
    for kw in (dict(), dict(max_rays_in_flight=8192)):          # one batch / several batches per frame
        ctx = gpu.Context(sc, max_depth=6, seed=4, tuning=gpu.TUNE_FRAME_GRAPH, **kw)
        for frame in range(4):
            st = ctx.render()
            rows, ids, _ = ctx.download(want_ids=True)
            assert st["host_levels"] == 0 and st["rays_total"] == st0["rays_total"], (kw, frame)
            for k in ("rays_primary", "rays_secondary", "rays_shadow", "paths"):
                assert st[k] == st0[k]
            assert np.array_equal(ids, ids0) and (rows != rows0).mean() < 1e-5
        ctx.close()
    plain.close()
    # rt_render_multi keeps its contexts and re-binds them: unchanged scene -> graph, moved spheres -> plain launches, same frames
    data = None
    first = 0
    for ob in desc["objects"]:
        if ob["shape"] == "spheres":
            data = np.ascontiguousarray(ob["data"], dtype=np.float64)
            break
        first += 1
    G = dict(gather=gpu.GATHER_DIRECT, max_depth=6, seed=4, tuning=gpu.TUNE_FRAME_GRAPH)
    frames = [gpu.raytrace.render_multi(sc, [0, 0], **G) for _ in range(3)]
    moved = data.copy(); moved[:, 1] += 0.25
    sc.update_spheres(first, moved)
    up = [gpu.raytrace.render_multi(sc, [0, 0], **G) for _ in range(3)]
    sc.update_spheres(first, data)
    back = gpu.raytrace.render_multi(sc, [0, 0], **G)
    for fr in frames[1:] + [back]:
        assert np.array_equal(fr.ids, frames[0].ids) and (fr.rows != frames[0].rows).mean() < 1e-5
    assert np.array_equal(frames[0].ids, ref["ids"])
    for fr in up[1:]:
        assert np.array_equal(fr.ids, up[0].ids) and (fr.rows != up[0].rows).mean() < 1e-5
    assert not np.array_equal(up[0].rows, frames[0].rows)
