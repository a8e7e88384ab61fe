"""Pins the CPU oracle (oracle/oracle.cpp) before anything is allowed to trust it.

Upstream ships no tests, so the pins are
  * the literals upstream's behaviour is defined by (sRGB tables of src/color.rs, BMP header bytes of out.bmp) --
    committed under tests/golden/ by tests/golden/make_golden.py;
  * known answers derived by hand from the cited reference lines;
  * the statistical golden out.bmp (tests/test_oracle_outbmp.py).
"""
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def S():
    import importlib
    return importlib.import_module("rust-raytrace_b200.scene")


def tiny_scene(objects, materials=None, lights=(), camera=None, background=(0, 0, 0), options=(8, 8, 1)):
    s = S()
    b = s.SceneBuilder()
    mats = materials or [s.PhongMaterial((0.5, 0.5, 0.5), (0, 0, 0), 1.0, (0, 0, 0))]
    for m in mats:
        b.add_material(m)
    for shape, mi in objects:
        b.add_object(shape, mi)
    for l in lights:
        b.add_light(l)
    b.camera = camera or s.SimplePerspectiveCamera.new((0, 3, 17), (0, 0, -1), (0, 1, 0), 3.6)
    b.background = s.SolidColorBackground(background)
    b.options = s.Options(*options)
    return b.build()


# ---- quantiser + tables (src/color.rs:75-332, 335-591, 593-600) -------------------------------------------
def test_srgb_tables_equal_upstream_literals(orc):
    g = np.load(os.path.join(GOLDEN, "srgb_tables.npz"))
    v, a = orc.srgb_tables()
    assert np.array_equal(v, g["values"]), "SRGB_VALUES differ from the literals of src/color.rs:75-332"
    assert np.array_equal(a, g["average"]), "SRGB_AVERAGE differ from the literals of src/color.rs:335-591"
    assert np.array_equal(a, (v[:-1] + v[1:]) / 2)  # exact midpoints
    assert np.all(np.diff(a) > 0)
    d = np.diff(a)
    assert abs(d.min() - 3.0353e-4) < 1e-7 and abs(d.max() - 8.8747e-3) < 1e-6


def test_quantiser_known_answers(orc):
    v, a = orc.srgb_tables()
    q = lambda x: int(orc.quantize_bgr([[x, x, x]])[0, 0])
    assert q(0.0) == 0 and q(-1.0) == 0 and q(-0.0) == 0
    assert q(a[0]) == 1 and q(np.nextafter(a[0], -np.inf)) == 0  # `<`, not `<=` (src/color.rs:595)
    assert abs(a[0] - 0.000151763491774418758) < 1e-18
    assert q(0.051) == 64 and q(0.18) == 118 and q(0.5) == 188
    assert q(1.0) == 255 and q(5.0) == 255 and q(np.inf) == 255 and q(np.nan) == 255 and q(-np.inf) == 0
    for i in range(255):  # every threshold, both sides
        assert q(a[i]) == i + 1 and q(np.nextafter(a[i], -np.inf)) == i
    # write_bgr stores b, g, r (src/color.rs:628-632)
    assert orc.quantize_bgr([[0.0, 0.18, 1.0]]).tolist() == [[255, 118, 0]]


def test_quantiser_matches_linear_search(orc):
    v, a = orc.srgb_tables()
    x = np.random.default_rng(0).uniform(-0.05, 1.1, 30000)
    want = np.array([next((i for i in range(255) if val < a[i]), 255) for val in x[:3000]], dtype=np.uint8)
    got = orc.quantize_bgr(np.repeat(x[:3000, None], 3, axis=1))[:, 0]
    assert np.array_equal(got, want)


# ---- BMP (src/bmp.rs:10-61) -------------------------------------------------------------------------------
def test_bmp_header_equals_outbmp_header(orc):
    g = np.load(os.path.join(GOLDEN, "outbmp_block_means.npz"))
    assert orc.bmp_header(800, 800) == g["header"].tobytes(), "header differs from the first 122 bytes of upstream out.bmp"
    want = bytes.fromhex("424d7a4c1d00" "00000000" "7a000000" "6c000000" "20030000" "20030000" "0100" "1800" "00000000" "004c1d00"
                         "130b0000" "130b0000") + bytes(8) + bytes(16) + bytes.fromhex("42475273") + bytes(48)
    assert orc.bmp_header(800, 800) == want
    assert [orc.bytewidth(w) for w in (1, 2, 3, 4, 5, 800, 3840, 7680)] == [4, 8, 12, 12, 16, 2400, 11520, 23040]
    h = orc.bmp_header(3841, 7)
    assert int.from_bytes(h[2:6], "little") == 122 + 11524 * 7 and int.from_bytes(h[34:38], "little") == 11524 * 7


# ---- RNG conventions (rand 0.3; unpinned upstream) ------------------------------------
def test_xorshift128_recurrence(orc):
    import ctypes as C
    x, y, z, w = 0x193A6754, 0xA8A7D469, 0x97830E05, 0x113BA7BB  # rand 0.3 XorShiftRng::new_unseeded constants
    out = np.zeros(64, dtype=np.uint32)
    orc.lib().orc_xorshift(x, y, z, w, 64, out.ctypes.data_as(C.POINTER(C.c_uint32)))
    ref = []
    for _ in range(64):
        t = (x ^ (x << 11)) & 0xFFFFFFFF
        x, y, z = y, z, w
        w = (w ^ (w >> 19) ^ (t ^ (t >> 8))) & 0xFFFFFFFF
        ref.append(w)
    assert out.tolist() == ref


def test_counter_rng_is_splitmix_finaliser(orc):
    M = (1 << 64) - 1
    G = 0x9E3779B97F4A7C15

    def mix(z):
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
        return z ^ (z >> 31)

    L = orc.lib()
    for seed, pixel, aa in [(1, 0, 0), (3, 8294399, 0), (2**63 + 5, 33177599, 1023)]:
        k = mix(seed ^ 0x243F6A8885A308D3)
        k = mix((k + G * (pixel + 1)) & M)
        k = mix((k + G * (aa + 1)) & M)
        assert L.orc_key_sample(seed, pixel, aa) == k
        assert L.orc_child_key(k, 1) == mix((k + G * 2) & M)
        assert L.orc_draw_bits(k, 5) == mix(((k ^ 0xD6E8FEB86659FD93) + G * 6) & M)


# ---- camera (src/camera.rs:51-79) -------------------------------------------------------------------------
def test_camera_known_answers(orc):
    sc = orc.OracleScene(tiny_scene([]))
    pos, m = sc.camera()
    assert pos.tolist() == [0, 3, 17]
    # columns u = (1,0,0), v = (0,1,0), w = (0,0,-3.6)  (test_scene.txt:99-104)
    assert np.array_equal(m, np.array([[1, 0, 0], [0, 1, 0], [0, 0, -3.6]]))
    r = sc.project([[0.0, 0.0], [1.0, 0.0], [0.0, -1.0]])
    assert r[0].tolist() == [0, 3, 17, 0, 0, -1]
    n = np.sqrt(1 + 3.6 * 3.6)
    assert np.array_equal(r[1, 3:], np.array([1 / n, 0, -3.6 / n]))
    assert np.array_equal(r[2, 3:], np.array([0, -1 / n, -3.6 / n]))


def test_look_at_camera(orc):
    s = S()
    cam = s.SimplePerspectiveCamera.look_at((0.0, 1.0, 0.0), (0.0, 0.0, -1.0), (0.0, 1.0, 0.0), np.deg2rad(90.0), 2.0)
    sc = orc.OracleScene(tiny_scene([], camera=cam))
    pos, m = sc.camera()
    # look_at (src/camera.rs:67-73): im_dist = 1/tan(pov/2), position = focus - normalize(look) * h * im_dist
    im = 1.0 / np.tan(np.deg2rad(90.0) / 2)
    assert np.allclose(pos, [0, 1, 2.0 * im], rtol=0, atol=1e-15)
    assert np.allclose(m[:, 2], [0, 0, -im], rtol=0, atol=1e-15)


# ---- shapes + Scene::intersect (src/shapes.rs:50-112, src/scene.rs:223-249) ---------------------------------
def test_intersect_known_answers(orc):
    s = S()
    desc = tiny_scene([(s.Plane((0, 0, -3), (0, 0, 1)), 0),      # 0
                       (s.Sphere((0, 1.5, 0), 1.5), 0),           # 1
                       (s.Sphere((0, 1.5, 0), 1.5), 0),           # 2 coincident with 1
                       (s.Plane((0, -1, 0), (0, 1, 0)), 0),       # 3
                       (s.Sphere((5, 5, 5), 1.0), 0)])            # 4
    sc = orc.OracleScene(desc)
    rays = np.array([
        [0, 3, 17, 0, 0, -1],      # tangent to sphere 1: disc == 0 -> miss (src/shapes.rs:66); plane 0 at t = 20
        [0, 1.5, 17, 0, 0, -1],    # through the centre: first root 15.5, tie between 1 and 2 -> 1
        [0, 1.5, 0, 0, 0, 1],      # inside: near root negative -> far root 1.5 (src/shapes.rs:76-83)
        [0, 1.5, 0, 0, 0, -2.5],   # inside, un-normalised direction: t = 1.5/2.5
        [0, 5, 5, 1, 0, 0],        # parallel to plane 3, numerator -6 -> -inf -> rejected; sphere 4 at t = 4
        [0, -5, 5, 1, 0, 0],       # parallel, numerator +4 -> t = +inf accepted (src/shapes.rs:103 rejects only t <= 0)
        [0, -1, 5, 1, 0, 0],       # in plane 3: 0/0 = NaN is a hit whose key sorts first (src/scene.rs:227-233)
        [5, 20, 5, 0, -1, 0],      # sphere 4 (t = 14) before plane 3 (t = 21)
        [0, 0, -3, 0, 0, -1],      # starts on plane 0: t = 0 rejected; plane 3 gives -inf -> miss
    ], dtype=np.float64)
    ids, t, n = sc.trace_rays(rays)
    assert ids.tolist() == [0, 1, 1, 1, 4, 3, 3, 4, -1]
    assert t[0] == 20.0 and n[0].tolist() == [0, 0, 1]
    assert t[1] == 15.5 and n[1].tolist() == [0, 0, 1]
    assert t[2] == 1.5 and n[2].tolist() == [0, 0, 1]
    assert t[3] == 0.6
    assert t[4] == 4.0 and np.isinf(t[5]) and t[5] > 0 and np.isnan(t[6]) and np.isinf(t[8])
    assert t[7] == 14.0 and n[7].tolist() == [0, 1, 0]
    # plane normal is returned as stored: not normalised, not flipped (src/shapes.rs:108)
    sc2 = orc.OracleScene(tiny_scene([(s.Plane((0, 0, 0), (0, -2, 0)), 0)]))
    ids, t, n = sc2.trace_rays([[0, 5, 0, 0, -1, 0]])
    assert ids[0] == 0 and t[0] == 5.0 and n[0].tolist() == [0, -2, 0]


def test_intersect_against_numpy_restating_shapes_rs(orc, scenes):
    """Independent numpy restatement of Sphere::intersect / Plane::intersect + min_by_key over 300 objects."""
    desc = scenes.random_spheres(300, 42, 10.0, 8, 8)
    data = desc["objects"][1]["data"]
    rng = np.random.default_rng(7)
    n = 4000
    o = np.stack([rng.uniform(-12, 12, n), rng.uniform(0.1, 8, n), rng.uniform(-12, 12, n)], 1)
    d = rng.normal(size=(n, 3))
    ids, t, _ = orc.OracleScene(desc).trace_rays(np.concatenate([o, d], 1))
    best_t = np.full(n, np.inf)
    best = np.full(n, -1)
    with np.errstate(all="ignore"):
        tp = (0 * (0 - o[:, 0]) + 1 * (0 - o[:, 1]) + 0 * (0 - o[:, 2])) / (0 * d[:, 0] + 1 * d[:, 1] + 0 * d[:, 2])
        hit = ~(tp <= 0)
        best_t[hit], best[hit] = tp[hit], 0
        a = d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]
        for k in range(data.shape[0]):
            oc = o - data[k, :3]
            b = 2.0 * (d[:, 0] * oc[:, 0] + d[:, 1] * oc[:, 1] + d[:, 2] * oc[:, 2])
            c = (oc[:, 0] * oc[:, 0] + oc[:, 1] * oc[:, 1] + oc[:, 2] * oc[:, 2]) - data[k, 3] * data[k, 3]
            disc = b * b - 4.0 * a * c
            sq = np.sqrt(disc)
            t1, t2 = (-b - sq) / (2.0 * a), (-b + sq) / (2.0 * a)
            th = np.where(t1 > 0, t1, np.where(t2 > 0, t2, np.nan))
            ok = (disc > 0) & ~np.isnan(th) & (th < best_t)
            best_t[ok], best[ok] = th[ok], k + 1
    assert np.array_equal(ids, best)
    assert np.array_equal(t, best_t)


# ---- frame loop + materials (src/main.rs:39-59, src/raytrace.rs:30-276) -------------------------------------
def test_pixel_mapping_and_row_order(orc):
    s = S()
    # a wall facing the camera, emissive: top half of the view red-ish sphere absent -> just check orientation with a plane split
    mats = [s.PhongMaterial((0, 0, 0), (0, 0, 0), 1.0, (1.0, 0.0, 0.0)), s.PhongMaterial((0, 0, 0), (0, 0, 0), 1.0, (0.0, 0.0, 1.0))]
    desc = tiny_scene([(s.Plane((0, 0, 0), (0, 1, 0)), 0), (s.Plane((0, 0, -50), (0, 0, 1)), 1)], materials=mats,
                      camera=s.SimplePerspectiveCamera.new((0, 1, 10), (0, 0, -1), (0, 1, 0), 1.0), options=(16, 8, 1))
    r = orc.OracleScene(desc).render()
    bgr = r["bgr"][:, :48].reshape(8, 16, 3)
    # BMP row 0 is y = 0 = image bottom (src/main.rs:45-58): the floor (red ambient) is at the bottom rows, the wall (blue) on top
    assert (bgr[0, :, 2] == 255).all() and (bgr[0, :, 0] == 0).all()
    assert (bgr[7, :, 0] == 255).all() and (bgr[7, :, 2] == 0).all()
    assert (r["ids"][0] == 0).all() and (r["ids"][7] == 1).all()
    # 3840x2160: scale = 1/1080 (short side spans [-1, 1)), x in [-1.778, 1.778)
    sc = orc.OracleScene(tiny_scene([], options=(3840, 2160, 1)))
    assert max(1 / 1920.0, 1 / 1080.0) == 1 / 1080.0


def test_depth_cut_returns_ambient_only(orc):
    s = S()
    # two facing mirrors with ambient 0.01 and specular 1: pixel = ambient * (number of shaded bounces + 1)
    mir = s.PhongMaterial((0, 0, 0), (1.0, 1.0, 1.0), 1.0, (0.01, 0.01, 0.01))
    desc = tiny_scene([(s.Plane((0, 0, -5), (0, 0, 1)), 0), (s.Plane((0, 0, 5), (0, 0, -1)), 0)], materials=[mir],
                      camera=s.SimplePerspectiveCamera.new((0, 0, 0), (0, 0, -1), (0, 1, 0), 1.0), options=(2, 2, 1))
    for md in (0, 1, 4, 8):
        acc = orc.OracleScene(desc).render(max_depth=md)["accum"]
        # depths 0..md are shaded and reflect; the depth md+1 hit returns ambient only (src/raytrace.rs:33: `depth > MAX_DEPTH`)
        assert np.allclose(acc, 0.01 * (md + 2), rtol=1e-12, atol=0), (md, acc[0, 0])
        st = orc.OracleScene(desc).render(max_depth=md)["stats"]
        assert st["rays"] == 4 * (md + 2)


def test_significance_cull(orc):
    s = S()
    # specular significance 0.3 per bounce: 0.3^k * 1 > 1/512 holds for k <= 5 (0.3^5 = 2.43e-3 > 1.95e-3 > 0.3^6)
    mir = s.PhongMaterial((0, 0, 0), (0.1, 0.1, 0.1), 1.0, (0, 0, 0))
    desc = tiny_scene([(s.Plane((0, 0, -5), (0, 0, 1)), 0), (s.Plane((0, 0, 5), (0, 0, -1)), 0)], materials=[mir],
                      camera=s.SimplePerspectiveCamera.new((0, 0, 0), (0, 0, -1), (0, 1, 0), 1.0), options=(1, 1, 1))
    st = orc.OracleScene(desc).render(max_depth=50)["stats"]
    # hit k has sig 0.3^k; it reflects iff 0.3 * 0.3^k > 1/512 (src/raytrace.rs:36), i.e. k <= 4 -> 6 closest-hit rays
    assert st["rays"] == 6


def test_phong_direct_light_known_answer(orc):
    s = S()
    m = s.PhongMaterial((0.5, 0.25, 1.0), (0, 0, 0), 1.0, (0.1, 0.1, 0.1))
    desc = tiny_scene([(s.Plane((0, 0, 0), (0, 1, 0)), 0)], materials=[m], lights=[s.DirectionalLight((0, -2, 0), (1, 1, 1))],
                      camera=s.SimplePerspectiveCamera.new((0, 5, 0), (0, -1, 0), (0, 0, -1), 1.0), options=(1, 1, 1))
    acc = orc.OracleScene(desc).render()["accum"][0, 0]
    # ldir = -direction un-normalised (src/scene.rs:137): dot(ldir, n) = 2; diffuse * lcol * 2 * FRAC_1_PI + ambient
    want = np.array([0.5, 0.25, 1.0]) * 2.0 * (1 / np.pi) + 0.1
    assert np.allclose(acc, want, rtol=1e-15, atol=0)
    # point light: no distance fall-off (src/raytrace.rs:52); an occluder counts only within range (t*t < r2, :46-49).
    # im_dist 1e7 makes the jittered primary ray vertical to ~1e-7, so dot(ldir, n) = 1 to ~1e-14.
    cam = s.SimplePerspectiveCamera.new((0, 0.5, 0), (0, -1, 0), (0, 0, -1), 1e7)  # below the occluder
    desc = tiny_scene([(s.Plane((0, 0, 0), (0, 1, 0)), 0)], materials=[m], lights=[s.PointLight((0, 4, 0), (1, 1, 1))], camera=cam, options=(1, 1, 1))
    lit = np.array([0.5, 0.25, 1.0]) * (1 / np.pi) + 0.1
    assert np.allclose(orc.OracleScene(desc).render()["accum"][0, 0], lit, rtol=1e-12, atol=0)
    desc["lights"] = [s.PointLight((0, 400, 0), (1, 1, 1))]
    assert np.allclose(orc.OracleScene(desc).render()["accum"][0, 0], lit, rtol=1e-12, atol=0)
    far = tiny_scene([(s.Plane((0, 0, 0), (0, 1, 0)), 0), (s.Sphere((0, 10, 0), 1.0), 0)], materials=[m], lights=[s.PointLight((0, 4, 0), (1, 1, 1))],
                     camera=cam, options=(1, 1, 1))
    assert np.allclose(orc.OracleScene(far).render()["accum"][0, 0], lit, rtol=1e-12, atol=0)  # occluder beyond the light: not occluded
    far["objects"][1]["center"] = (0.0, 2.0, 0.0)                                               # between floor and light: occluded
    r = orc.OracleScene(far).render()
    assert np.allclose(r["accum"][0, 0], [0.1, 0.1, 0.1], rtol=0, atol=0)
    assert r["stats"]["rays"] == 2 and r["stats"]["shadow_rays"] == 1


def test_indirect_specular_is_nan_white(orc):
    s = S()
    # upstream's shadowed `ray` makes the specular term of IndirectPhong normalize(0) = NaN (src/raytrace.rs:108,115) -> code 255
    m = s.IndirectPhongMaterial((0.5, 0.5, 0.5), (0.5, 0.5, 0.5), 10.0, (0, 0, 0), 1)
    desc = tiny_scene([(s.Plane((0, 0, 0), (0, 1, 0)), 0)], materials=[m],
                      camera=s.SimplePerspectiveCamera.new((0, 5, 0), (0, -1, 0), (0, 0, -1), 1.0), options=(2, 2, 1))
    r = orc.OracleScene(desc).render()
    assert np.isnan(r["accum"]).all() and (r["bgr"][:, :6] == 255).all()


def test_antialias_zero_is_nan_white(orc):
    desc = tiny_scene([], options=(4, 4, 0))
    r = orc.OracleScene(desc).render(antialias=0)
    assert (r["bgr"][:, :12] == 255).all()  # 0/0 (src/main.rs:56) -> NaN -> 255 (src/color.rs:593-600)


def test_counter_rng_render_is_deterministic_and_thread_independent(orc, scenes):
    desc = scenes.materials_showcase(64, 48, 2)
    a = orc.OracleScene(desc).render(max_depth=5, seed=4, threads=1)
    b = orc.OracleScene(desc).render(max_depth=5, seed=4, threads=4)
    assert np.array_equal(a["bgr"], b["bgr"]) and np.array_equal(a["accum"], b["accum"], equal_nan=True)
    px = np.array([0, 17, 64 * 48 - 1, 1000], dtype=np.uint32)
    p = orc.OracleScene(desc).render_pixels(px, max_depth=5, seed=4)
    assert np.array_equal(p["accum"], a["accum"].reshape(-1, 3)[px])
    assert np.array_equal(p["ids"], a["ids"].reshape(-1)[px])
    c = orc.OracleScene(desc).render(max_depth=5, seed=5)
    assert not np.array_equal(a["bgr"], c["bgr"])
