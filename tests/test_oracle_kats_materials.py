"""Known answers for the oracle branches the upstream artefact out.bmp never exercises (missing #3):
FresnelMaterial, TransparentMaterial (entering / exiting with the un-normalised refract vector / total internal
reflection / the `fresnel < 1.0` gate), SkyboxBackground (face choice on all six faces, ties -> BLACK, bilinear
y-then-x with clamping), AreaLight draw order, DepthOfFieldCamera draw order.

The expected values come from tests/ref_numpy.py -- a restatement in plain Python written from the cited Rust lines,
independent of oracle/oracle.cpp -- and, where a closed form exists, from numbers worked out by hand in the test.
Both restatements are driven by the SAME xorshift128 stream in the reference's draw order (sequential mode), so the
comparison also pins which draw feeds which quantity.
"""
import importlib
import math

import numpy as np
import pytest

import ref_numpy as R

S = importlib.import_module("rust-raytrace_b200.scene")

STATE = (0x193a6754, 0xa8a7d469, 0x97830e05, 0x113ba7bb)   # rand 0.3 XorShiftRng::new_unseeded
RTOL = 1e-13   # same IEEE operations in the same order; libm pow / sqrt may differ in the last place


def build(objects, materials, lights=(), background=(0.2, 0.3, 0.5), camera=None, skybox=None):
    """the same scene as a description for the oracle and as a ref_numpy.Scene"""
    b = S.SceneBuilder()
    rm = []
    for m in materials:
        if m["kind"] == "phong":
            b.add_material(S.PhongMaterial(m["diffuse"], m["specular"], m["exponent"], m["ambient"]))
        elif m["kind"] == "fresnel":
            b.add_material(S.FresnelMaterial(m["diffuse"], m["specular"], m["exponent"], m["ambient"], m["ior"]))
        else:
            b.add_material(S.TransparentMaterial(m["specular"], m["exponent"], m["ior"]))
        rm.append({k: (tuple(float(x) for x in v) if isinstance(v, (tuple, list)) else v) for k, v in m.items()})
    ro = []
    for o in objects:
        if o[0] == "sphere":
            b.add_object(S.Sphere(o[1], o[2]), o[3])
        else:
            b.add_object(S.Plane(o[1], o[2]), o[3])
        ro.append((o[0], tuple(float(x) for x in o[1]), (float(o[2]) if o[0] == "sphere" else tuple(float(x) for x in o[2])), o[3]))
    rl = []
    for l in lights:
        if l[0] == "point":
            b.add_light(S.PointLight(l[1], l[2]))
        elif l[0] == "directional":
            b.add_light(S.DirectionalLight(l[1], l[2]))
        else:
            b.add_light(S.AreaLight(l[1], l[2], l[3], l[4]))
        rl.append(tuple(tuple(float(x) for x in v) if isinstance(v, (tuple, list)) else v for v in l))
    b.camera = camera or S.SimplePerspectiveCamera.new((0, 0, 10), (0, 0, -1), (0, 1, 0), 2.0)
    if skybox is not None:
        b.background = S.SkyboxBackground(*skybox)
    else:
        b.background = S.SolidColorBackground(background)
    b.options = S.Options(8, 8, 1)
    return b.build(), (ro, rm, rl)


def both(orc, desc, parts, rays, sig=1.0, depth=0, max_depth=4, bg=None):
    """ray_color of every ray from the oracle and from ref_numpy, one shared sequential xorshift stream"""
    ro, rm, rl = parts
    osc = orc.OracleScene(desc)
    got, n_rays = osc.ray_color(rays, significance=sig, depth=depth, max_depth=max_depth, xorshift_state=STATE)
    if bg is None:
        bg = ("solid", tuple(float(x) for x in desc["background"]["color"]))
    sc = R.Scene(ro, rm, rl, bg, max_depth=max_depth)
    rng = R.XorShift(*STATE)
    want = np.array([R.ray_color(sc, tuple(r[:3]), tuple(r[3:]), sig, depth, rng) for r in np.asarray(rays, dtype=np.float64)])
    assert n_rays == sc.intersect_calls, "Scene::intersect call counts differ: oracle %d, restatement %d" % (n_rays, sc.intersect_calls)
    return got, want


def close(got, want, rtol=RTOL):
    np.testing.assert_allclose(got, want, rtol=rtol, atol=1e-300)


GLASS = {"kind": "transparent", "specular": (0.9, 0.8, 0.7), "exponent": 40.0, "ior": 1.5}
FRES = {"kind": "fresnel", "diffuse": (0.6, 0.5, 0.4), "specular": (0.7, 0.8, 0.9), "exponent": 30.0, "ambient": (0.01, 0.02, 0.03), "ior": 1.8}


# ---- FresnelMaterial (src/raytrace.rs:123-167) ---------------------------------------------------------------------
def test_fresnel_head_on_is_r0_by_hand(orc):
    """normal incidence, no lights: F = r0 = ((ior-1)/(ior+1))^2; colour = ambient + specular * background * F"""
    desc, parts = build([("sphere", (0, 0, 0), 1.0, 0)], [FRES], background=(0.2, 0.3, 0.5))
    got, want = both(orc, desc, parts, [[0, 0, 5, 0, 0, -1]])
    r0 = ((1.8 - 1.0) / (1.8 + 1.0)) ** 2
    hand = [0.01 + 0.7 * 0.2 * r0, 0.02 + 0.8 * 0.3 * r0, 0.03 + 0.9 * 0.5 * r0]
    close(got[0], hand, rtol=1e-14)
    close(got, want)


def test_fresnel_with_lights_shadow_and_both_faces(orc):
    """oblique hits from outside and from inside (nd uses the UN-flipped normal, |nd| in omcos), a lit and a shadowed point
    light, an un-normalised directional light; reflection children leave to the background or hit the floor"""
    mats = [FRES, {"kind": "phong", "diffuse": (0.5, 0.5, 0.5), "specular": (0.0, 0.0, 0.0), "exponent": 1.0, "ambient": (0.05, 0.05, 0.05)}]
    objs = [("sphere", (0, 0, 0), 1.0, 0), ("plane", (0, -1.5, 0), (0, 1, 0), 1), ("sphere", (3, 3, 0), 0.5, 1)]
    lights = [("point", (6.0, 6.0, 0.0), (0.9, 0.8, 0.7)),          # behind the small sphere for hits near (0.7, 0.7, 0)
              ("point", (-4.0, 5.0, 3.0), (0.5, 0.6, 0.7)),
              ("directional", (0.3, -2.0, -0.4), (0.2, 0.2, 0.3))]
    desc, parts = build(objs, mats, lights)
    rays = [[0, 0.5, 5, 0, 0, -1], [0.3, 0.2, 5, 0.02, 0.01, -1], [3, 3, 3, -2.3, -2.3, -3], [0, 0, 0, 0.3, 0.5, 0.8],   # last: from inside
            [0.95, 0, 5, 0, 0, -1], [-5, 4, 0, 5.7, -3.3, 0.01]]
    got, want = both(orc, desc, parts, rays, max_depth=4)
    close(got, want)
    assert np.all(got > 0) and len({tuple(np.round(g, 9)) for g in got}) == len(rays)


def test_fresnel_specular_flag_uses_the_fresnel_factor(orc):
    """specular = specular.significance() * fresnel * significance > 1/512 (src/raytrace.rs:138): at normal incidence
    2.4 * r0 * sig straddles the threshold for sig = 0.0099 / 0.0101, and the reflection child exists only above it"""
    desc, parts = build([("sphere", (0, 0, 0), 1.0, 0)], [FRES], background=(1.0, 1.0, 1.0))
    r0 = ((1.8 - 1.0) / (1.8 + 1.0)) ** 2
    for sig, child in ((0.0099, False), (0.0101, True)):
        assert (2.4 * r0 * sig > 1.0 / 512.0) == child
        got, want = both(orc, desc, parts, [[0, 0, 5, 0, 0, -1]], sig=sig)
        close(got, want)
        hand = np.array([0.01, 0.02, 0.03]) + (np.array([0.7, 0.8, 0.9]) * r0 if child else 0.0)
        close(got[0], hand, rtol=1e-14)


# ---- TransparentMaterial (src/raytrace.rs:169-226) -----------------------------------------------------------------
def test_transparent_terms_by_hand():
    """the refraction vector and Schlick factor of :174-192 for the three cases, against closed forms"""
    # entering at 60 degrees: n = 1/1.5, sin2 = n^2 (1 - cos^2)
    d = (math.sin(math.radians(60)), -math.cos(math.radians(60)), 0.0)
    normal, refr, F = R.transparent_terms(1.5, (0.0, 1.0, 0.0), d)
    n = 1.0 / 1.5
    nd = d[1]
    sin2 = n * n * (1.0 - nd * nd)
    assert normal == (0.0, 1.0, 0.0)
    np.testing.assert_allclose(refr, (d[0] * n, d[1] * n - (n * abs(nd) + math.sqrt(1.0 - sin2)), 0.0), rtol=1e-15)
    # NOT Snell's vector: upstream subtracts normal * (n |nd| + cos) from d * n, whose normal component is already -n |nd|, so
    # the vector's normal part is -(2 n |nd| + cos) and its length is not 1 (it is normalised before it is traced, :219).
    # The tangential part is Snell's: sin(theta_t) = sin(60 deg) / 1.5.  Kept as upstream has it: it defines the image.
    np.testing.assert_allclose(refr[0], math.sin(math.radians(60)) / 1.5, rtol=1e-15)
    np.testing.assert_allclose(refr[1], -(2 * n * 0.5 + math.sqrt(1.0 - sin2)), rtol=1e-15)
    assert math.hypot(*refr[:2]) > 1.5
    r0 = 0.04
    np.testing.assert_allclose(F, r0 + (1 - r0) * (1 - 0.5) ** 5, rtol=1e-15)      # entering: omcos = 1 - |nd|
    # exiting at 30 degrees (nd > 0): n = ior, omcos = 1 - normal . refract with the flipped normal and the raw vector (:190)
    d = (math.sin(math.radians(30)), math.cos(math.radians(30)), 0.0)
    normal, refr, F = R.transparent_terms(1.5, (0.0, 1.0, 0.0), d)
    assert normal == (-0.0, -1.0, -0.0)
    sin2 = 1.5 * 1.5 * (1.0 - d[1] * d[1])
    omcos = 1.0 - (-(d[1] * 1.5 - (-1.0) * (1.5 * d[1] + math.sqrt(1.0 - sin2))))
    np.testing.assert_allclose(F, min(1.0, r0 + (1 - r0) * omcos ** 5), rtol=1e-13)
    assert omcos > 1.0 and F == 1.0      # normal . refract < 0 upstream: the factor saturates, no refraction child (:214)
    # total internal reflection: exiting at 60 degrees, sin2 = 2.25 * 0.75 >= 1
    d = (math.sin(math.radians(60)), math.cos(math.radians(60)), 0.0)
    normal, refr, F = R.transparent_terms(1.5, (0.0, 1.0, 0.0), d)
    assert refr is None and F == 1.0


def test_transparent_entering_exiting_tir_against_restatement(orc):
    """glass ball over a floor, no lights and one light: rays enter, bounce inside (exit branch with F saturated, total internal
    reflection at grazing inner hits), children cut by depth and by the significance test; plus rays started inside"""
    mats = [GLASS, {"kind": "phong", "diffuse": (0.4, 0.5, 0.6), "specular": (0.0, 0.0, 0.0), "exponent": 1.0, "ambient": (0.03, 0.02, 0.01)}]
    objs = [("sphere", (0, 0, 0), 1.0, 0), ("plane", (0, -1.2, 0), (0, 1, 0), 1)]
    rays = [[0, 0, 5, 0, 0, -1], [0.5, 0.1, 5, 0, 0, -1], [0.9, 0.3, 5, -0.01, 0, -1], [0, 0, 0, 0, 0.6, 0.8], [0.2, 0.1, 0.0, 1, 0.05, 0.02],
            [0.99, 0, 5, 0, 0, -1], [0, 3, 0, 0.1, -1, 0.05]]
    for lights in ((), (("point", (3.0, 4.0, 5.0), (1.0, 0.9, 0.8)),)):
        desc, parts = build(objs, mats, lights, background=(0.3, 0.5, 0.9))
        for md in (1, 4, 8):
            got, want = both(orc, desc, parts, rays, max_depth=md)
            close(got, want)
    assert np.all(np.isfinite(got)) and got[:, 2].max() > 0.1


def test_transparent_single_interface_by_hand(orc):
    """a transparent PLANE seen head-on from the air side: F = r0 = 0.04; the reflection child sees the background behind the
    viewer, the refraction child goes straight on (refract = d*n - N*(n + 1) normalised = d) to the background:
    colour = specular * bg * F + bg * (1 - F)"""
    desc, parts = build([("plane", (0, 0, 0), (0, 0, 1), 0)], [GLASS], background=(0.25, 0.5, 0.75))
    got, want = both(orc, desc, parts, [[0.3, -0.2, 4, 0, 0, -1]])
    F = 0.04
    hand = [0.9 * 0.25 * F + 0.25 * (1 - F), 0.8 * 0.5 * F + 0.5 * (1 - F), 0.7 * 0.75 * F + 0.75 * (1 - F)]
    close(got[0], hand, rtol=1e-14)
    close(got, want)
    # from the glass side (nd > 0), 60 degrees: total internal reflection -> F = 1, only specular * background
    s60, c60 = math.sin(math.radians(60)), math.cos(math.radians(60))
    got, want = both(orc, desc, parts, [[0, 0, -3, s60, 0, c60]])
    close(got[0], [0.9 * 0.25, 0.8 * 0.5, 0.7 * 0.75], rtol=1e-14)
    close(got, want)


# ---- SkyboxBackground + Texture::sample (src/raytrace.rs:234-256, src/texture.rs:39-58) ----------------------------
def _faces(size=4):
    rng = np.random.default_rng(77)
    return [rng.integers(0, 256, size=(size, size + k, 3), dtype=np.uint8) for k in range(6)]   # six different widths


def test_skybox_face_choice_ties_and_bilinear(orc):
    faces = _faces()
    values, _ = orc.srgb_tables()
    desc, parts = build([], [], skybox=faces)
    e = 1e-9
    dirs = [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1),                      # face centres
            (1, 0.3, -0.4), (-1, 0.3, -0.4), (0.2, 1, 0.7), (0.2, -1, 0.7), (-0.6, 0.1, 1), (-0.6, 0.1, -1),
            (2.0, 1.999, -1.5), (0.001, -3.0, 2.999), (1 + e, 1, 0.5), (1, 1 + e, 0.5), (0.5, 1, 1 + e),     # near the edges
            (1, 1, 0), (1, 0, 1), (0, 1, 1), (1, 1, 1), (-1, 1, 0.2), (0, 0, 0), (3, -3, 3),              # ties -> BLACK (strict >)
            (0.9999, -1, 1e-12), (5e-4, 2e-4, -1e-3), (-7.0, 6.9, 6.8)]
    rays = [[0, 0, 0, *d] for d in dirs]
    got, want = both(orc, desc, parts, rays, bg=("skybox", faces, values))
    close(got, want, rtol=1e-15)
    ties = slice(17, 24)
    assert np.all(got[ties] == 0.0), "ties fall through every strict comparison to color::BLACK (src/raytrace.rs:254)"
    assert np.all(got[:17].sum(axis=1) > 0)
    # face centres by hand: sample(0.5, 0.5) on a 4-row face with widths 4..9 -> bilinear of the four middle texels, y first
    for k in range(6):
        f = faces[k]
        h, w = f.shape[:2]
        x, y = 0.5 * (w - 1), 0.5 * (h - 1)
        x0, y0 = int(x), int(y)
        xx, yy = x - x0, y - y0
        x1, y1 = min(x0 + 1, w - 1), min(y0 + 1, h - 1)
        lin = lambda xi, yi: values[f[yi, xi]]
        cx0 = lin(x0, y0) * (1 - yy) + lin(x0, y1) * yy
        cx1 = lin(x1, y0) * (1 - yy) + lin(x1, y1) * yy
        np.testing.assert_allclose(got[k], cx0 * (1 - xx) + cx1 * xx, rtol=1e-15)


def test_skybox_uv_orientation(orc):
    """which way u and v run on every face (src/raytrace.rs:250-252): X faces u = -dz/dx, v = -dy/|dx|; Y faces u = dx/|dy|,
    v = dz/dy; Z faces u = dx/dz, v = -dy/|dz|; row 0 of a face is v = 0.  One bright texel in a corner of each face."""
    values, _ = orc.srgb_tables()
    faces = []
    for k in range(6):
        f = np.zeros((3, 3, 3), dtype=np.uint8)
        f[0, 2] = (255, 255, 255)     # top-right corner: u = 1, v = 0
        faces.append(f)
    desc, parts = build([], [], skybox=faces)
    c = 0.999999
    corner = {0: (1, c, -c), 1: (-1, c, c), 2: (c, 1, -c), 3: (c, -1, c), 4: (c, c, 1), 5: (-c, c, -1)}   # u -> 1, v -> 0 on face k
    rays = [[0, 0, 0, *corner[k]] for k in range(6)]
    got, want = both(orc, desc, parts, rays, bg=("skybox", faces, values))
    close(got, want, rtol=1e-15)
    assert np.all(got > 0.99), got
    dom = {0: 0, 1: 0, 2: 1, 3: 1, 4: 2, 5: 2}
    opposite = [[0, 0, 0, *[v if a == dom[k] else -v for a, v in enumerate(corner[k])]] for k in range(6)]   # u -> 0, v -> 1: a dark texel
    got2, want2 = both(orc, desc, parts, opposite, bg=("skybox", faces, values))
    close(got2, want2, rtol=1e-15)
    assert np.all(got2 < 0.01)


def test_texture_sample_clamps_and_truncates():
    """sample() clamps to [0,1] before scaling by (w-1), (h-1); `as u32` truncates; the neighbour index clamps at the edge"""
    tex = np.arange(2 * 3 * 3, dtype=np.uint8).reshape(2, 3, 3) * 14
    values = np.linspace(0.0, 1.0, 256) ** 2.2
    at = lambda x, y: values[tex[y, x]]
    np.testing.assert_array_equal(R.texture_sample(tex, values, -3.0, 9.0), at(0, 1))
    np.testing.assert_array_equal(R.texture_sample(tex, values, 1.0, 0.0), at(2, 0))
    np.testing.assert_allclose(R.texture_sample(tex, values, 0.25, 0.0), at(0, 0) * 0.5 + at(1, 0) * 0.5, rtol=1e-15)
    np.testing.assert_allclose(R.texture_sample(tex, values, 0.75, 1.0), at(1, 1) * 0.5 + at(2, 1) * 0.5, rtol=1e-15)


# ---- AreaLight draw order (src/scene.rs:151-155) -------------------------------------------------------------------
def test_area_light_draws_side1_then_side2(orc):
    """origin + side1 * (first draw) + side2 * (second draw): an L-shaped pair of sides makes the two orders give different
    light positions, hence different diffuse terms; three lights in a row consume six draws in order"""
    mats = [{"kind": "phong", "diffuse": (0.8, 0.7, 0.6), "specular": (0.3, 0.3, 0.3), "exponent": 8.0, "ambient": (0.0, 0.0, 0.0)}]
    lights = [("area", (-2.0, 5.0, -1.0), (8.0, 0.0, 0.0), (0.0, 0.0, 3.0), (1.0, 0.9, 0.8)),
              ("area", (4.0, 2.0, 4.0), (0.0, 6.0, 0.0), (-5.0, 0.0, 0.0), (0.2, 0.3, 0.9)),
              ("point", (0.0, 9.0, 0.0), (0.1, 0.1, 0.1)),
              ("area", (-3.0, 1.0, 2.0), (0.0, 0.0, -4.0), (1.0, 7.0, 0.0), (0.6, 0.1, 0.1))]
    desc, parts = build([("plane", (0, 0, 0), (0, 1, 0), 0)], mats, lights, background=(0, 0, 0))
    rays = [[0.5, 4, 3, 0.1, -1, -0.2], [-2, 6, 1, 0.3, -1, 0.1]]
    got, want = both(orc, desc, parts, rays)
    close(got, want)
    # by hand for the first light of the first ray: the first two numbers of the stream
    rng = R.XorShift(*STATE)
    u1, u2 = rng.f64(), rng.f64()
    assert 0 <= u1 < 1 and 0 <= u2 < 1 and abs(u1 - u2) > 0.05
    loc = (-2.0 + 8.0 * u1, 5.0, -1.0 + 3.0 * u2)
    pt = R.add((0.5, 4.0, 3.0), R.mul((0.1, -1.0, -0.2), 4.0))
    ldir = R.normalize(R.sub(loc, pt))
    swapped = R.normalize(R.sub((-2.0 + 8.0 * u2, 5.0, -1.0 + 3.0 * u1), pt))
    assert abs(ldir[1] - swapped[1]) > 1e-3
    only_first, _ = both(orc, *build([("plane", (0, 0, 0), (0, 1, 0), 0)], [dict(mats[0], specular=(0.0, 0.0, 0.0))], lights[:1], background=(0, 0, 0)),
                         rays[:1])
    hand = np.array([0.8, 0.7, 0.6]) * np.array([1.0, 0.9, 0.8]) * max(0.0, ldir[1]) * R.FRAC_1_PI
    close(only_first[0], hand, rtol=1e-14)


def test_area_light_counter_rng_dimensions(orc):
    """under the counter RNG light i draws dimensions 2+2i (side1) and 3+2i (side2) of the node's key"""
    mats = [{"kind": "phong", "diffuse": (0.8, 0.7, 0.6), "specular": (0.0, 0.0, 0.0), "exponent": 1.0, "ambient": (0.0, 0.0, 0.0)}]
    lights = [("point", (0.0, 9.0, 0.0), (0.1, 0.1, 0.1)), ("area", (-2.0, 5.0, -1.0), (8.0, 0.0, 0.0), (0.0, 0.0, 3.0), (1.0, 0.9, 0.8))]
    desc, parts = build([("plane", (0, 0, 0), (0, 1, 0), 0)], mats, lights, background=(0, 0, 0))
    key = 0x0123456789ABCDEF
    L = orc.lib()
    u = lambda dim: float(L.orc_draw_bits(key, dim) >> 11) * 2.0 ** -53
    got, _ = orc.OracleScene(desc).ray_color([[0.5, 4, 3, 0.1, -1, -0.2]], keys=[key])
    ro, rm, rl = parts
    sc = R.Scene(ro, rm, rl, ("solid", (0.0, 0.0, 0.0)))
    want = R.ray_color(sc, (0.5, 4.0, 3.0), (0.1, -1.0, -0.2), 1.0, 0, R.FixedDraws([u(4), u(5)]))   # light index 1 -> dims 4, 5
    close(got[0], want)


# ---- DepthOfFieldCamera (src/camera.rs:97-122) ---------------------------------------------------------------------
def test_dof_camera_draw_order_and_geometry(orc):
    """theta = first draw * 2 pi, radius = sqrt(Closed01 second draw) * aperture; origin on the image plane displaced along the
    camera's u, v; every ray of one image position passes through the same focal point"""
    look, up = (0.1, -0.2, -1.0), (0.0, 1.0, 0.0)
    pos0 = (1.0, 2.0, 9.0)
    cam = S.DepthOfFieldCamera.new(S.SimplePerspectiveCamera.new(pos0, look, up, 2.5), 12.0, 0.4, 3)
    desc, _ = build([], [], camera=cam)
    osc = orc.OracleScene(desc)
    P = [(0.0, 0.0), (0.3, -0.2), (-0.9, 0.7), (0.3, -0.2), (0.3, -0.2)]
    got = osc.project_sequential(P, STATE)
    m = R.spc_matrix(look, up, 2.5)
    _, om = osc.camera()
    np.testing.assert_allclose(np.array(m), om.reshape(3, 3), rtol=1e-15, atol=1e-17)
    rng = R.XorShift(*STATE)
    for i, p in enumerate(P):
        o, d = R.dof_project(pos0, m, 12.0, 0.4, p, rng)
        np.testing.assert_allclose(got[i, :3], o, rtol=1e-15)
        np.testing.assert_allclose(got[i, 3:], d, rtol=1e-14)
    assert rng.draws == 2 * len(P)
    # by hand for the first ray: the lens offset has length sqrt(r2) * aperture * |u| (u, v are unit vectors)
    rng = R.XorShift(*STATE)
    theta, r2 = rng.f64() * 2 * math.pi, rng.closed01()
    ip = R.add(pos0, R.mat_vec(m, (0.0, 0.0, 1.0)))
    off = R.sub(tuple(got[0, :3]), ip)
    np.testing.assert_allclose(math.sqrt(R.dot(off, off)), math.sqrt(r2) * 0.4, rtol=1e-12)
    u = (m[0][0], m[1][0], m[2][0])
    np.testing.assert_allclose(R.dot(off, u), math.cos(theta) * math.sqrt(r2) * 0.4, rtol=1e-12)
    # the three rays of image position (0.3, -0.2) meet at its focal point: position + dir * focus / |w|
    dirv = R.mat_vec(m, (0.3, -0.2, 1.0))
    fp = R.add(pos0, R.mul(dirv, 12.0 / 2.5))
    for i in (1, 3, 4):
        o, d = got[i, :3], got[i, 3:]
        t = np.dot(np.array(fp) - o, d)
        np.testing.assert_allclose(o + d * t, fp, rtol=1e-12, atol=1e-12)
    assert len({tuple(got[i, :3]) for i in (1, 3, 4)}) == 3


def test_restatement_agrees_with_oracle_on_a_mixed_scene(orc):
    """everything at once (Phong + Fresnel + Transparent, plane + spheres, point + directional + area lights, depth 6):
    40 rays, every Scene::intersect call and every draw in the same order in both restatements"""
    mats = [{"kind": "phong", "diffuse": (0.7, 0.6, 0.5), "specular": (0.2, 0.2, 0.2), "exponent": 12.0, "ambient": (0.02, 0.02, 0.02)}, FRES, GLASS]
    objs = [("plane", (0, 0, 0), (0, 1, 0), 0), ("sphere", (-1.5, 1.0, 0), 1.0, 1), ("sphere", (1.2, 0.8, 0.5), 0.8, 2), ("sphere", (0, 0.4, 2.0), 0.4, 0)]
    lights = [("point", (4.0, 6.0, 5.0), (0.9, 0.9, 0.8)), ("directional", (0.2, -1.0, -0.3), (0.3, 0.3, 0.4)),
              ("area", (-4.0, 5.0, -2.0), (2.0, 0.0, 0.0), (0.0, 0.0, 2.0), (0.5, 0.4, 0.3))]
    desc, parts = build(objs, mats, lights, background=(0.2, 0.3, 0.5))
    g = np.random.default_rng(5)
    o = np.stack([g.uniform(-3, 3, 40), g.uniform(0.5, 4, 40), g.uniform(4, 7, 40)], 1)
    tgt = np.stack([g.uniform(-2, 2, 40), g.uniform(0, 1.5, 40), g.uniform(-0.5, 2.2, 40)], 1)
    rays = np.concatenate([o, tgt - o], 1)
    got, want = both(orc, desc, parts, rays, max_depth=6)
    close(got, want, rtol=1e-12)
