"""Independent restatement, in plain Python floats (IEEE f64, same operation order), of the parts of the reference the
KATs in tests/test_oracle_kats_materials.py pin -- written from the cited Rust lines, NOT from oracle/oracle.cpp:

    Scene::intersect            /root/reference/src/scene.rs:223-249, src/shapes.rs:50-112
    Phong/Fresnel/Transparent   /root/reference/src/raytrace.rs:30-67, :123-167, :169-226
    lights                      /root/reference/src/scene.rs:122-155
    backgrounds + Texture       /root/reference/src/raytrace.rs:228-256, src/texture.rs:39-58, src/color.rs:611-613
    DepthOfFieldCamera          /root/reference/src/camera.rs:97-122
    XorShiftRng + gen::<f64>    rand 0.3 (published algorithm; SURVEY.md 8c states the assumption)

Test infrastructure only.  Pure-Python recursion: meant for a handful of rays.  Random numbers come from an explicit
`rng` object whose draws are consumed in the reference's order (`rng.f64()`, `rng.closed01()`), so a test can hand both
this file and the oracle the same xorshift128 stream and compare results draw for draw.
"""
import math

MIN_SIGNIFICANCE = 1.0 / 256.0 / 2.0   # src/raytrace.rs:17
MAX_DEPTH = 4                          # src/raytrace.rs:18
FRAC_1_PI = 0.318309886183790671537767526745028724
M32 = 0xFFFFFFFF


class XorShift:
    """rand 0.3 XorShiftRng: t = x ^ (x << 11); x, y, z = y, z, w; w = w ^ (w >> 19) ^ (t ^ (t >> 8))"""

    def __init__(self, x, y, z, w):
        self.s = [x & M32, y & M32, z & M32, w & M32]
        self.draws = 0

    def next_u32(self):
        x, y, z, w = self.s
        t = (x ^ (x << 11)) & M32
        w2 = (w ^ (w >> 19) ^ (t ^ (t >> 8))) & M32
        self.s = [y, z, w, w2]
        return w2

    def next_u64(self):
        hi = self.next_u32()
        lo = self.next_u32()
        return (hi << 32) | lo

    def f64(self):          # gen::<f64>() in [0, 1): 53 bits
        self.draws += 1
        return float(self.next_u64() >> 11) * 2.0 ** -53

    def closed01(self):     # Closed01<f64>: [0, 1)-value * 2^53 / (2^53 - 1)
        self.draws += 1
        return float(self.next_u64() >> 11) / 9007199254740991.0


class FixedDraws:
    """hands out a prepared list of numbers (for tests that choose the draws themselves)"""

    def __init__(self, values):
        self.v = list(values)
        self.draws = 0

    def f64(self):
        self.draws += 1
        return self.v.pop(0)

    closed01 = f64


# ---- vectors -----------------------------------------------------------------------------------------------
def add(a, b): return (a[0] + b[0], a[1] + b[1], a[2] + b[2])
def sub(a, b): return (a[0] - b[0], a[1] - b[1], a[2] - b[2])
def mul(a, s): return (a[0] * s, a[1] * s, a[2] * s)
def neg(a): return (-a[0], -a[1], -a[2])
def dot(a, b): return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]
def cmul(a, b): return (a[0] * b[0], a[1] * b[1], a[2] * b[2])


def normalize(a):
    n = math.sqrt(dot(a, a))
    if n == 0.0:
        return (math.nan, math.nan, math.nan)
    return (a[0] / n, a[1] / n, a[2] / n)


def clamp_zero(x): return 0.0 if x < 0.0 else x
def clamp_one(x): return 1.0 if x > 1.0 else x
def significance(c): return c[0] + c[1] + c[2]


def powf(x, e):
    if x == 0.0:
        return 1.0 if e == 0.0 else 0.0
    return math.pow(x, e)


# ---- shapes (src/shapes.rs:50-89, :100-112) ---------------------------------------------------------------
def sphere_intersect(center, radius, o, d):
    oc = sub(o, center)
    a = dot(d, d)
    b = 2.0 * dot(d, oc)
    c = dot(oc, oc) - radius * radius
    disc = b * b - 4.0 * a * c
    if disc > 0.0:
        dsqrt = math.sqrt(disc)
        t = (-b - dsqrt) / (2.0 * a)
        if not (t > 0.0):
            t = (-b + dsqrt) / (2.0 * a)
            if not (t > 0.0):
                return None
        return t, normalize(sub(add(o, mul(d, t)), center))
    return None


def plane_intersect(point, normal, o, d):
    den = dot(normal, d)
    num = dot(normal, sub(point, o))
    t = num / den if den != 0.0 else (math.nan if num == 0.0 else math.copysign(math.inf, num) * math.copysign(1.0, den))
    if t <= 0.0:
        return None
    return t, normal


class Scene:
    """objects: list of ("sphere", center, radius, material) | ("plane", point, normal, material); materials: dicts with
    kind/diffuse/specular/ambient/exponent/ior; lights: ("point", loc, color) | ("directional", dir, color) |
    ("area", origin, side1, side2, color); background: ("solid", rgb) | ("skybox", [six (h, w, 3) uint8 arrays], srgb_values)"""

    def __init__(self, objects, materials, lights, background, max_depth=MAX_DEPTH, min_significance=MIN_SIGNIFICANCE):
        self.objects, self.materials, self.lights, self.background = objects, materials, lights, background
        self.max_depth, self.min_significance = max_depth, min_significance
        self.intersect_calls = 0

    # src/scene.rs:247-249: min_by_key over Option<FloatNotNan(t)>: NaN (None) first, then smallest t, ties -> first index
    def intersect(self, o, d):
        self.intersect_calls += 1
        best = None
        for i, ob in enumerate(self.objects):
            r = sphere_intersect(ob[1], ob[2], o, d) if ob[0] == "sphere" else plane_intersect(ob[1], ob[2], o, d)
            if r is None:
                continue
            t, n = r
            if best is None:
                best = (i, t, n)
                continue
            bt = best[1]
            if bt != bt:
                continue                      # a NaN already holds the minimum; later NaNs tie and lose to the first
            if t != t or t < bt:
                best = (i, t, n)
        return best


# ---- lights (src/scene.rs:122-155) ------------------------------------------------------------------------
def light_dir_and_sq_range(light, pt, rng):
    kind = light[0]
    if kind == "directional":
        return neg(light[1]), None
    if kind == "area":
        u1 = rng.f64()                        # side1's factor is drawn first (src/scene.rs:153, left to right)
        u2 = rng.f64()
        loc = add(add(light[1], mul(light[2], u1)), mul(light[3], u2))
    else:
        loc = light[1]
    diff = (loc[0] - pt[0], loc[1] - pt[1], loc[2] - pt[2])
    return normalize(diff), dot(sub(loc, pt), sub(loc, pt))


def _shadowed(scene, pt, ldir, sqrange):
    hit = scene.intersect(add(pt, mul(ldir, 0.00001)), ldir)
    if hit is None:
        return False
    return True if sqrange is None else hit[1] * hit[1] < sqrange


# ---- texture + backgrounds (src/texture.rs:39-58, src/raytrace.rs:228-256) ------------------------------
def _clamp01(x): return 0.0 if x < 0.0 else (1.0 if x > 1.0 else x)


def texture_sample(tex, values, xs, ys):
    h, w = tex.shape[0], tex.shape[1]

    def at(x, y):
        p = tex[y, x]
        return (float(values[p[0]]), float(values[p[1]]), float(values[p[2]]))
    x = _clamp01(xs) * float(w - 1)
    y = _clamp01(ys) * float(h - 1)
    x0 = int(x)
    x1 = w - 1 if x0 >= w - 1 else x0 + 1
    xx = x - float(x0)
    y0 = int(y)
    y1 = h - 1 if y0 >= h - 1 else y0 + 1
    yy = y - float(y0)
    cx0 = add(mul(at(x0, y0), 1.0 - yy), mul(at(x0, y1), yy))      # y first ...
    cx1 = add(mul(at(x1, y0), 1.0 - yy), mul(at(x1, y1), yy))
    return add(mul(cx0, 1.0 - xx), mul(cx1, xx))                    # ... then x


def background_color(scene, d):
    bg = scene.background
    if bg[0] == "solid":
        return bg[1]
    faces, values = bg[1], bg[2]       # px nx py ny pz nz
    ax, ay, az = abs(d[0]), abs(d[1]), abs(d[2])
    if ax > az and ax > ay:
        px, py = -d[2] / d[0], -d[1] / ax
        return texture_sample(faces[0 if d[0] > 0.0 else 1], values, px * 0.5 + 0.5, py * 0.5 + 0.5)
    if ay > ax and ay > az:
        px, py = d[0] / ay, d[2] / d[1]
        return texture_sample(faces[2 if d[1] > 0.0 else 3], values, px * 0.5 + 0.5, py * 0.5 + 0.5)
    if az > ax and az > ay:
        px, py = d[0] / d[2], -d[1] / az
        return texture_sample(faces[4 if d[2] > 0.0 else 5], values, px * 0.5 + 0.5, py * 0.5 + 0.5)
    return (0.0, 0.0, 0.0)


# ---- materials (src/raytrace.rs:30-67, :123-167, :169-226) ----------------------------------------------
def ray_color(scene, o, d, sig, depth, rng):
    hit = scene.intersect(o, d)
    if hit is None:
        return background_color(scene, d)
    i, t, rn = hit
    m = scene.materials[scene.objects[i][3]]
    return {"phong": _phong, "fresnel": _fresnel, "transparent": _transparent}[m["kind"]](scene, m, t, rn, o, d, sig, depth, rng)


def _phong(scene, m, t, rn, o, d, sig, depth, rng):
    res = m["ambient"]
    if depth > scene.max_depth:
        return res
    pt = add(o, mul(d, t))
    diffuse = significance(m["diffuse"]) * sig > scene.min_significance
    specular = significance(m["specular"]) * sig > scene.min_significance
    normal = neg(rn) if dot(rn, d) > 0.0 else rn
    for light in scene.lights:
        if diffuse or specular:
            ldir, sq = light_dir_and_sq_range(light, pt, rng)
            if _shadowed(scene, pt, ldir, sq):
                continue
            lcol = light[-1]
            if diffuse:
                res = add(res, mul(mul(cmul(m["diffuse"], lcol), clamp_zero(dot(ldir, normal))), FRAC_1_PI))
            if specular:
                res = add(res, mul(cmul(m["specular"], lcol), powf(clamp_zero(dot(normal, normalize(sub(ldir, d)))), m["exponent"])))
    if specular:
        rd = sub(d, mul(normal, 2.0 * dot(d, normal)))
        child = ray_color(scene, add(pt, mul(rd, 0.00001)), rd, sig * significance(m["specular"]), depth + 1, rng)
        res = add(res, cmul(m["specular"], child))
    return res


def schlick(ior, omcos):
    r0 = (ior - 1.0) / (ior + 1.0)
    r0 = r0 * r0
    omcos2 = omcos * omcos
    return clamp_one(r0 + (1.0 - r0) * omcos2 * omcos2 * omcos)


def _fresnel(scene, m, t, rn, o, d, sig, depth, rng):
    res = m["ambient"]
    if depth > scene.max_depth:
        return res
    pt = add(o, mul(d, t))
    nd = dot(rn, d)
    normal = neg(rn) if nd > 0.0 else rn
    fresnel = schlick(m["ior"], 1.0 - abs(nd))
    diffuse = significance(m["diffuse"]) * sig > scene.min_significance
    specular = significance(m["specular"]) * fresnel * sig > scene.min_significance
    for light in scene.lights:
        if diffuse or specular:
            ldir, sq = light_dir_and_sq_range(light, pt, rng)
            if _shadowed(scene, pt, ldir, sq):
                continue
            lcol = light[-1]
            if diffuse:
                res = add(res, mul(mul(cmul(m["diffuse"], lcol), clamp_zero(dot(ldir, normal))), FRAC_1_PI))
            if specular:
                res = add(res, mul(mul(cmul(m["specular"], lcol), fresnel), powf(clamp_zero(dot(normal, normalize(sub(ldir, d)))), m["exponent"])))
    if specular:
        rd = sub(d, mul(normal, 2.0 * dot(d, normal)))
        child = ray_color(scene, add(pt, mul(rd, 0.00001)), rd, fresnel * sig * significance(m["specular"]), depth + 1, rng)
        res = add(res, mul(cmul(m["specular"], child), fresnel))
    return res


def transparent_terms(ior, rn, d):
    """(normal, refract or None, fresnel) of src/raytrace.rs:174-192"""
    nd = dot(rn, d)
    normal = neg(rn) if nd > 0.0 else rn
    n = ior if nd > 0.0 else 1.0 / ior
    sin2 = n * n * (1.0 - nd * nd)
    refract = None
    if sin2 < 1.0:
        cos = math.sqrt(1.0 - sin2)
        refract = sub(mul(d, n), mul(normal, n * abs(nd) + cos))
    if nd > 0.0:
        omcos = 1.0 - dot(normal, refract) if refract is not None else 0.0    # the un-normalised vector, :190
    else:
        omcos = 1.0 - abs(nd)
    fresnel = schlick(ior, omcos) if refract is not None else 1.0
    return normal, refract, fresnel


def _transparent(scene, m, t, rn, o, d, sig, depth, rng):
    res = (0.0, 0.0, 0.0)
    if depth > scene.max_depth:
        return res
    pt = add(o, mul(d, t))
    normal, refract, fresnel = transparent_terms(m["ior"], rn, d)
    ndv = dot(normal, d)
    specular = significance(m["specular"]) * fresnel * sig > scene.min_significance
    for light in scene.lights:
        if specular:
            ldir, sq = light_dir_and_sq_range(light, pt, rng)
            if _shadowed(scene, pt, ldir, sq):
                continue
            res = add(res, mul(mul(cmul(m["specular"], light[-1]), fresnel), powf(clamp_zero(dot(normal, normalize(sub(ldir, d)))), m["exponent"])))
    if specular:
        rd = sub(d, mul(normal, 2.0 * ndv))
        child = ray_color(scene, add(pt, mul(rd, 0.00001)), rd, fresnel * sig * significance(m["specular"]), depth + 1, rng)
        res = add(res, mul(cmul(m["specular"], child), fresnel))
    if fresnel < 1.0 and refract is not None:
        omf = clamp_one(1.0 - fresnel)
        rf = normalize(refract)
        child = ray_color(scene, add(pt, mul(rf, 0.00001)), rf, omf * sig, depth + 1, rng)
        res = add(res, mul(child, omf))
    return res


# ---- DepthOfFieldCamera (src/camera.rs:51-63, :97-122) ---------------------------------------------------
def cross(a, b): return (a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0])


def spc_matrix(look, up, im_dist):
    """columns u, v, w of SimplePerspectiveCamera::new, as rows of a row-major Mat3 (src/camera.rs:51-63)"""
    u = normalize(cross(look, up))
    v = normalize(cross(u, look))
    w = mul(normalize(look), im_dist)
    return ((u[0], v[0], w[0]), (u[1], v[1], w[1]), (u[2], v[2], w[2]))


def mat_vec(m, v):
    return (m[0][0] * v[0] + m[0][1] * v[1] + m[0][2] * v[2], m[1][0] * v[0] + m[1][1] * v[1] + m[1][2] * v[2],
            m[2][0] * v[0] + m[2][1] * v[1] + m[2][2] * v[2])


def dof_project(position, m, focus, aperture, pos, rng):
    im = mat_vec(m, (0.0, 0.0, 1.0))
    im_dist = math.sqrt(dot(im, im))
    dirv = mat_vec(m, (pos[0], pos[1], 1.0))
    ip = add(position, dirv)
    fp = add(position, mul(dirv, focus / im_dist))
    theta = rng.f64() * (2.0 * math.pi)          # the angle is drawn first
    r2 = rng.closed01()
    r = math.sqrt(r2) * aperture
    orig = add(ip, mat_vec(m, (math.cos(theta) * r, math.sin(theta) * r, 0.0)))
    return orig, normalize(sub(fp, orig))
