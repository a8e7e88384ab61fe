// oracle.cpp -- CPU restatement of rust-raytrace's per-pixel tracing path.  TEST INFRASTRUCTURE ONLY.
//
// This file is the *checker* for the CUDA product in rust-raytrace_b200/.  Nothing under
// rust-raytrace_b200/ may include, link or call it; only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py load liboracle.so.
//
// What it restates (reference file:line, all relative to the upstream crate):
//   frame loop ................ src/main.rs:39-59
//   raytrace / ray_color ...... src/raytrace.rs:261-276
//   four materials ............ src/raytrace.rs:30-226
//   two backgrounds ........... src/raytrace.rs:228-256
//   lights .................... src/scene.rs:117-155
//   Scene::intersect .......... src/scene.rs:223-249   (closest hit, NaN-first, first-index ties)
//   Sphere / Plane ............ src/shapes.rs:18-25, 50-89, 100-112
//   cameras ................... src/camera.rs:51-79, 97-122
//   colour ops + quantiser .... src/color.rs:27-67, 593-639   (tables regenerated, see tools/gen_srgb_tables.py)
//   bilinear texture .......... src/texture.rs:26-58
//   BMP header ................ src/bmp.rs:10-61
//
// Arithmetic is f64 in the reference's exact operation order; build with -ffp-contract=off so that
// gcc never fuses a*b+c (Rust/x86-64 never does).
//
// PARITY UNPINNED: the upstream crate cannot be built here (no rustc/cargo, crates
// nalgebra ^0.4 / rand ^0.3 / image ^0.6 are not vendored, Cargo.lock is git-ignored) and ships no
// tests.  The vector primitives below restate nalgebra's published definitions (dot = x*x'+y*y'+z*z'
// left to right; normalize = componentwise division by sqrt(dot(v,v)); Mat3*Vec3 row-wise starting
// from zero; standard cross).  The only upstream artefact that pins behaviour is out.bmp, used as a
// statistical golden (tests/test_oracle_outbmp.py).
//
// RNG: the reference threads one OS-seeded XorShiftRng through the whole frame (src/main.rs:43), so no
// run of it is reproducible.  Two modes are provided:
//   ORC_RNG_SEQUENTIAL  one xorshift128 stream consumed in the reference's call order (single thread);
//   ORC_RNG_COUNTER     the counter-based definition of DESIGN.md section "RNG", shared with the GPU.

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/rt_srgb_tables.inc"

namespace {

// ---------------------------------------------------------------------------------------------
// value types (src/types.rs:10-25, src/color.rs:15-67)
// ---------------------------------------------------------------------------------------------
struct V3 {
    double x, y, z;
};
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator/(V3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }
inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline double sqnorm(V3 a) { return dot(a, a); }
inline double norm(V3 a) { return std::sqrt(sqnorm(a)); }
inline V3 normalize(V3 a) { return a / norm(a); }
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

struct Color {
    double r, g, b;
};
const Color BLACK = {0.0, 0.0, 0.0};
inline Color operator+(Color a, Color b) { return {a.r + b.r, a.g + b.g, a.b + b.b}; }
inline Color operator*(Color a, Color b) { return {a.r * b.r, a.g * b.g, a.b * b.b}; }
inline Color operator*(Color a, double s) { return {a.r * s, a.g * s, a.b * s}; }
inline Color operator/(Color a, double s) { return {a.r / s, a.g / s, a.b / s}; }
inline double significance(Color c) { return c.r + c.g + c.b; }  // src/color.rs:637-639

const double SRGB_VALUES[256] = RT_SRGB_VALUES_INIT;
const double SRGB_AVERAGE[255] = RT_SRGB_AVERAGE_INIT;

// src/color.rs:593-600 -- smallest i with val < SRGB_AVERAGE[i], else 255 (NaN -> 255, negative -> 0)
inline uint8_t to_srgb(double val) {
    for (int i = 0; i < 255; i++)
        if (val < SRGB_AVERAGE[i]) return (uint8_t)i;
    return 255;
}

// ---------------------------------------------------------------------------------------------
// RNG
// ---------------------------------------------------------------------------------------------
enum { ORC_RNG_COUNTER = 0, ORC_RNG_SEQUENTIAL = 1 };

inline uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
const uint64_t GOLD = 0x9E3779B97F4A7C15ULL;
inline uint64_t key_sample(uint64_t seed, uint64_t pixel, uint64_t aa) {
    uint64_t k = mix64(seed ^ 0x243F6A8885A308D3ULL);
    k = mix64(k + GOLD * (pixel + 1));
    return mix64(k + GOLD * (aa + 1));
}
inline uint64_t child_key(uint64_t k, uint64_t code) { return mix64(k + GOLD * (code + 1)); }
inline uint64_t draw_bits(uint64_t k, uint64_t dim) { return mix64((k ^ 0xD6E8FEB86659FD93ULL) + GOLD * (dim + 1)); }
const uint64_t CODE_REFLECT = 1, CODE_REFRACT = 2, CODE_INDIRECT0 = 3, CODE_CAMERA0 = 1ULL << 32;

struct Rng {
    int mode;
    uint32_t x, y, z, w;  // xorshift128 state (rand 0.3 XorShiftRng), sequential mode only
    uint32_t next_u32() {
        uint32_t t = x ^ (x << 11);
        x = y; y = z; z = w;
        w = w ^ (w >> 19) ^ (t ^ (t >> 8));
        return w;
    }
    uint64_t next_u64() {
        uint64_t hi = next_u32();
        uint64_t lo = next_u32();
        return (hi << 32) | lo;
    }
    uint64_t bits(uint64_t key, uint64_t dim) { return mode == ORC_RNG_SEQUENTIAL ? next_u64() : draw_bits(key, dim); }
    // gen::<f64>() in [0,1): 53 random bits
    double f64(uint64_t key, uint64_t dim) { return (double)(bits(key, dim) >> 11) * 0x1.0p-53; }
    // Closed01<f64> in [0,1]
    double closed01(uint64_t key, uint64_t dim) { return (double)(bits(key, dim) >> 11) / 9007199254740991.0; }
};

// ---------------------------------------------------------------------------------------------
// scene model (src/scene.rs:32-212, src/shapes.rs:43-98, src/camera.rs:31-106, src/texture.rs:20-24)
// ---------------------------------------------------------------------------------------------
enum { SHAPE_SPHERE = 0, SHAPE_PLANE = 1 };
enum { MAT_PHONG = 0, MAT_INDIRECT = 1, MAT_FRESNEL = 2, MAT_TRANSPARENT = 3 };
enum { LIGHT_POINT = 0, LIGHT_DIRECTIONAL = 1, LIGHT_AREA = 2 };
enum { CAM_SPC = 0, CAM_DOF = 1 };
enum { BG_SOLID = 0, BG_SKYBOX = 1 };

struct Object {
    int kind;
    V3 p;           // sphere centre | plane point
    V3 n;           // plane normal
    double radius;  // sphere
    uint32_t material;
};
struct Material {
    int kind;
    Color diffuse, specular, ambient;
    double exponent, ior;
    uint32_t samples;
};
struct Light {
    int kind;
    V3 a, s1, s2;  // location | direction | origin, side1, side2
    Color color;
};
struct Camera {
    int kind = CAM_SPC;
    V3 position{0, 0, 0};
    double m[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};  // row-major Mat3
    double focus = 0, aperture = 0, im_dist = 0;
    uint32_t samples = 1;
};
struct Texture {
    uint32_t width = 0, height = 0;
    std::vector<uint8_t> data;
};
struct Scene {
    std::vector<Object> objects;
    std::vector<Material> materials;
    std::vector<Light> lights;
    Camera camera;
    int bg_kind = BG_SOLID;
    Color bg_color = BLACK;
    Texture faces[6];  // px nx py ny pz nz
    uint32_t width = 0, height = 0, antialias = 1;
};

struct Ray {
    V3 origin, direction;
};
inline V3 cast(const Ray& r, double t) { return r.origin + r.direction * t; }  // src/shapes.rs:22-24

struct Hit {
    double t;
    V3 normal;
};

// src/shapes.rs:50-89
inline bool sphere_intersect(const Object& s, const Ray& ray, Hit* out) {
    V3 ominusc = ray.origin - s.p;
    double a = sqnorm(ray.direction);
    double b = 2.0 * dot(ray.direction, ominusc);
    double c = sqnorm(ominusc) - s.radius * s.radius;
    double discriminant = b * b - 4.0 * a * c;
    if (discriminant > 0.0) {
        double dsqrt = std::sqrt(discriminant);
        double t = (-b - dsqrt) / (2.0 * a);
        if (t > 0.0) {
            out->t = t;
            out->normal = normalize(cast(ray, t) - s.p);
            return true;
        }
        double t2 = (-b + dsqrt) / (2.0 * a);
        if (t2 > 0.0) {
            out->t = t2;
            out->normal = normalize(cast(ray, t2) - s.p);
            return true;
        }
    }
    return false;
}

// src/shapes.rs:100-112 -- note `t <= 0` rejects, so +inf and NaN are hits
inline bool plane_intersect(const Object& p, const Ray& ray, Hit* out) {
    double t = dot(p.n, p.p - ray.origin) / dot(p.n, ray.direction);
    if (t <= 0.0) return false;
    out->t = t;
    out->normal = p.n;
    return true;
}

struct Stats {
    uint64_t rays = 0;         // Scene::intersect calls (closest + shadow)
    uint64_t shadow_rays = 0;  // of which shadow queries
    uint64_t shape_tests = 0;  // Shape::intersect calls
};

// src/scene.rs:247-249: filter_map(...).min_by_key(|o| FloatNotNan::new(t)).  Key type is
// Option<FloatNotNan>; None (NaN) < Some(_); min_by_key keeps the FIRST minimum.
inline int scene_intersect(const Scene& sc, const Ray& ray, Hit* best, Stats* st) {
    int best_i = -1;
    bool best_nan = false;
    st->rays++;
    const size_t n = sc.objects.size();
    st->shape_tests += n;
    for (size_t i = 0; i < n; i++) {
        const Object& o = sc.objects[i];
        Hit h{};
        bool ok = o.kind == SHAPE_SPHERE ? sphere_intersect(o, ray, &h) : plane_intersect(o, ray, &h);
        if (!ok) continue;
        bool h_nan = h.t != h.t;
        bool better;
        if (best_i < 0) better = true;
        else if (best_nan) better = false;          // nothing is strictly less than None
        else if (h_nan) better = true;              // None < Some
        else better = h.t < best->t;                // strict: ties keep the first
        if (better) { best_i = (int)i; *best = h; best_nan = h_nan; }
    }
    return best_i;
}

// ---------------------------------------------------------------------------------------------
// cameras (src/camera.rs)
// ---------------------------------------------------------------------------------------------
inline V3 mat_mul(const double* m, V3 v) {  // nalgebra Mat3 * Vec3: res_i = ((0 + m_i1 x) + m_i2 y) + m_i3 z
    return {m[0] * v.x + m[1] * v.y + m[2] * v.z, m[3] * v.x + m[4] * v.y + m[5] * v.z, m[6] * v.x + m[7] * v.y + m[8] * v.z};
}

// src/camera.rs:51-63
void camera_spc_new(Camera* cam, V3 position, V3 look, V3 up, double im_dist) {
    V3 u = normalize(cross(look, up));
    V3 v = normalize(cross(u, look));
    V3 w = normalize(look) * im_dist;
    cam->kind = CAM_SPC;
    cam->position = position;
    double m[9] = {u.x, v.x, w.x, u.y, v.y, w.y, u.z, v.z, w.z};
    std::memcpy(cam->m, m, sizeof m);
    cam->samples = 1;
}
// src/camera.rs:67-73
void camera_spc_look_at(Camera* cam, V3 focus, V3 look, V3 up, double pov, double h) {
    double cot = 1.0 / std::tan(pov / 2.0);
    double im_dist = cot;
    double d = h * cot;
    V3 position = focus - normalize(look) * d;
    camera_spc_new(cam, position, look, up, im_dist);
}
// src/camera.rs:97-106
void camera_make_dof(Camera* cam, double focus, double aperture, uint32_t samples) {
    cam->im_dist = norm(mat_mul(cam->m, V3{0.0, 0.0, 1.0}));
    cam->kind = CAM_DOF;
    cam->focus = focus;
    cam->aperture = aperture;
    cam->samples = samples;
}
// src/camera.rs:77-79, 110-121
inline Ray camera_project(const Camera& cam, double px, double py, Rng& rng, uint64_t key) {
    if (cam.kind == CAM_SPC) return Ray{cam.position, normalize(mat_mul(cam.m, V3{px, py, 1.0}))};
    V3 dir = mat_mul(cam.m, V3{px, py, 1.0});
    V3 ip = cam.position + dir;
    V3 fp = cam.position + dir * (cam.focus / cam.im_dist);
    double theta = rng.f64(key, 0) * (2.0 * M_PI);
    double r2 = rng.closed01(key, 1);
    double r = std::sqrt(r2) * cam.aperture;
    V3 orig = ip + mat_mul(cam.m, V3{std::cos(theta) * r, std::sin(theta) * r, 0.0});
    return Ray{orig, normalize(fp - orig)};
}

// ---------------------------------------------------------------------------------------------
// texture + backgrounds (src/texture.rs:26-58, src/raytrace.rs:228-256, src/color.rs:611-613)
// ---------------------------------------------------------------------------------------------
inline double clamp01(double x) { return x < 0.0 ? 0.0 : (x > 1.0 ? 1.0 : x); }
inline uint32_t f64_as_u32(double x) {  // Rust `as u32`: saturating, NaN -> 0
    if (!(x > 0.0)) return 0;
    if (x >= 4294967295.0) return 4294967295u;
    return (uint32_t)x;
}
inline Color tex_at(const Texture& t, uint32_t x, uint32_t y) {
    size_t idx = 3 * ((size_t)x + (size_t)y * t.width);
    return Color{SRGB_VALUES[t.data[idx]], SRGB_VALUES[t.data[idx + 1]], SRGB_VALUES[t.data[idx + 2]]};
}
inline Color tex_sample(const Texture& t, double xs, double ys) {
    double x = clamp01(xs) * (double)(t.width - 1);
    double y = clamp01(ys) * (double)(t.height - 1);
    uint32_t x0 = f64_as_u32(x);
    uint32_t x1 = x0 >= t.width - 1 ? t.width - 1 : x0 + 1;
    double xx = x - (double)x0;
    uint32_t y0 = f64_as_u32(y);
    uint32_t y1 = y0 >= t.height - 1 ? t.height - 1 : y0 + 1;
    double yy = y - (double)y0;
    Color cx0 = tex_at(t, x0, y0) * (1.0 - yy) + tex_at(t, x0, y1) * yy;
    Color cx1 = tex_at(t, x1, y0) * (1.0 - yy) + tex_at(t, x1, y1) * yy;
    return cx0 * (1.0 - xx) + cx1 * xx;
}
inline Color background_color(const Scene& sc, const Ray& ray) {
    if (sc.bg_kind == BG_SOLID) return sc.bg_color;
    V3 d = ray.direction;
    double ax = std::fabs(d.x), ay = std::fabs(d.y), az = std::fabs(d.z);
    if (ax > az && ax > ay) {
        double px = -d.z / d.x, py = -d.y / ax;
        return tex_sample(sc.faces[d.x > 0.0 ? 0 : 1], px * 0.5 + 0.5, py * 0.5 + 0.5);
    }
    if (ay > ax && ay > az) {
        double px = d.x / ay, py = d.z / d.y;
        return tex_sample(sc.faces[d.y > 0.0 ? 2 : 3], px * 0.5 + 0.5, py * 0.5 + 0.5);
    }
    if (az > ax && az > ay) {
        double px = d.x / d.z, py = -d.y / az;
        return tex_sample(sc.faces[d.z > 0.0 ? 4 : 5], px * 0.5 + 0.5, py * 0.5 + 0.5);
    }
    return BLACK;
}

// ---------------------------------------------------------------------------------------------
// lights (src/scene.rs:117-155)
// ---------------------------------------------------------------------------------------------
inline void point_light_dir(V3 location, V3 pt, V3* ldir, double* r2) {
    *ldir = normalize(V3{location.x - pt.x, location.y - pt.y, location.z - pt.z});
    *r2 = sqnorm(location - pt);  // sqdist
}
// returns has_range
inline bool light_dir_and_sq_range(const Light& l, V3 pt, Rng& rng, uint64_t key, uint64_t dim0, V3* ldir, double* r2) {
    switch (l.kind) {
    case LIGHT_POINT:
        point_light_dir(l.a, pt, ldir, r2);
        return true;
    case LIGHT_DIRECTIONAL:
        *ldir = -l.a;
        return false;
    default: {
        double u1 = rng.f64(key, dim0);
        double u2 = rng.f64(key, dim0 + 1);
        V3 loc = l.a + l.s1 * u1 + l.s2 * u2;
        point_light_dir(loc, pt, ldir, r2);
        return true;
    }
    }
}

// ---------------------------------------------------------------------------------------------
// tracing (src/raytrace.rs)
// ---------------------------------------------------------------------------------------------
struct TraceCtx {
    const Scene* scene;
    Rng rng;
    Stats stats;
    uint32_t max_depth;       // src/raytrace.rs:18 (4 upstream)
    double min_significance;  // src/raytrace.rs:17 (1/512 upstream)
    int32_t* primary_id;      // where to record the depth-0 hit of this sample (or null)
};

inline double clamp_zero(double x) { return x < 0.0 ? 0.0 : x; }
inline double clamp_one(double x) { return x > 1.0 ? 1.0 : x; }

Color ray_color(TraceCtx& c, const Ray& ray, double sig, uint32_t depth, uint64_t key);

// shadow test shared by all four materials (src/raytrace.rs:43-50, 83-90, 143-150, 198-205)
inline bool in_shadow(TraceCtx& c, V3 pt, V3 ldir, bool has_range, double r2) {
    Hit h{};
    c.stats.shadow_rays++;
    int i = scene_intersect(*c.scene, Ray{pt + ldir * 0.00001, ldir}, &h, &c.stats);
    if (i < 0) return false;
    return has_range ? (h.t * h.t < r2) : true;
}

Color material_color(TraceCtx& c, const Material& m, const Hit& result, const Ray& ray, double sig, uint32_t depth, uint64_t key) {
    const Scene& sc = *c.scene;
    const size_t L = sc.lights.size();
    const double FRAC_1_PI = 0.318309886183790671537767526745028724;
    switch (m.kind) {
    case MAT_PHONG: {  // src/raytrace.rs:30-67
        Color res = m.ambient;
        if (depth > c.max_depth) return res;
        V3 pt = cast(ray, result.t);
        bool diffuse = significance(m.diffuse) * sig > c.min_significance;
        bool specular = significance(m.specular) * sig > c.min_significance;
        V3 normal = dot(result.normal, ray.direction) > 0.0 ? -result.normal : result.normal;
        for (size_t li = 0; li < L; li++) {
            if (diffuse || specular) {
                const Light& light = sc.lights[li];
                V3 ldir; double r2 = 0;
                bool has = light_dir_and_sq_range(light, pt, c.rng, key, 2 + 2 * li, &ldir, &r2);
                if (in_shadow(c, pt, ldir, has, r2)) continue;
                if (diffuse) res = res + m.diffuse * light.color * clamp_zero(dot(ldir, normal)) * FRAC_1_PI;
                if (specular) res = res + m.specular * light.color * std::pow(clamp_zero(dot(normal, normalize(ldir - ray.direction))), m.exponent);
            }
        }
        if (specular) {
            V3 d = ray.direction;
            V3 rd = d - normal * (2.0 * dot(d, normal));
            Ray reflect{pt + rd * 0.00001, rd};
            res = res + m.specular * ray_color(c, reflect, sig * significance(m.specular), depth + 1, child_key(key, CODE_REFLECT));
        }
        return res;
    }
    case MAT_INDIRECT: {  // src/raytrace.rs:69-121
        Color res = m.ambient;
        if (depth > c.max_depth) return res;
        V3 pt = cast(ray, result.t);
        bool diffuse = significance(m.diffuse) * sig > c.min_significance;
        bool specular = significance(m.specular) * sig > c.min_significance;
        V3 normal = dot(result.normal, ray.direction) > 0.0 ? -result.normal : result.normal;
        if (diffuse || specular) {
            for (size_t li = 0; li < L; li++) {
                const Light& light = sc.lights[li];
                V3 ldir; double r2 = 0;
                bool has = light_dir_and_sq_range(light, pt, c.rng, key, 2 + 2 * li, &ldir, &r2);
                if (in_shadow(c, pt, ldir, has, r2)) continue;
                if (diffuse) res = res + m.diffuse * light.color * clamp_zero(dot(ldir, normal)) * FRAC_1_PI;
                if (specular) res = res + m.specular * light.color * std::pow(clamp_zero(dot(normal, normalize(ldir - ray.direction))), m.exponent);
            }
            for (uint32_t k = 0; k < m.samples; k++) {
                double r1 = c.rng.f64(key, 2 + 2 * L + 2 * k) * 2.0 - 1.0;
                double r2 = c.rng.f64(key, 2 + 2 * L + 2 * k + 1) * (2.0 * M_PI);
                double sin_theta = 1.0 - r1 * r1;  // sic: no sqrt upstream (src/raytrace.rs:103)
                double phi = r2;
                double x = sin_theta * std::cos(phi);
                double z = sin_theta * std::sin(phi);
                V3 d0{x, r1, z};
                V3 dir = dot(d0, normal) >= 0.0 ? d0 : -d0;
                Ray child{pt + dir * 0.00001, dir};  // shadows `ray` upstream (src/raytrace.rs:108)
                Color col = ray_color(c, child, sig, depth + 1, child_key(key, CODE_INDIRECT0 + k));
                double fac = (double)m.samples * 0.5;
                if (diffuse) res = res + m.diffuse * col * dot(normal, dir) / fac;
                // (dir - child.direction) is the zero vector -> normalize gives NaN (src/raytrace.rs:115)
                if (specular) res = res + m.specular * col * std::pow(clamp_zero(dot(normal, normalize(dir - child.direction))), m.exponent) / fac;
            }
        }
        return res;
    }
    case MAT_FRESNEL: {  // src/raytrace.rs:123-167
        Color res = m.ambient;
        if (depth > c.max_depth) return res;
        V3 pt = cast(ray, result.t);
        double nd = dot(result.normal, ray.direction);
        V3 normal = nd > 0.0 ? -result.normal : result.normal;
        double r0 = (m.ior - 1.0) / (m.ior + 1.0);
        r0 = r0 * r0;
        double omcos = 1.0 - std::fabs(nd);
        double omcos2 = omcos * omcos;
        double fresnel = clamp_one(r0 + (1.0 - r0) * omcos2 * omcos2 * omcos);
        bool diffuse = significance(m.diffuse) * sig > c.min_significance;
        bool specular = significance(m.specular) * fresnel * sig > c.min_significance;
        for (size_t li = 0; li < L; li++) {
            if (diffuse || specular) {
                const Light& light = sc.lights[li];
                V3 ldir; double r2 = 0;
                bool has = light_dir_and_sq_range(light, pt, c.rng, key, 2 + 2 * li, &ldir, &r2);
                if (in_shadow(c, pt, ldir, has, r2)) continue;
                if (diffuse) res = res + m.diffuse * light.color * clamp_zero(dot(ldir, normal)) * FRAC_1_PI;
                if (specular) res = res + m.specular * light.color * fresnel * std::pow(clamp_zero(dot(normal, normalize(ldir - ray.direction))), m.exponent);
            }
        }
        if (specular) {
            V3 d = ray.direction;
            V3 rd = d - normal * (2.0 * dot(d, normal));
            Ray reflect{pt + rd * 0.00001, rd};
            res = res + m.specular * ray_color(c, reflect, fresnel * sig * significance(m.specular), depth + 1, child_key(key, CODE_REFLECT)) * fresnel;
        }
        return res;
    }
    default: {  // MAT_TRANSPARENT, src/raytrace.rs:169-226
        Color res = BLACK;
        if (depth > c.max_depth) return res;
        V3 pt = cast(ray, result.t);
        double nd = dot(result.normal, ray.direction);
        V3 normal = nd > 0.0 ? -result.normal : result.normal;
        double ndv = dot(normal, ray.direction);
        double n = nd > 0.0 ? m.ior : 1.0 / m.ior;
        double sin2 = n * n * (1.0 - nd * nd);
        bool has_refract = sin2 < 1.0;
        V3 refract{0, 0, 0};
        if (has_refract) {
            double cs = std::sqrt(1.0 - sin2);
            refract = ray.direction * n - normal * (n * std::fabs(nd) + cs);
        }
        double r0 = (m.ior - 1.0) / (m.ior + 1.0);
        r0 = r0 * r0;
        double omcos = nd > 0.0 ? (has_refract ? 1.0 - dot(normal, refract) : 0.0) : 1.0 - std::fabs(nd);
        double omcos2 = omcos * omcos;
        double fresnel = has_refract ? clamp_one(r0 + (1.0 - r0) * omcos2 * omcos2 * omcos) : 1.0;
        bool specular = significance(m.specular) * fresnel * sig > c.min_significance;
        for (size_t li = 0; li < L; li++) {
            if (specular) {
                const Light& light = sc.lights[li];
                V3 ldir; double r2 = 0;
                bool has = light_dir_and_sq_range(light, pt, c.rng, key, 2 + 2 * li, &ldir, &r2);
                if (in_shadow(c, pt, ldir, has, r2)) continue;
                res = res + m.specular * light.color * fresnel * std::pow(clamp_zero(dot(normal, normalize(ldir - ray.direction))), m.exponent);
            }
        }
        if (specular) {
            V3 rd = ray.direction - normal * (2.0 * ndv);
            Ray reflect{pt + rd * 0.00001, rd};
            res = res + m.specular * ray_color(c, reflect, fresnel * sig * significance(m.specular), depth + 1, child_key(key, CODE_REFLECT)) * fresnel;
        }
        if (fresnel < 1.0 && has_refract) {
            double omf = clamp_one(1.0 - fresnel);
            V3 rn = normalize(refract);
            res = res + ray_color(c, Ray{pt + rn * 0.00001, rn}, omf * sig, depth + 1, child_key(key, CODE_REFRACT)) * omf;
        }
        return res;
    }
    }
}

// src/raytrace.rs:261-267
Color ray_color(TraceCtx& c, const Ray& ray, double sig, uint32_t depth, uint64_t key) {
    Hit h{};
    int i = scene_intersect(*c.scene, ray, &h, &c.stats);
    if (depth == 0 && c.primary_id) { *c.primary_id = i; c.primary_id = nullptr; }
    if (i >= 0) return material_color(c, c.scene->materials[c.scene->objects[i].material], h, ray, sig, depth, key);
    return background_color(*c.scene, ray);
}

// src/raytrace.rs:270-276
Color raytrace(TraceCtx& c, double px, double py, double sig, uint64_t sample_key) {
    Color res = BLACK;
    const uint32_t n = c.scene->camera.samples;
    for (uint32_t s = 0; s < n; s++) {
        uint64_t ck = child_key(sample_key, CODE_CAMERA0 + s);
        Ray r = camera_project(c.scene->camera, px, py, c.rng, ck);
        res = res + ray_color(c, r, sig, 0, ck);
    }
    return res / (double)n;
}

struct Opts {
    uint32_t width, height, antialias;
    uint32_t max_depth;
    double min_significance;
    uint64_t seed;
    int32_t rng_mode;
    int32_t threads;  // 0 = all
};

// one pixel of src/main.rs:47-56
inline Color render_pixel(TraceCtx& c, const Opts& o, uint32_t x, uint32_t y, int32_t* primary_id) {
    const double halfwidth = (double)o.width / 2.0;
    const double halfheight = (double)o.height / 2.0;
    const double scale = std::fmax(1.0 / halfwidth, 1.0 / halfheight);
    Color res = BLACK;
    const uint64_t pixel = (uint64_t)y * o.width + x;
    for (uint32_t s = 0; s < o.antialias; s++) {
        uint64_t sk = key_sample(o.seed, pixel, s);
        double u1 = c.rng.f64(sk, 0);
        double u2 = c.rng.f64(sk, 1);
        double px = (((double)x + u1) - halfwidth) * scale;
        double py = (((double)y + u2) - halfheight) * scale;
        c.primary_id = (s == 0) ? primary_id : nullptr;
        res = res + raytrace(c, px, py, 1.0, sk);
    }
    return res / (double)o.antialias;
}

inline void write_bgr(Color c, uint8_t* buf, size_t i) {  // src/color.rs:628-632
    buf[i * 3] = to_srgb(c.b);
    buf[i * 3 + 1] = to_srgb(c.g);
    buf[i * 3 + 2] = to_srgb(c.r);
}

inline void seed_rng(Rng& r, const Opts& o) {
    r.mode = o.rng_mode;
    // rand 0.3 XorShiftRng::new_unseeded constants, perturbed by the seed
    uint64_t s = mix64(o.seed + 1);
    r.x = 0x193a6754u ^ (uint32_t)s;
    r.y = 0xa8a7d469u ^ (uint32_t)(s >> 32);
    r.z = 0x97830e05u;
    r.w = 0x113ba7bbu;
    if ((r.x | r.y | r.z | r.w) == 0) r.w = 1;
}

inline void add_stats(Stats& a, const Stats& b) {
    a.rays += b.rays; a.shadow_rays += b.shadow_rays; a.shape_tests += b.shape_tests;
}

}  // namespace

// =================================================================================================
// C API (ctypes)
// =================================================================================================
extern "C" {

struct orc_opts {
    uint32_t width, height, antialias;
    uint32_t max_depth;
    double min_significance;
    uint64_t seed;
    int32_t rng_mode;
    int32_t threads;
};
struct orc_stats {
    uint64_t rays, shadow_rays, shape_tests;
    double seconds;
    int32_t threads;
    int32_t _pad;
};

void* orc_scene_new() { return new Scene(); }
void orc_scene_free(void* s) { delete (Scene*)s; }

static Color col3(const double* c) { return Color{c[0], c[1], c[2]}; }
static V3 vec3(const double* v) { return V3{v[0], v[1], v[2]}; }

int orc_add_material(void* s, int kind, const double* diffuse, const double* specular, double exponent, const double* ambient, double ior, uint32_t samples) {
    Scene* sc = (Scene*)s;
    Material m{};
    m.kind = kind;
    m.diffuse = diffuse ? col3(diffuse) : BLACK;
    m.specular = specular ? col3(specular) : BLACK;
    m.ambient = ambient ? col3(ambient) : BLACK;
    m.exponent = exponent; m.ior = ior; m.samples = samples;
    sc->materials.push_back(m);
    return (int)sc->materials.size() - 1;
}
int orc_add_spheres(void* s, uint64_t n, const double* cxyz_r, const uint32_t* material) {
    Scene* sc = (Scene*)s;
    for (uint64_t i = 0; i < n; i++) {
        Object o{};
        o.kind = SHAPE_SPHERE;
        o.p = V3{cxyz_r[4 * i], cxyz_r[4 * i + 1], cxyz_r[4 * i + 2]};
        o.radius = cxyz_r[4 * i + 3];
        o.material = material[i];
        if (o.material >= sc->materials.size()) return -1;
        sc->objects.push_back(o);
    }
    return 0;
}
int orc_add_plane(void* s, const double* point, const double* normal, uint32_t material) {
    Scene* sc = (Scene*)s;
    if (material >= sc->materials.size()) return -1;
    Object o{};
    o.kind = SHAPE_PLANE; o.p = vec3(point); o.n = vec3(normal); o.material = material;
    sc->objects.push_back(o);
    return 0;
}
int orc_add_light(void* s, int kind, const double* a, const double* s1, const double* s2, const double* color) {
    Scene* sc = (Scene*)s;
    Light l{};
    l.kind = kind; l.a = vec3(a);
    if (s1) l.s1 = vec3(s1);
    if (s2) l.s2 = vec3(s2);
    l.color = col3(color);
    sc->lights.push_back(l);
    return 0;
}
void orc_set_camera_spc_new(void* s, const double* pos, const double* look, const double* up, double im_dist) {
    camera_spc_new(&((Scene*)s)->camera, vec3(pos), vec3(look), vec3(up), im_dist);
}
void orc_set_camera_spc_look_at(void* s, const double* focus, const double* look, const double* up, double pov, double h) {
    camera_spc_look_at(&((Scene*)s)->camera, vec3(focus), vec3(look), vec3(up), pov, h);
}
void orc_set_camera_dof(void* s, double focus, double aperture, uint32_t samples) {
    camera_make_dof(&((Scene*)s)->camera, focus, aperture, samples);
}
void orc_get_camera(void* s, double* position3, double* matrix9) {
    Scene* sc = (Scene*)s;
    position3[0] = sc->camera.position.x; position3[1] = sc->camera.position.y; position3[2] = sc->camera.position.z;
    std::memcpy(matrix9, sc->camera.m, sizeof sc->camera.m);
}
void orc_set_background_solid(void* s, const double* color) {
    Scene* sc = (Scene*)s;
    sc->bg_kind = BG_SOLID; sc->bg_color = col3(color);
}
int orc_set_skybox_face(void* s, int face, uint32_t w, uint32_t h, const uint8_t* rgb) {
    Scene* sc = (Scene*)s;
    if (face < 0 || face > 5 || w == 0 || h == 0) return -1;
    sc->bg_kind = BG_SKYBOX;
    sc->faces[face].width = w; sc->faces[face].height = h;
    sc->faces[face].data.assign(rgb, rgb + (size_t)3 * w * h);
    return 0;
}
void orc_set_options(void* s, uint32_t w, uint32_t h, uint32_t aa) {
    Scene* sc = (Scene*)s;
    sc->width = w; sc->height = h; sc->antialias = aa;
}

uint32_t orc_bmp_bytewidth(uint32_t width) { return (3 * width + 3) & 0xFFFFFFFCu; }

// src/bmp.rs:10-61 -- 14-byte file header + 108-byte BITMAPV4HEADER
void orc_bmp_header(uint32_t width, uint32_t height, uint8_t* out122) {
    uint32_t bytewidth = orc_bmp_bytewidth(width);
    uint32_t pasize = bytewidth * height;
    uint32_t fsize = 14 + 108 + pasize;
    std::memset(out122, 0, 122);
    auto le32 = [&](int off, uint32_t v) { for (int i = 0; i < 4; i++) out122[off + i] = (uint8_t)(v >> (8 * i)); };
    out122[0] = 0x42; out122[1] = 0x4D;
    le32(2, fsize);
    le32(10, 0x7A);
    le32(14, 0x6C);
    le32(18, width);
    le32(22, height);
    out122[26] = 1; out122[28] = 0x18;
    le32(34, pasize);
    le32(38, 0x0B13);
    le32(42, 0x0B13);
    out122[70] = 0x42; out122[71] = 0x47; out122[72] = 0x52; out122[73] = 0x73;
}

void orc_quantize_bgr(const double* rgb, uint64_t n, uint8_t* bgr) {
    for (uint64_t i = 0; i < n; i++) write_bgr(Color{rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]}, bgr, i);
}
void orc_srgb_tables(double* values256, double* average255) {
    std::memcpy(values256, SRGB_VALUES, sizeof SRGB_VALUES);
    std::memcpy(average255, SRGB_AVERAGE, sizeof SRGB_AVERAGE);
}

// batch closest hit: Scene::intersect on n rays (origin xyz, direction xyz)
void orc_trace_rays(void* s, uint64_t n, const double* origin_dir, int32_t* ids, double* t, double* normal) {
    Scene* sc = (Scene*)s;
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        Stats st;
        Hit h{std::numeric_limits<double>::quiet_NaN(), V3{0, 0, 0}};
        Ray r{V3{origin_dir[6 * i], origin_dir[6 * i + 1], origin_dir[6 * i + 2]}, V3{origin_dir[6 * i + 3], origin_dir[6 * i + 4], origin_dir[6 * i + 5]}};
        int id = scene_intersect(*sc, r, &h, &st);
        ids[i] = id;
        if (t) t[i] = id >= 0 ? h.t : std::numeric_limits<double>::infinity();
        if (normal) { normal[3 * i] = id >= 0 ? h.normal.x : 0; normal[3 * i + 1] = id >= 0 ? h.normal.y : 0; normal[3 * i + 2] = id >= 0 ? h.normal.z : 0; }
    }
}

// camera.project for tests: pos (n x 2) -> rays (n x 6), counter RNG with key_sample(seed, i, 0) / camera sample 0
void orc_project(void* s, uint64_t n, const double* pos, uint64_t seed, double* origin_dir) {
    Scene* sc = (Scene*)s;
    Rng rng{}; rng.mode = ORC_RNG_COUNTER;
    for (uint64_t i = 0; i < n; i++) {
        uint64_t ck = child_key(key_sample(seed, i, 0), CODE_CAMERA0);
        Ray r = camera_project(sc->camera, pos[2 * i], pos[2 * i + 1], rng, ck);
        origin_dir[6 * i] = r.origin.x; origin_dir[6 * i + 1] = r.origin.y; origin_dir[6 * i + 2] = r.origin.z;
        origin_dir[6 * i + 3] = r.direction.x; origin_dir[6 * i + 4] = r.direction.y; origin_dir[6 * i + 5] = r.direction.z;
    }
}

// ray_color for tests (src/raytrace.rs:261-267): colour of each given ray at the given significance and recursion depth.
// rng_mode ORC_RNG_COUNTER: ray i draws from keys[i].  ORC_RNG_SEQUENTIAL: ONE xorshift128 stream with the given state is
// threaded through all rays in order, exactly as upstream threads `rng` (so a test can predict every draw, in order).
// rays_out (optional): Scene::intersect calls made, shadow queries included.
void orc_ray_color(void* s, uint64_t n, const double* origin_dir, double significance, uint32_t depth, const uint64_t* keys, uint32_t max_depth,
                   double min_significance, int32_t rng_mode, const uint32_t* xorshift_state4, double* rgb_out, uint64_t* rays_out) {
    const Scene* sc = (const Scene*)s;
    TraceCtx c{};
    c.scene = sc;
    c.max_depth = max_depth;
    c.min_significance = min_significance;
    c.primary_id = nullptr;
    c.rng.mode = rng_mode;
    if (xorshift_state4) { c.rng.x = xorshift_state4[0]; c.rng.y = xorshift_state4[1]; c.rng.z = xorshift_state4[2]; c.rng.w = xorshift_state4[3]; }
    for (uint64_t i = 0; i < n; i++) {
        Ray r{V3{origin_dir[6 * i], origin_dir[6 * i + 1], origin_dir[6 * i + 2]}, V3{origin_dir[6 * i + 3], origin_dir[6 * i + 4], origin_dir[6 * i + 5]}};
        Color col = ray_color(c, r, significance, depth, keys ? keys[i] : 0);
        rgb_out[3 * i] = col.r; rgb_out[3 * i + 1] = col.g; rgb_out[3 * i + 2] = col.b;
    }
    if (rays_out) *rays_out = c.stats.rays;
}

// camera.project with the reference's sequential RNG (src/camera.rs:110-121 draws theta, then the radius)
void orc_project_sequential(void* s, uint64_t n, const double* pos, const uint32_t* xorshift_state4, double* origin_dir) {
    Scene* sc = (Scene*)s;
    Rng rng{}; rng.mode = ORC_RNG_SEQUENTIAL;
    rng.x = xorshift_state4[0]; rng.y = xorshift_state4[1]; rng.z = xorshift_state4[2]; rng.w = xorshift_state4[3];
    for (uint64_t i = 0; i < n; i++) {
        Ray r = camera_project(sc->camera, pos[2 * i], pos[2 * i + 1], rng, 0);
        origin_dir[6 * i] = r.origin.x; origin_dir[6 * i + 1] = r.origin.y; origin_dir[6 * i + 2] = r.origin.z;
        origin_dir[6 * i + 3] = r.direction.x; origin_dir[6 * i + 4] = r.direction.y; origin_dir[6 * i + 5] = r.direction.z;
    }
}

static void fill_opts(const Scene* sc, const orc_opts* in, Opts* o) {
    o->width = in->width ? in->width : sc->width;
    o->height = in->height ? in->height : sc->height;
    o->antialias = in->antialias != 0xFFFFFFFFu ? in->antialias : sc->antialias;
    o->max_depth = in->max_depth;
    o->min_significance = in->min_significance;
    o->seed = in->seed;
    o->rng_mode = in->rng_mode;
    o->threads = in->threads;
}

static double now_s() {
#ifdef _OPENMP
    return omp_get_wtime();
#else
    return (double)clock() / CLOCKS_PER_SEC;
#endif
}

// Full frame: src/main.rs:39-59.  bgr_rows = bytewidth*height bytes, row 0 = y 0 (BMP bottom-up order).
// primary_ids (optional) = width*height object index of the first sample's primary ray, -1 = miss.
// accum (optional) = width*height*3 linear RGB doubles (the value handed to write_bgr).
int orc_render(void* s, const orc_opts* in, uint8_t* bgr_rows, int32_t* primary_ids, double* accum, orc_stats* stats) {
    const Scene* sc = (const Scene*)s;
    Opts o; fill_opts(sc, in, &o);
    const uint32_t bytewidth = orc_bmp_bytewidth(o.width);
    Stats total;
    int threads = 1;
    double t0 = now_s();
    if (o.rng_mode == ORC_RNG_SEQUENTIAL) {
        TraceCtx c{sc, Rng{}, Stats{}, o.max_depth, o.min_significance, nullptr};
        seed_rng(c.rng, o);
        for (uint32_t y = 0; y < o.height; y++) {
            uint8_t* row = bgr_rows + (size_t)y * bytewidth;
            std::memset(row, 0, bytewidth);
            for (uint32_t x = 0; x < o.width; x++) {
                size_t p = (size_t)y * o.width + x;
                Color res = render_pixel(c, o, x, y, primary_ids ? primary_ids + p : nullptr);
                if (accum) { accum[3 * p] = res.r; accum[3 * p + 1] = res.g; accum[3 * p + 2] = res.b; }
                write_bgr(res, row, x);
            }
        }
        total = c.stats;
    } else {
#ifdef _OPENMP
        threads = o.threads > 0 ? o.threads : omp_get_max_threads();
#endif
#pragma omp parallel num_threads(threads)
        {
            TraceCtx c{sc, Rng{}, Stats{}, o.max_depth, o.min_significance, nullptr};
            c.rng.mode = ORC_RNG_COUNTER;
#pragma omp for schedule(dynamic, 1)
            for (int64_t y = 0; y < (int64_t)o.height; y++) {
                uint8_t* row = bgr_rows + (size_t)y * bytewidth;
                std::memset(row, 0, bytewidth);
                for (uint32_t x = 0; x < o.width; x++) {
                    size_t p = (size_t)y * o.width + x;
                    Color res = render_pixel(c, o, x, (uint32_t)y, primary_ids ? primary_ids + p : nullptr);
                    if (accum) { accum[3 * p] = res.r; accum[3 * p + 1] = res.g; accum[3 * p + 2] = res.b; }
                    write_bgr(res, row, x);
                }
            }
#pragma omp critical
            add_stats(total, c.stats);
        }
    }
    if (stats) {
        stats->rays = total.rays; stats->shadow_rays = total.shadow_rays; stats->shape_tests = total.shape_tests;
        stats->seconds = now_s() - t0; stats->threads = threads;
    }
    return 0;
}

// Sparse pixel set (counter RNG only): pixels[i] = y*width + x.  bgr = 3n bytes, ids = n, accum = 3n doubles.
int orc_render_pixels(void* s, const orc_opts* in, uint64_t n, const uint32_t* pixels, uint8_t* bgr, int32_t* ids, double* accum, orc_stats* stats) {
    const Scene* sc = (const Scene*)s;
    Opts o; fill_opts(sc, in, &o);
    if (o.rng_mode != ORC_RNG_COUNTER) return -1;
    Stats total;
    int threads = 1;
#ifdef _OPENMP
    threads = o.threads > 0 ? o.threads : omp_get_max_threads();
#endif
    double t0 = now_s();
#pragma omp parallel num_threads(threads)
    {
        TraceCtx c{sc, Rng{}, Stats{}, o.max_depth, o.min_significance, nullptr};
        c.rng.mode = ORC_RNG_COUNTER;
#pragma omp for schedule(dynamic, 16)
        for (int64_t i = 0; i < (int64_t)n; i++) {
            uint32_t p = pixels[i];
            Color res = render_pixel(c, o, p % o.width, p / o.width, ids ? ids + i : nullptr);
            if (accum) { accum[3 * i] = res.r; accum[3 * i + 1] = res.g; accum[3 * i + 2] = res.b; }
            if (bgr) write_bgr(res, bgr, (size_t)i);
        }
#pragma omp critical
        add_stats(total, c.stats);
    }
    if (stats) {
        stats->rays = total.rays; stats->shadow_rays = total.shadow_rays; stats->shape_tests = total.shape_tests;
        stats->seconds = now_s() - t0; stats->threads = threads;
    }
    return 0;
}

// test hooks for the RNG definition
uint64_t orc_key_sample(uint64_t seed, uint64_t pixel, uint64_t aa) { return key_sample(seed, pixel, aa); }
uint64_t orc_child_key(uint64_t k, uint64_t code) { return child_key(k, code); }
uint64_t orc_draw_bits(uint64_t k, uint64_t dim) { return draw_bits(k, dim); }
void orc_xorshift(uint32_t x, uint32_t y, uint32_t z, uint32_t w, uint32_t n, uint32_t* out) {
    Rng r{ORC_RNG_SEQUENTIAL, x, y, z, w};
    for (uint32_t i = 0; i < n; i++) out[i] = r.next_u32();
}

}  // extern "C"
