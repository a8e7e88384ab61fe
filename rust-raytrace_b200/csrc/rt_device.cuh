// rt_device.cuh -- per-ray arithmetic of the tracing path: RNG, cameras, shape tests, BVH traversal, shading,
// backgrounds, quantiser.  Everything here is a pure function over rt_types.h records.
//
// The functions are `__host__ __device__` so that tests/emu/ can compile them with g++ and run the wavefront
// formulation serially on the CPU *as a test of this header* (the product library never contains a CPU render
// path; see tests/emu/README.md).  In the product they are only ever called from the kernels in kernels.cu.
//
// Arithmetic order follows the reference expression by expression (file:line cited per function).  The
// translation unit is compiled twice: -fmad=false (namespace rt::strict, bit-parity with the reference's
// unfused x86-64 arithmetic) and -fmad=true (namespace rt::fast).
#pragma once
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "rt_types.h"

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#else
#define RT_HD inline
#endif

#if defined(__CUDA_ARCH__)
#define RT_LDG(p) __ldg(p)
#define RT_PREFETCH_L1(p) asm volatile("prefetch.global.L1 [%0];" ::"l"(p))
#else
#define RT_LDG(p) (*(p))
#define RT_PREFETCH_L1(p) ((void)0)
#endif

#ifndef RT_NS
#define RT_NS strict
#endif

// Bounds checks of our own (compute-sanitizer is closed on the GPU pool this was developed on).  A build with
// -DRT_CHECK_BOUNDS (tools/build_variant.sh checked -DRT_CHECK_BOUNDS) traps -- the launch fails, the render returns
// RT_ERR_CUDA -- on a node / leaf / stack / queue / pixel index outside its buffer; the GPU test-suite and
// tools/gpu_sanitize.py are then run through that library (profiles/r02_bounds_checked_build.txt).  Off: no code.
#if defined(RT_CHECK_BOUNDS)
#if defined(__CUDA_ARCH__)
#define RT_CHECK(c) do { if (!(c)) __trap(); } while (0)
#else
#define RT_CHECK(c) do { if (!(c)) abort(); } while (0)
#endif
#else
#define RT_CHECK(c) ((void)0)
#endif

namespace rt {
namespace RT_NS {

// ---------------------------------------------------------------------------------------------------------
// value types (src/types.rs:10-25, src/color.rs:15-67)
// ---------------------------------------------------------------------------------------------------------
struct V3 {
    double x, y, z;
};
RT_HD V3 v3(double x, double y, double z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
RT_HD V3 v3(const double* p) { return v3(p[0], p[1], p[2]); }
RT_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
RT_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
RT_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
RT_HD V3 operator*(V3 a, double s) { return v3(a.x * s, a.y * s, a.z * s); }
RT_HD V3 operator/(V3 a, double s) { return v3(a.x / s, a.y / s, a.z / s); }
RT_HD double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
RT_HD double sqnorm(V3 a) { return dot(a, a); }
RT_HD V3 normalize(V3 a) { return a / sqrt(sqnorm(a)); }

struct C3 {
    double r, g, b;
};
RT_HD C3 c3(double r, double g, double b) { C3 c; c.r = r; c.g = g; c.b = b; return c; }
RT_HD C3 c3(const double* p) { return c3(p[0], p[1], p[2]); }
RT_HD C3 operator+(C3 a, C3 b) { return c3(a.r + b.r, a.g + b.g, a.b + b.b); }
RT_HD C3 operator*(C3 a, C3 b) { return c3(a.r * b.r, a.g * b.g, a.b * b.b); }
RT_HD C3 operator*(C3 a, double s) { return c3(a.r * s, a.g * s, a.b * s); }
RT_HD C3 operator/(C3 a, double s) { return c3(a.r / s, a.g / s, a.b / s); }
RT_HD double significance(C3 c) { return c.r + c.g + c.b; }  // src/color.rs:637-639

RT_HD double clamp_zero(double x) { return x < 0.0 ? 0.0 : x; }  // src/raytrace.rs:21-23
RT_HD double clamp_one(double x) { return x > 1.0 ? 1.0 : x; }   // src/raytrace.rs:25-28

// ---------------------------------------------------------------------------------------------------------
// counter RNG (DESIGN.md "RNG"); replaces the sequential XorShiftRng of src/types.rs:27 / src/main.rs:43
// ---------------------------------------------------------------------------------------------------------
RT_HD uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
#define RT_GOLD 0x9E3779B97F4A7C15ULL
RT_HD uint64_t key_sample(uint64_t seed, uint64_t pixel, uint64_t aa) {
    uint64_t k = mix64(seed ^ 0x243F6A8885A308D3ULL);
    k = mix64(k + RT_GOLD * (pixel + 1));
    return mix64(k + RT_GOLD * (aa + 1));
}
RT_HD uint64_t child_key(uint64_t k, uint64_t code) { return mix64(k + RT_GOLD * (code + 1)); }
RT_HD uint64_t draw_bits(uint64_t k, uint64_t dim) { return mix64((k ^ 0xD6E8FEB86659FD93ULL) + RT_GOLD * (dim + 1)); }
RT_HD double draw_f64(uint64_t k, uint64_t dim) { return (double)(draw_bits(k, dim) >> 11) * 0x1.0p-53; }
RT_HD double draw_closed01(uint64_t k, uint64_t dim) { return (double)(draw_bits(k, dim) >> 11) / 9007199254740991.0; }
constexpr uint64_t CODE_REFLECT = 1, CODE_REFRACT = 2, CODE_INDIRECT0 = 3, CODE_CAMERA0 = 1ULL << 32;

#define RT_PI 3.14159265358979323846264338327950288
#define RT_FRAC_1_PI 0.318309886183790671537767526745028724

// ---------------------------------------------------------------------------------------------------------
// cameras (src/camera.rs:77-79, :110-121)
// ---------------------------------------------------------------------------------------------------------
RT_HD V3 mat_mul(const double* m, V3 v) {
    return v3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[3] * v.x + m[4] * v.y + m[5] * v.z, m[6] * v.x + m[7] * v.y + m[8] * v.z);
}

RT_HD void camera_project(const CameraRec& cam, double px, double py, uint64_t key, V3& origin, V3& direction) {
    V3 dir = mat_mul(cam.m, v3(px, py, 1.0));
    if (cam.kind == CAM_SPC) {
        origin = v3(cam.pos);
        direction = normalize(dir);
        return;
    }
    V3 pos = v3(cam.pos);
    V3 ip = pos + dir;
    V3 fp = pos + dir * (cam.focus / cam.im_dist);
    double theta = draw_f64(key, 0) * (2.0 * RT_PI);
    double r2 = draw_closed01(key, 1);
    double r = sqrt(r2) * cam.aperture;
    V3 orig = ip + mat_mul(cam.m, v3(cos(theta) * r, sin(theta) * r, 0.0));
    origin = orig;
    direction = normalize(fp - orig);
}

// pixel -> normalised image position with jitter (src/main.rs:39-53)
RT_HD void pixel_position(const FrameParams& fp, uint32_t x, uint32_t y, uint64_t sample_key, double& px, double& py) {
    double halfwidth = (double)fp.width / 2.0;
    double halfheight = (double)fp.height / 2.0;
    double scale = fmax(1.0 / halfwidth, 1.0 / halfheight);
    double u1 = draw_f64(sample_key, 0);
    double u2 = draw_f64(sample_key, 1);
    px = (((double)x + u1) - halfwidth) * scale;
    py = (((double)y + u2) - halfheight) * scale;
}

// ---------------------------------------------------------------------------------------------------------
// shapes (src/shapes.rs:50-89, :100-112).  `a`, `fa` = 4a, `ta` = 2a are per-ray invariants of the quadratic.
// ---------------------------------------------------------------------------------------------------------
struct RayQ {
    V3 o, d;
    double a, fa, ta;
};
RT_HD RayQ make_rayq(V3 o, V3 d) {
    RayQ r;
    r.o = o; r.d = d;
    r.a = sqnorm(d);
    r.fa = 4.0 * r.a;
    r.ta = 2.0 * r.a;
    return r;
}

RT_HD bool sphere_hit(const RayQ& r, double cx, double cy, double cz, double r2, double& t_out) {
    V3 oc = v3(r.o.x - cx, r.o.y - cy, r.o.z - cz);
    double b = 2.0 * dot(r.d, oc);
    double c = sqnorm(oc) - r2;
    double disc = b * b - r.fa * c;
    if (disc > 0.0) {
        double dsqrt = sqrt(disc);
        double t = (-b - dsqrt) / r.ta;
        if (t > 0.0) { t_out = t; return true; }
        double t2 = (-b + dsqrt) / r.ta;
        if (t2 > 0.0) { t_out = t2; return true; }
    }
    return false;
}

// `t <= 0` rejects, so +inf and NaN are hits (src/shapes.rs:102-104)
RT_HD bool plane_hit(const RayQ& r, V3 p, V3 n, double& t_out) {
    const double num = dot(n, p - r.o), den = dot(n, r.d);
    // opposite signs or a zero numerator over a non-zero denominator: the quotient is negative or +-0 and `t <= 0` rejects
    // it -- no need to form it (the f64 division is ~30 instructions; a shadow ray leaving the floor towards a light, the
    // common case, ends here).  Everything else (same signs, a zero denominator, NaN) goes through the division as upstream.
    if ((num < 0.0 && den > 0.0) || (num > 0.0 && den < 0.0) || (num == 0.0 && den != 0.0 && den == den)) return false;
    double t = num / den;
    if (t <= 0.0) return false;
    t_out = t;
    return true;
}

// closest-hit bookkeeping with the reference's ordering (src/scene.rs:223-249):
// key = Option<FloatNotNan(t)>, None (NaN) sorts first, ties keep the lowest object index.
struct Closest {
    double t;
    int32_t obj;
    bool nan;
};
RT_HD Closest closest_none() { Closest c; c.t = (double)INFINITY; c.obj = 0x7FFFFFFF; c.nan = false; return c; }
RT_HD void closest_update(Closest& c, double t, int32_t obj) {
    if (c.nan) {
        if (t != t && obj < c.obj) c.obj = obj;
        return;
    }
    if (t != t) { c.nan = true; c.t = t; c.obj = obj; return; }
    if (t < c.t || (t == c.t && obj < c.obj)) { c.t = t; c.obj = obj; }
}
RT_HD int32_t closest_id(const Closest& c) { return c.obj == 0x7FFFFFFF ? -1 : c.obj; }

RT_HD void planes_closest(const DeviceScene& sc, const RayQ& r, Closest& best) {
    for (uint32_t i = 0; i < sc.n_planes; i++) {
        const PlaneRec* p = sc.planes + i;
        double t;
        if (plane_hit(r, v3(RT_LDG(&p->p[0]), RT_LDG(&p->p[1]), RT_LDG(&p->p[2])), v3(RT_LDG(&p->n[0]), RT_LDG(&p->n[1]), RT_LDG(&p->n[2])), t))
            closest_update(best, t, RT_LDG(&p->obj));
    }
}

// Shadow query semantics (src/raytrace.rs:43-50): occluded iff the CLOSEST hit exists and (no range or t*t < r2).
// `range2 < 0` encodes "no range".  Equivalent any-hit form: a NaN plane hit is the closest hit by the ordering
// above (occluded iff no range); otherwise occluded iff some hit has t*t < r2 (squaring is monotone for t > 0).
struct Shadow {
    bool nan, any;
};
RT_HD bool shadow_hit_counts(double t, double range2) { return range2 < 0.0 ? true : (t * t < range2); }
RT_HD void planes_shadow(const DeviceScene& sc, const RayQ& r, double range2, Shadow& s) {
    for (uint32_t i = 0; i < sc.n_planes; i++) {
        const PlaneRec* p = sc.planes + i;
        double t;
        if (plane_hit(r, v3(RT_LDG(&p->p[0]), RT_LDG(&p->p[1]), RT_LDG(&p->p[2])), v3(RT_LDG(&p->n[0]), RT_LDG(&p->n[1]), RT_LDG(&p->n[2])), t)) {
            if (t != t) s.nan = true;
            else if (shadow_hit_counts(t, range2)) s.any = true;
        }
    }
}
RT_HD bool shadow_result(const Shadow& s, double range2) { return s.nan ? (range2 < 0.0) : s.any; }

// ---------------------------------------------------------------------------------------------------------
// BVH traversal.
//
// The box test is only a conservative FILTER: it may pass a box the ray misses, never reject a box whose sphere
// the exact test would hit.  It runs in f32 (native FFMA/FMNMX) on the f32 boxes that the build rounded outward.
// The leaf test is the same f64 expression as the linear scan (sphere_hit), so the hit, its t and the tie-break
// are the scan's, bit for bit.
//
// Conservativeness of the f32 filter (DESIGN.md "BVH == linear scan"):
//   advance          -- the ray is first advanced (in f64) to where it enters the bounds of all sphere boxes, so the
//                       filter works on an origin of scene magnitude however far away the ray started (ground-plane
//                       hits near the horizon are 10^6..10^9 units out); filter distances are t - t0;
//   additive errors  -- rounding the advanced origin to f32, the rounding of (o +- pad)*inv, the error of the advance
//                       itself, and the rounding error of the exact test's discriminant (which makes the scan "hit"
//                       spheres inflated by up to min(E/2r_min, sqrt(E)), E = 64 eps S^2, S the magnitude of the
//                       ORIGINAL origin) -- are covered by widening every box by `pad`;
//   relative errors  -- inv = fl32(1/d), the f32 rounding of the per-ray slope and the FFMA rounding scale a slab distance
//                       by (1 +- 3*2^-24).  A slab distance times |d| is the displacement from o' to the plane along that
//                       axis, at most 2S because o' lies inside the bounds: the error is at most 6*2^-24 S in space, and
//                       is covered by the same pad (S * 2^-20) -- no per-box slack multiplications.
// ---------------------------------------------------------------------------------------------------------
#define RT_BVH_TSLACK 1.000000000001   // relative slack when a t of the exact test bounds the filter

// Traversal stacks cannot overflow, and the walks do not spend instructions checking: a walk postpones at most one
// child per level of the path it is on (binary) / three per two levels (wide), the Karras tree over 63 Morton bits +
// 32 tie-break bits is at most 95 levels deep, and the build MEASURES the depth of the tree it made and refuses a tree
// deeper than BVH_STACK (bvh_build.cu k_tree_depth -> rt_context_create fails) -- so a full stack is a build-time error,
// never a silently pruned walk.
constexpr int BVH_STACK = 96;          // >= 63 Morton bits + 32 tie-break bits
constexpr int32_t TRAV_EMPTY = (int32_t)0x80000000;  // "no node left"; never a valid ~leaf (n_spheres < 2^31 - 16)

struct TraverseCount {
    uint32_t nodes, prims;  // nodes fetched, exact sphere tests done
};
RT_HD TraverseCount traverse_count_zero() { TraverseCount t; t.nodes = 0; t.prims = 0; return t; }

struct BoxRay32 {
    // slab distance of a box plane at coordinate v: v * s - off, with one slope per axis and two offsets:
    //   f32 nodes : a = offset of lower planes ((o' + pad) * inv), b = offset of upper planes ((o' - pad) * inv);
    //   grid nodes: a = offset of the plane the ray ENTERS through, b = of the plane it LEAVES through, and sn* / sf*
    //               are the byte-permute selectors that pick that plane's 16-bit coordinate out of the node word.
    float sx, sy, sz, ax, ay, az, bx, by, bz;
    uint32_t snx, sny, snz, sfx, sfy, sfz;
    double t0;         // ray parameter of the advanced origin o' = o + d * t0 (0 when o is inside the sphere bounds)
    bool enter;        // false: no sphere can be hit (non-finite ray, or the ray misses the bounds) -> skip the tree
};
RT_HD BoxRay32 boxray32_none() {
    BoxRay32 b;
    b.sx = b.sy = b.sz = b.ax = b.ay = b.az = b.bx = b.by = b.bz = 0.0f;
    b.snx = b.sny = b.snz = b.sfx = b.sfy = b.sfz = 0u;
    b.t0 = 0.0; b.enter = false;
    return b;
}

#if defined(__CUDA_ARCH__)
RT_HD float f64_to_f32_down(double x) { return __double2float_rd(x); }
RT_HD float f64_to_f32_up(double x) { return __double2float_ru(x); }
#else
RT_HD float f64_to_f32_down(double x) {
    float f = (float)x;
    return ((double)f > x) ? nextafterf(f, -INFINITY) : f;
}
RT_HD float f64_to_f32_up(double x) {
    float f = (float)x;
    return ((double)f < x) ? nextafterf(f, INFINITY) : f;
}
#endif

template <bool QNODES>
RT_HD BoxRay32 make_boxray32_t(const DeviceScene& sc, V3 o, V3 d) {
    BoxRay32 b;
    const double s = fmax(fmax(fabs(o.x), fabs(o.y)), fabs(o.z));
    const double dm = fmax(fmax(fabs(d.x), fabs(d.y)), fabs(d.z));
    // any non-finite origin/direction makes every sphere discriminant NaN (=> miss); skip the tree
    b.enter = (s < 1e30) && (dm < 1e30) && (o.x == o.x) && (o.y == o.y) && (o.z == o.z) && (d.x == d.x) && (d.y == d.y) && (d.z == d.z);
    b.t0 = 0.0;
    const double St = s + sc.bvh_cmax;
    const double E = 64.0 * 2.220446049250313e-16 * St * St;  // bound on the rounding error of the discriminant, in r^2 units
    // min(E / 2 r_min, sqrt(E)) without a division or a square root: sqrt(64 * 2^-52 * St^2) = 2^-23 * St, and
    // bvh_inv2rmin = 0.5 / r_min rounded up on the host
    const double infl = fmin(E * sc.bvh_inv2rmin, St * (0x1.0p-23 * (1.0 + 0x1.0p-50)));
    const double da[3] = {d.x, d.y, d.z};
    float sl[3], fa[3], fb[3];
    uint32_t sn[3] = {0u, 0u, 0u}, sf[3] = {0u, 0u, 0u};
    const double m = infl + St * 0x1.0p-40;
    if (QNODES && b.enter && s < 4096.0 && dm < 1e18) {
        // Origins of scene magnitude -- every ray but the few that start 10^4+ units out (ground-plane hits near the horizon),
        // which take the f64 path below: the clip against the bounds of all sphere boxes and the per-ray constants of the
        // filter are formed in f32.  (In f64 they were a quarter of the traversal kernels' instructions: three f64 reciprocals,
        // an f64 slab test, f64 offsets; and a warp whose lanes split between "inside the bounds" and "outside" paid for
        // two versions.)  Error budget, all absorbed by padding so that the filter stays conservative:
        //   clip        planes widened by mg = 2^-20 St + m in space (roundings of fl32(o), of the differences: < 2^-22 St),
        //               entry shrunk / exit grown by 2^-20 (reciprocal 2^-23, product 2^-24): t0 <= the true entry distance,
        //               and a ray is only declared a miss when it misses the widened box;
        //   reciprocal  fl32(1 / fl32(d)): relative error 2^-23 against 1 / d;
        //   slope       fl32(fl32(cell) * inv): 2^-22 against cell / d; over at most 65535 cells = the extent <= 2 S that is
        //               2^-21 S in space;
        //   offset      g = fl32(q_org - o') (one rounding of the f64 difference; o' = o + d t0 in f64), (g -+ padq) and its
        //               product with inv: 2^-22 relative on a distance of at most 2 S + padq: 2^-21 S in space; the rounding
        //               of the final FMA (2^23 slope - ...) is the half cell the 0.75 cell in padq covers (the 2^23 slope
        //               term itself is exact);
        //   plane FMA   2^-24 on a slab distance, 2^-23 S in space.
        // Together below 5 * 2^-22 S; the pad is S * 2^-18 = 16 * 2^-22 S (a box grows by 5e-4 units where the spheres' radii
        // are 0.2 and up: nothing the walk can measure).  S = St + |d| t0 bounds the magnitude of o' and of the boxes.
        const float mg = f64_to_f32_up(St * 0x1.0p-20 + m);
        const double oa0[3] = {o.x, o.y, o.z};
        float inv[3];
        float te = 0.0f, tx = INFINITY;
        for (int a = 0; a < 3; a++) {
            inv[a] = fabs(da[a]) < 1e-30 ? (float)copysign(1e30, da[a]) : 1.0f / (float)da[a];
            const float of = (float)oa0[a];
            float t1 = ((sc.bvh_lo_f[a] - mg) - of) * inv[a], t2 = ((sc.bvh_hi_f[a] + mg) - of) * inv[a];
            if (t1 > t2) { const float w = t1; t1 = t2; t2 = w; }
            te = fmaxf(te, t1);
            tx = fminf(tx, t2);
        }
        if (!(te * (1.0f - 0x1.0p-20f) <= tx * (1.0f + 0x1.0p-20f))) {  // misses the bounds (or a NaN: 0 * 1e30 cannot occur, inf - inf neither)
            BoxRay32 none = boxray32_none();
            return none;
        }
        const float t0f = te > 0.0f ? te * (1.0f - 0x1.0p-20f) : 0.0f;
        b.t0 = (double)t0f;
        const double oa[3] = {o.x + d.x * b.t0, o.y + d.y * b.t0, o.z + d.z * b.t0};
        const float padf = f64_to_f32_up((St + dm * b.t0) * 0x1.0p-18 + infl);
        for (int a = 0; a < 3; a++) {
            const float iv = inv[a];
            const float cellf = (float)sc.q_cell[a];
            const float padq = padf + 0.75f * cellf * 1.0000002f;
            const bool up = iv > 0.0f;  // the ray runs upward on this axis: it enters boxes through their lower plane
            sl[a] = cellf * iv;
            const float base = 8388608.0f * sl[a];  // exact: a power of two
            const float g = (float)(sc.q_org[a] - oa[a]);
            const float lo_off = fmaf(-(g - padq), iv, base), hi_off = fmaf(-(g + padq), iv, base);
            fa[a] = up ? lo_off : hi_off;
            fb[a] = up ? hi_off : lo_off;
            sn[a] = up ? 0x7610u : 0x7632u;  // node word = lower | upper << 16 (BvhQNode)
            sf[a] = up ? 0x7632u : 0x7610u;
        }
        b.sx = sl[0]; b.sy = sl[1]; b.sz = sl[2]; b.ax = fa[0]; b.ay = fa[1]; b.az = fa[2]; b.bx = fb[0]; b.by = fb[1]; b.bz = fb[2];
        b.snx = sn[0]; b.sny = sn[1]; b.snz = sn[2]; b.sfx = sf[0]; b.sfy = sf[1]; b.sfz = sf[2];
        return b;
    }
    // one reciprocal per axis serves the advance and (rounded to f32) the filter
    double rcp[3];
    float inv[3];
    for (int a = 0; a < 3; a++) {
        rcp[a] = fabs(da[a]) < 1e-300 ? 0.0 : 1.0 / da[a];
        inv[a] = fabs(da[a]) < 1e-30 ? (float)copysign(1e30, da[a]) : (float)rcp[a];
    }
    V3 oo = o;
    if (b.enter) {
        // f64 slab test against the sphere bounds grown by the inflation: where does the ray enter them?
        const double oa[3] = {o.x, o.y, o.z};
        double te = 0.0, tx = (double)INFINITY;
        bool miss = false;
        for (int a = 0; a < 3; a++) {
            const double lo = sc.bvh_lo[a] - m, hi = sc.bvh_hi[a] + m;
            if (fabs(da[a]) < 1e-300) {
                if (oa[a] < lo || oa[a] > hi) miss = true;
            } else {
                double t1 = (lo - oa[a]) * rcp[a], t2 = (hi - oa[a]) * rcp[a];  // within 2 ulp of the quotients; slack below is 2^-40
                if (t1 > t2) { const double w = t1; t1 = t2; t2 = w; }
                if (t1 > te) te = t1;
                if (t2 < tx) tx = t2;
            }
        }
        if (miss || te * (1.0 - 0x1.0p-40) > tx * (1.0 + 0x1.0p-40)) {
            b.enter = false;
        } else if (te > 0.0) {
            b.t0 = te * (1.0 - 0x1.0p-40);
            oo = v3(o.x + d.x * b.t0, o.y + d.y * b.t0, o.z + d.z * b.t0);
        }
    }
    const double s2 = fmax(fmax(fabs(oo.x), fabs(oo.y)), fabs(oo.z));
    const double S = s2 + sc.bvh_cmax;
    const double pad = S * 0x1.0p-20 + infl + (s + dm * b.t0) * 0x1.0p-50;
    const double oa[3] = {oo.x, oo.y, oo.z};
    for (int a = 0; a < 3; a++) {
        const double iv = (double)inv[a];
        if (!QNODES) {
            sl[a] = inv[a];
            fa[a] = (float)((oa[a] + pad) * iv);
            fb[a] = (float)((oa[a] - pad) * iv);
        } else {
            // grid form (see qplane).  The 2^23 * slope term of the offset is built from the ROUNDED slope, so it cancels
            // exactly in the FFMA; what is left is the rounding of the offset itself, < 0.51 cells (|offset| <
            // (2^23 + 2^16) slope, o' being inside the bounds): 0.75 more cells of padding.
            const double padq = pad + 0.75 * sc.q_cell[a];
            const bool up = inv[a] > 0.0f;  // the ray runs upward on this axis: it enters boxes through their lower plane
            sl[a] = (float)(sc.q_cell[a] * iv);
            const double base = 8388608.0 * (double)sl[a];
            const double lo_off = base - (sc.q_org[a] - oa[a] - padq) * iv, hi_off = base - (sc.q_org[a] - oa[a] + padq) * iv;
            fa[a] = (float)(up ? lo_off : hi_off);
            fb[a] = (float)(up ? hi_off : lo_off);
            sn[a] = up ? 0x7610u : 0x7632u;  // node word = lower | upper << 16 (BvhQNode)
            sf[a] = up ? 0x7632u : 0x7610u;
        }
    }
    b.sx = sl[0]; b.sy = sl[1]; b.sz = sl[2]; b.ax = fa[0]; b.ay = fa[1]; b.az = fa[2]; b.bx = fb[0]; b.by = fb[1]; b.bz = fb[2];
    b.snx = sn[0]; b.sny = sn[1]; b.snz = sn[2]; b.sfx = sf[0]; b.sfy = sf[1]; b.sfz = sf[2];
    return b;
}
RT_HD BoxRay32 make_boxray32(const DeviceScene& sc, V3 o, V3 d) {
    return sc.bvh_qnodes ? make_boxray32_t<true>(sc, o, d) : make_boxray32_t<false>(sc, o, d);
}

// can the padded box be hit within [0, bound]?  tn = conservative entry distance (>= 0, from the advanced origin)
RT_HD bool box_hit32(const BoxRay32& b, const float* lo, const float* hi, float bound, float& tn) {
    const float lx = fmaf(lo[0], b.sx, -b.ax), hx = fmaf(hi[0], b.sx, -b.bx);
    const float ly = fmaf(lo[1], b.sy, -b.ay), hy = fmaf(hi[1], b.sy, -b.by);
    const float lz = fmaf(lo[2], b.sz, -b.az), hz = fmaf(hi[2], b.sz, -b.bz);
    tn = fmaxf(fmaxf(fminf(lx, hx), fminf(ly, hy)), fmaxf(fminf(lz, hz), 0.0f));
    const float tf = fminf(fminf(fmaxf(lx, hx), fmaxf(ly, hy)), fmaxf(lz, hz));
    return tn <= tf && tn <= bound;
}

// ---- 32-byte nodes: box planes on the scene's 16-bit grid -------------------------------------------------
// largest q with org + q*cell <= v (lower bounds) / smallest q with org + q*cell >= v (upper bounds), evaluated with
// the same f64 expression the check uses; `ok` is cleared when v lies off the grid (the caller keeps the f32 nodes)
RT_HD uint16_t quant_lo(float v, double org, double cell, bool& ok) {
    double q = floor(((double)v - org) / cell);
    if (!(q >= 0.0)) { q = 0.0; }
    if (q > 65535.0) q = 65535.0;
    while (q > 0.0 && org + q * cell > (double)v) q -= 1.0;
    if (!(org + q * cell <= (double)v)) ok = false;
    return (uint16_t)q;
}
RT_HD uint16_t quant_hi(float v, double org, double cell, bool& ok) {
    double q = ceil(((double)v - org) / cell);
    if (!(q >= 0.0)) { q = 0.0; }
    if (q > 65535.0) q = 65535.0;
    while (q < 65535.0 && org + q * cell < (double)v) q += 1.0;
    if (!(org + q * cell >= (double)v)) ok = false;
    return (uint16_t)q;
}
RT_HD BvhQNode quantise_node(const BvhNode& n, const double* org, const double* cell, bool& ok) {
    BvhQNode o;
    for (int a = 0; a < 3; a++) {  // per child and axis: lower plane in the low half-word, upper plane in the high one
        o.q[2 * a] = quant_lo(n.lo0[a], org[a], cell[a], ok);
        o.q[2 * a + 1] = quant_hi(n.hi0[a], org[a], cell[a], ok);
        o.q[6 + 2 * a] = quant_lo(n.lo1[a], org[a], cell[a], ok);
        o.q[6 + 2 * a + 1] = quant_hi(n.hi1[a], org[a], cell[a], ok);
    }
    o.c0 = n.c0; o.c1 = n.c1;
    return o;
}

// The filter never decodes a grid coordinate: a plane at org + q*cell has slab distance
//   ((org + q*cell) - (o' +- pad)) * inv  =  (2^23 + q) * (cell*inv)  -  (2^23 * cell*inv - (org - (o' +- pad)) * inv),
// so with the per-ray slope cell*inv and the bracket as offset it is one FFMA on the float 2^23 + q, whose bit pattern
// is 0x4B000000 | q -- one byte permute of the node word.  The permute's selector also picks WHICH plane (lower or upper
// half-word), so the ray's direction signs are spent there and the slab test needs no min/max per axis:
// entry = max of the three entry planes, exit = min of the three exit planes.
struct QNodeWords { uint32_t w[8]; };  // child 0: x y z | child 1: x y z | c0 c1
#if defined(__CUDA_ARCH__)
RT_HD QNodeWords load_qnode(const BvhQNode* p) {
    QNodeWords n;
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(n.w[0]), "=r"(n.w[1]), "=r"(n.w[2]), "=r"(n.w[3]), "=r"(n.w[4]), "=r"(n.w[5]), "=r"(n.w[6]), "=r"(n.w[7]) : "l"(p));
    return n;
}
RT_HD float qplane(uint32_t word, uint32_t sel) {  // prmt directly: __byte_perm would mask the selector on every call
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(word), "r"(0x4B000000u), "r"(sel));
    return __uint_as_float(r);
}
#else
RT_HD QNodeWords load_qnode(const BvhQNode* p) {
    QNodeWords n;
    memcpy(n.w, p, 32);
    return n;
}
RT_HD float qplane(uint32_t word, uint32_t sel) {  // the two selectors in use: 0x7610 lower, 0x7632 upper half-word
    return (float)(8388608u + (sel == 0x7610u ? (word & 0xFFFFu) : (word >> 16)));
}
#endif
RT_HD bool box_hit_q(const BoxRay32& b, uint32_t wx, uint32_t wy, uint32_t wz, float bound, float& tn) {
    const float nx = fmaf(qplane(wx, b.snx), b.sx, -b.ax), fx = fmaf(qplane(wx, b.sfx), b.sx, -b.bx);
    const float ny = fmaf(qplane(wy, b.sny), b.sy, -b.ay), fy = fmaf(qplane(wy, b.sfy), b.sy, -b.by);
    const float nz = fmaf(qplane(wz, b.snz), b.sz, -b.az), fz = fmaf(qplane(wz, b.sfz), b.sz, -b.bz);
    tn = fmaxf(fmaxf(nx, ny), fmaxf(nz, 0.0f));
    const float tf = fminf(fminf(fx, fy), fz);
    return tn <= tf && tn <= bound;
}

#if defined(__CUDA_ARCH__)
// the 64-byte node in two 256-bit loads (sm_100 LDG.E.ENL2.256): a divergent node fetch costs the L1 data pipe one
// wavefront per lane per load, so two loads instead of four halve the pipe's work
RT_HD BvhNode load_node(const BvhNode* p) {
    BvhNode n;
    float a0, a1, a2, a3, a4, a5, a6, a7, b0, b1, b2, b3, b4, b5, b6, b7;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a0), "=f"(a1), "=f"(a2), "=f"(a3), "=f"(a4), "=f"(a5), "=f"(a6), "=f"(a7) : "l"(p));
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(b0), "=f"(b1), "=f"(b2), "=f"(b3), "=f"(b4), "=f"(b5), "=f"(b6), "=f"(b7) : "l"(reinterpret_cast<const char*>(p) + 32));
    n.lo0[0] = a0; n.lo0[1] = a1; n.lo0[2] = a2; n.hi0[0] = a3;
    n.hi0[1] = a4; n.hi0[2] = a5; n.lo1[0] = a6; n.lo1[1] = a7;
    n.lo1[2] = b0; n.hi1[0] = b1; n.hi1[1] = b2; n.hi1[2] = b3;
    n.c0 = __float_as_int(b4); n.c1 = __float_as_int(b5);
    return n;
}
#else
RT_HD BvhNode load_node(const BvhNode* p) { return *p; }
#endif

// traversal stack: internal nodes (>= 0) and leaves (~slot < 0) alike, each with its conservative entry distance
// Postponed far children: (node, entry distance) packed in one 8-byte word so that a push is one store and a pop
// one load.  The array is indexed dynamically and therefore lives in local memory; the depth `sp` is passed beside it
// so that it stays in a register.
struct TravStack {
    unsigned long long e[BVH_STACK];
};
RT_HD unsigned long long trav_pack(int32_t node, float t) {
    uint32_t tb;
    memcpy(&tb, &t, 4);
    return (unsigned long long)(uint32_t)node | ((unsigned long long)tb << 32);
}

RT_HD void trav_push(TravStack& st, int& sp, int32_t node, float t) { RT_CHECK(sp >= 0 && sp < BVH_STACK); st.e[sp++] = trav_pack(node, t); }
RT_HD int32_t trav_pop(const TravStack& st, int& sp, float bound) {
    while (sp > 0) {
        const unsigned long long e = st.e[--sp];
        const uint32_t tb = (uint32_t)(e >> 32);
        float t;
        memcpy(&t, &tb, 4);
        if (t <= bound) return (int32_t)(uint32_t)e;
    }
    return TRAV_EMPTY;
}
// Shadow queries: the bound (the light's distance) never shrinks during the walk, so a child that passed the box test when
// it was postponed still passes when it is popped -- 4-byte entries, no entry distance, no test on pop.
struct TravStackFixed {
    int32_t e[BVH_STACK];
};
RT_HD void trav_push(TravStackFixed& st, int& sp, int32_t node, float) { RT_CHECK(sp >= 0 && sp < BVH_STACK); st.e[sp++] = node; }
RT_HD int32_t trav_pop(const TravStackFixed& st, int& sp, float) { return sp > 0 ? st.e[--sp] : TRAV_EMPTY; }

// Walk internal nodes from `cur` (near child first, far child pushed) until the next thing to do is a leaf test
// (returns ~slot < 0) or nothing is left (TRAV_EMPTY).  Leaves are NOT tested here: postponing them keeps the lanes
// of a warp in the same code (box tests together, sphere tests together).
// `max_steps` bounds the walk (the kernels come back to refill finished lanes; an unfinished walk returns cur >= 0).
// ---- wide nodes (shadow queries) --------------------------------------------------------------------------
// Measured on the B200 (profiles/r02_v1_wide_ab.txt): 9.0 instead of 16.5 visits per shadow ray at 10k spheres, 14.9
// instead of 26.2 at 1M; k_shadow_bvh 6.37 -> 6.19 ms and 7.72 -> 7.47 ms.  The closest-hit counterpart (nearest of four
// boxes first, entry distances on the stack) halved the visits too but its sorting network made the launches SLOWER
// (4.96 -> 5.40 ms, 5.12 -> 5.28 ms): removed.
// One half of wide node i from binary node `parent`: the child `which` of it is either an internal node (its own two
// children become the half) or a leaf (the half holds the leaf itself and an empty slot that no ray can hit).
RT_HD BvhQNode widen_half(const BvhQNode& parent, int which, const BvhQNode* qnodes) {
    const int32_t c = which ? parent.c1 : parent.c0;
    if (c >= 0) return qnodes[c];
    BvhQNode h;
    for (int a = 0; a < 3; a++) {
        h.q[2 * a] = parent.q[6 * which + 2 * a];
        h.q[2 * a + 1] = parent.q[6 * which + 2 * a + 1];
        h.q[6 + 2 * a] = 0xFFFFu;  // lower plane above the upper one on every axis
        h.q[6 + 2 * a + 1] = 0u;
    }
    h.c0 = c;
    h.c1 = TRAV_EMPTY;
    return h;
}
RT_HD BvhWNode widen_node(const BvhQNode* qnodes, uint32_t i) {
    BvhWNode w;
    const BvhQNode p = qnodes[i];
    w.h[0] = widen_half(p, 0, qnodes);
    w.h[1] = widen_half(p, 1, qnodes);
    return w;
}
// up to three hit children are postponed per visit, one visit per two binary levels (3 * ceil(95 / 2)), plus one
constexpr int BVH_STACK_WIDE = 148;
struct TravStackWide {
    int32_t e[BVH_STACK_WIDE];
};
RT_HD int32_t trav_pop(const TravStackWide& st, int& sp, float) { return sp > 0 ? st.e[--sp] : TRAV_EMPTY; }
RT_HD void trav_push_if(TravStackWide& st, int& sp, bool hit, int32_t node) {
    RT_CHECK(!hit || (sp >= 0 && sp < BVH_STACK_WIDE));
    if (hit) st.e[sp++] = node;  // two predicated instructions, no branch
}

// The walk of a shadow query over wide nodes.  No order among the children: the query only asks whether anything is
// hit within the bound, which never shrinks.  Returns like trav_descend (leaf to test, TRAV_EMPTY, or an unfinished walk).
RT_HD int32_t trav_descend_wide(const DeviceScene& sc, const BoxRay32& br, int32_t cur, float bound, TravStackWide& st, int& sp, TraverseCount& tc,
                                int max_steps = 0x7FFFFFFF) {
    while (cur >= 0 && max_steps-- > 0) {
        RT_CHECK((uint32_t)cur < sc.n_bvh_nodes);
        const BvhQNode* p = sc.bvh_wnodes[cur].h;
        const QNodeWords a = load_qnode(p), b = load_qnode(p + 1);
        float t;
        const int32_t c0 = (int32_t)a.w[6], c1 = (int32_t)a.w[7], c2 = (int32_t)b.w[6], c3 = (int32_t)b.w[7];
        const bool h0 = box_hit_q(br, a.w[0], a.w[1], a.w[2], bound, t) && c0 != TRAV_EMPTY;
        const bool h1 = box_hit_q(br, a.w[3], a.w[4], a.w[5], bound, t) && c1 != TRAV_EMPTY;
        const bool h2 = box_hit_q(br, b.w[0], b.w[1], b.w[2], bound, t) && c2 != TRAV_EMPTY;
        const bool h3 = box_hit_q(br, b.w[3], b.w[4], b.w[5], bound, t) && c3 != TRAV_EMPTY;
        tc.nodes++;
        // The first hit child (in slot order: the query is unordered) is walked next straight out of its register; the other
        // hit children are pushed; only a visit that hits nothing pops.  (Pushing all four and popping the top put a local-
        // memory store -> load round trip on the critical path of EVERY visit: 12 % of the kernel's stall samples sat on the
        // popped word, as many as on the node fetch itself -- profiles/r02_v3_ncu_k_shadow_bvh.txt.)  Still no branch on
        // which of the four were hit: three predicated stores, selects, one predicated load.
        const bool any01 = h0 || h1, any012 = any01 || h2;
        trav_push_if(st, sp, h3 && any012, c3);
        trav_push_if(st, sp, h2 && any01, c2);
        trav_push_if(st, sp, h1 && h0, c1);
        const int32_t first = h0 ? c0 : (h1 ? c1 : (h2 ? c2 : c3));
        if (any012 || h3) cur = first;
        else cur = trav_pop(st, sp, bound);
    }
    // the walk ends at a leaf: its sphere record is fetched while the other lanes of the warp finish their walks (the leaf
    // tests of a warp run together, afterwards)
    if (cur < 0 && cur != TRAV_EMPTY) RT_PREFETCH_L1(sc.bvh_sph + ~cur);
    return cur;
}

#ifndef RT_DESCEND_MIN_LANES
#define RT_DESCEND_MIN_LANES 0  // leave the bounded walk when fewer lanes are still in it; 8 / 12 / 16 measured: closest-hit launches
#endif                          // 4.94 -> 5.02 / 5.12 / 5.28 ms (10k), 5.10 -> 5.75+ ms (1M); only the 1M shadow launch gained (7.73 -> 7.40)
template <bool QNODES, typename STACK>
RT_HD int32_t trav_descend(const DeviceScene& sc, const BoxRay32& br, int32_t cur, float bound, STACK& st, int& sp, TraverseCount& tc,
                           int max_steps = 0x7FFFFFFF) {
    while (cur >= 0 && max_steps-- > 0) {
        RT_CHECK((uint32_t)cur < sc.n_bvh_nodes);
        float t0, t1;
        bool h0, h1;
        int32_t c0, c1;
        if (QNODES) {
            const QNodeWords n = load_qnode(sc.bvh_qnodes + cur);
            h0 = box_hit_q(br, n.w[0], n.w[1], n.w[2], bound, t0);
            h1 = box_hit_q(br, n.w[3], n.w[4], n.w[5], bound, t1);
            c0 = (int32_t)n.w[6]; c1 = (int32_t)n.w[7];
        } else {
            const BvhNode n = load_node(sc.bvh_nodes + cur);
            h0 = box_hit32(br, n.lo0, n.hi0, bound, t0);
            h1 = box_hit32(br, n.lo1, n.hi1, bound, t1);
            c0 = n.c0; c1 = n.c1;
        }
        tc.nodes++;
        // nearer hit child next, the other one postponed; selects instead of branches: the lanes of a warp take all
        // four cases at once
        const bool both = h0 && h1;
        const bool second = h1 && (!h0 || t1 < t0);
        const int32_t nearc = second ? c1 : c0;
        if (both) trav_push(st, sp, second ? c0 : c1, second ? t0 : t1);
        if (h0 || h1) cur = nearc;
        else cur = trav_pop(st, sp, bound);
#if defined(__CUDA_ARCH__) && RT_DESCEND_MIN_LANES > 0
        // few lanes left walking: go back to the warp (leaf tests, refill) instead of finishing the bounded walk with the
        // others idle.  A scheduling heuristic only -- what the mask holds cannot change a result.
        if (__popc(__activemask()) < RT_DESCEND_MIN_LANES) break;
#endif
    }
    if (cur < 0 && cur != TRAV_EMPTY) RT_PREFETCH_L1(sc.bvh_sph + ~cur);  // see trav_descend_wide
    return cur;
}

// filter bound (distance from the advanced origin) that no hit closer than `t` of the exact test can exceed
RT_HD float closest_bound32(double t, double t0) { return f64_to_f32_up((t - t0) * RT_BVH_TSLACK + (fabs(t) + t0) * 1e-12); }
RT_HD float shadow_bound32(double range2, double t0) {
    if (range2 < 0.0) return INFINITY;
    if (t0 == 0.0) {
        // the usual case (origin inside the bounds): an upper bound of sqrt(range2) * (1 + 2e-12) without the f64 square root --
        // range2 rounded up to f32, its f32 square root (correctly rounded: within 2^-24), two parts in 2^23 on top
        const float r = sqrtf(f64_to_f32_up(range2));
        return r + r * 0x1.0p-22f;
    }
    const double r = sqrt(range2);
    return f64_to_f32_up((r - t0) * RT_BVH_TSLACK + (r + t0) * 1e-12);
}


// exact test of Morton-slot k against the running closest hit; returns the new filter bound
RT_HD void leaf_closest(const DeviceScene& sc, const RayQ& r, int32_t k, Closest& best, TraverseCount& tc) {
    RT_CHECK(k >= 0 && (uint32_t)k < sc.n_spheres);
    const SphereRec* s = sc.bvh_sph + k;
    double t;
    tc.prims++;
    if (sphere_hit(r, RT_LDG(&s->cx), RT_LDG(&s->cy), RT_LDG(&s->cz), RT_LDG(&s->r2), t)) closest_update(best, t, RT_LDG(sc.bvh_obj + k));
}
RT_HD bool leaf_shadow(const DeviceScene& sc, const RayQ& r, int32_t k, double range2, TraverseCount& tc) {
    RT_CHECK(k >= 0 && (uint32_t)k < sc.n_spheres);
    const SphereRec* s = sc.bvh_sph + k;
    double t;
    tc.prims++;
    return sphere_hit(r, RT_LDG(&s->cx), RT_LDG(&s->cy), RT_LDG(&s->cz), RT_LDG(&s->r2), t) && shadow_hit_counts(t, range2);
}

// Scene::intersect through the BVH (one ray, start to finish; the kernels run the same pieces lane by lane)
RT_HD void closest_bvh(const DeviceScene& sc, const RayQ& r, int32_t& id, double& t, TraverseCount& tc) {
    Closest best = closest_none();
    planes_closest(sc, r, best);
    if (!best.nan && sc.n_spheres > 0) {
        const BoxRay32 br = make_boxray32(sc, r.o, r.d);
        if (br.enter) {
            TravStack st;
            int sp = 0;
            int32_t cur = 0;
            float bound = closest_bound32(best.t, br.t0);
            while (true) {
                cur = sc.bvh_qnodes ? trav_descend<true>(sc, br, cur, bound, st, sp, tc) : trav_descend<false>(sc, br, cur, bound, st, sp, tc);
                if (cur == TRAV_EMPTY) break;
                leaf_closest(sc, r, ~cur, best, tc);
                bound = closest_bound32(best.t, br.t0);
                cur = trav_pop(st, sp, bound);
            }
        }
    }
    id = closest_id(best);
    t = best.t;
}
RT_HD bool shadow_bvh(const DeviceScene& sc, const RayQ& r, double range2, TraverseCount& tc) {
    Shadow s; s.nan = false; s.any = false;
    planes_shadow(sc, r, range2, s);
    if (!s.nan && !s.any && sc.n_spheres > 0) {
        const BoxRay32 br = make_boxray32(sc, r.o, r.d);
        if (br.enter && sc.bvh_wnodes) {
            TravStackWide st;
            int sp = 0;
            int32_t cur = 0;
            const float bound = shadow_bound32(range2, br.t0);
            while (true) {
                cur = trav_descend_wide(sc, br, cur, bound, st, sp, tc);
                if (cur == TRAV_EMPTY) break;
                if (leaf_shadow(sc, r, ~cur, range2, tc)) { s.any = true; break; }
                cur = trav_pop(st, sp, bound);
            }
        } else if (br.enter) {
            TravStackFixed st;
            int sp = 0;
            int32_t cur = 0;
            const float bound = shadow_bound32(range2, br.t0);
            while (true) {
                cur = sc.bvh_qnodes ? trav_descend<true>(sc, br, cur, bound, st, sp, tc) : trav_descend<false>(sc, br, cur, bound, st, sp, tc);
                if (cur == TRAV_EMPTY) break;
                if (leaf_shadow(sc, r, ~cur, range2, tc)) { s.any = true; break; }
                cur = trav_pop(st, sp, bound);
            }
        }
    }
    return shadow_result(s, range2);
}

// plain scan over global memory (emulation harness and tiny batches; the kernel stages tiles through smem)
RT_HD void closest_linear(const DeviceScene& sc, const RayQ& r, int32_t& id, double& t) {
    Closest best = closest_none();
    planes_closest(sc, r, best);
    if (!best.nan) {
        for (uint32_t k = 0; k < sc.n_spheres; k++) {
            const SphereRec* s = sc.sph + k;
            double th;
            if (sphere_hit(r, RT_LDG(&s->cx), RT_LDG(&s->cy), RT_LDG(&s->cz), RT_LDG(&s->r2), th)) closest_update(best, th, RT_LDG(sc.sph_obj + k));
        }
    }
    id = closest_id(best);
    t = best.t;
}
RT_HD bool shadow_linear(const DeviceScene& sc, const RayQ& r, double range2) {
    Shadow s; s.nan = false; s.any = false;
    planes_shadow(sc, r, range2, s);
    if (!s.nan && !s.any) {
        for (uint32_t k = 0; k < sc.n_spheres; k++) {
            const SphereRec* q = sc.sph + k;
            double th;
            if (sphere_hit(r, RT_LDG(&q->cx), RT_LDG(&q->cy), RT_LDG(&q->cz), RT_LDG(&q->r2), th) && shadow_hit_counts(th, range2)) { s.any = true; break; }
        }
    }
    return shadow_result(s, range2);
}

// normal the shape's `intersect` returns (src/shapes.rs:72,79 / :108)
RT_HD V3 shape_normal(const DeviceScene& sc, int32_t obj, V3 o, V3 d, double t) {
    uint32_t sh = RT_LDG(sc.obj_shape + obj);
    if (sh & 0x80000000u) {
        const PlaneRec* p = sc.planes + (sh & 0x7FFFFFFFu);
        return v3(RT_LDG(&p->n[0]), RT_LDG(&p->n[1]), RT_LDG(&p->n[2]));
    }
    const SphereRec* s = sc.sph + sh;
    V3 c = v3(RT_LDG(&s->cx), RT_LDG(&s->cy), RT_LDG(&s->cz));
    return normalize((o + d * t) - c);
}

// ---------------------------------------------------------------------------------------------------------
// backgrounds + bilinear texture (src/raytrace.rs:228-256, src/texture.rs:26-58, src/color.rs:611-613)
// ---------------------------------------------------------------------------------------------------------
RT_HD double clamp01(double x) { return x < 0.0 ? 0.0 : (x > 1.0 ? 1.0 : x); }
RT_HD uint32_t f64_as_u32(double x) {  // Rust `as u32`: saturating, NaN -> 0
    if (!(x > 0.0)) return 0u;
    if (x >= 4294967295.0) return 4294967295u;
    return (uint32_t)x;
}
RT_HD C3 tex_at(const DeviceScene& sc, const TextureRec& t, uint32_t x, uint32_t y) {
    RT_CHECK(x < t.width && y < t.height);
    size_t idx = 3 * ((size_t)x + (size_t)y * t.width);
    return c3(RT_LDG(sc.srgb_values + RT_LDG(t.data + idx)), RT_LDG(sc.srgb_values + RT_LDG(t.data + idx + 1)), RT_LDG(sc.srgb_values + RT_LDG(t.data + idx + 2)));
}
RT_HD C3 tex_sample(const DeviceScene& sc, const TextureRec& t, double xs, double ys) {
    double x = clamp01(xs) * (double)(t.width - 1);
    double y = clamp01(ys) * (double)(t.height - 1);
    uint32_t x0 = f64_as_u32(x);
    uint32_t x1 = x0 >= t.width - 1 ? t.width - 1 : x0 + 1;
    double xx = x - (double)x0;
    uint32_t y0 = f64_as_u32(y);
    uint32_t y1 = y0 >= t.height - 1 ? t.height - 1 : y0 + 1;
    double yy = y - (double)y0;
    C3 cx0 = tex_at(sc, t, x0, y0) * (1.0 - yy) + tex_at(sc, t, x0, y1) * yy;
    C3 cx1 = tex_at(sc, t, x1, y0) * (1.0 - yy) + tex_at(sc, t, x1, y1) * yy;
    return cx0 * (1.0 - xx) + cx1 * xx;
}
RT_HD C3 background_color(const DeviceScene& sc, V3 d) {
    if (sc.bg_kind == BG_SOLID) return c3(sc.bg_color);
    double ax = fabs(d.x), ay = fabs(d.y), az = fabs(d.z);
    if (ax > az && ax > ay) {
        double px = -d.z / d.x, py = -d.y / ax;
        return tex_sample(sc, sc.faces[d.x > 0.0 ? 0 : 1], px * 0.5 + 0.5, py * 0.5 + 0.5);
    }
    if (ay > ax && ay > az) {
        double px = d.x / ay, py = d.z / d.y;
        return tex_sample(sc, sc.faces[d.y > 0.0 ? 2 : 3], px * 0.5 + 0.5, py * 0.5 + 0.5);
    }
    if (az > ax && az > ay) {
        double px = d.x / d.z, py = -d.y / az;
        return tex_sample(sc, sc.faces[d.z > 0.0 ? 4 : 5], px * 0.5 + 0.5, py * 0.5 + 0.5);
    }
    return c3(0.0, 0.0, 0.0);
}

// ---------------------------------------------------------------------------------------------------------
// quantiser (src/color.rs:593-600): smallest i with val < SRGB_AVERAGE[i], else 255.  The table is strictly
// increasing, so a binary search returns the same index as the reference's linear scan (NaN -> 255).
// ---------------------------------------------------------------------------------------------------------
RT_HD uint8_t to_srgb(const double* avg, double val) {
    int lo = 0, hi = 255;  // answer in [lo, hi]
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (val < avg[mid]) hi = mid;
        else lo = mid + 1;
    }
    return (uint8_t)lo;
}

// ---------------------------------------------------------------------------------------------------------
// shading (src/raytrace.rs:30-226) in path-throughput form.
//
// The reference's colour recursion is linear in every child colour, so a node's contribution to its pixel is
// throughput (x) local terms, and each child inherits throughput (x) weight.  shade_begin does the
// part before the light loop, shade_light one iteration of the light loop, shade_child one spawned ray.
// ---------------------------------------------------------------------------------------------------------
struct ShadeCtx {
    V3 d;           // incoming direction
    V3 pt, normal;  // hit point, viewer-facing normal
    C3 T;           // throughput
    C3 mdiff, mspec;
    double exponent, sig, ssig, fresnel;
    double ndv, omf;
    V3 refr;  // normalised refraction direction
    uint64_t key;
    uint32_t depth, samples;
    int32_t kind;
    bool alive;  // hit, depth <= max_depth
    bool diffuse, specular, lit, has_refr;
};

struct ShadowRay {
    V3 o, d;
    double range2;
    C3 c;
};
struct ChildRay {
    V3 o, d;
    C3 T;
    double sig;
    uint64_t key;
};

// Returns the immediate pixel contribution (throughput x ambient, or throughput x background on a miss).
RT_HD C3 shade_begin(const DeviceScene& sc, const FrameParams& fp, V3 o, V3 d, C3 T, double sig, uint64_t key, uint32_t depth, int32_t hit_id,
                     double hit_t, ShadeCtx& s) {
    s.alive = false; s.lit = false; s.diffuse = false; s.specular = false; s.has_refr = false;
    s.kind = -1; s.samples = 0;
    if (hit_id < 0) return T * background_color(sc, d);
    RT_CHECK((uint32_t)hit_id < sc.n_objects);
    const MaterialRec* m = sc.materials + RT_LDG(sc.obj_material + hit_id);
    const int32_t kind = RT_LDG(&m->kind);
    C3 ambient = kind == MAT_TRANSPARENT ? c3(0.0, 0.0, 0.0) : c3(RT_LDG(&m->ambient[0]), RT_LDG(&m->ambient[1]), RT_LDG(&m->ambient[2]));
    C3 add = T * ambient;
    if (depth > fp.max_depth) return add;  // src/raytrace.rs:33,72,126,172
    s.alive = true;
    s.kind = kind;
    s.d = d; s.T = T; s.sig = sig; s.key = key; s.depth = depth;
    s.mdiff = c3(RT_LDG(&m->diffuse[0]), RT_LDG(&m->diffuse[1]), RT_LDG(&m->diffuse[2]));
    s.mspec = c3(RT_LDG(&m->specular[0]), RT_LDG(&m->specular[1]), RT_LDG(&m->specular[2]));
    s.exponent = RT_LDG(&m->exponent);
    s.samples = RT_LDG(&m->samples);
    s.ssig = significance(s.mspec);
    s.pt = o + d * hit_t;
    V3 rn = shape_normal(sc, hit_id, o, d, hit_t);
    double nd = dot(rn, d);
    s.normal = nd > 0.0 ? -rn : rn;
    s.fresnel = 1.0;
    s.omf = 0.0;
    s.ndv = 0.0;
    if (kind == MAT_PHONG || kind == MAT_INDIRECT) {
        s.diffuse = significance(s.mdiff) * sig > fp.min_significance;
        s.specular = s.ssig * sig > fp.min_significance;
        s.lit = s.diffuse || s.specular;
    } else {
        double ior = RT_LDG(&m->ior);
        double r0 = (ior - 1.0) / (ior + 1.0);
        r0 = r0 * r0;
        if (kind == MAT_FRESNEL) {
            double omcos = 1.0 - fabs(nd);
            double omcos2 = omcos * omcos;
            s.fresnel = clamp_one(r0 + (1.0 - r0) * omcos2 * omcos2 * omcos);
            s.diffuse = significance(s.mdiff) * sig > fp.min_significance;
            s.specular = s.ssig * s.fresnel * sig > fp.min_significance;
            s.lit = s.diffuse || s.specular;
        } else {  // MAT_TRANSPARENT, src/raytrace.rs:169-226
            s.ndv = dot(s.normal, d);
            double n = nd > 0.0 ? ior : 1.0 / ior;
            double sin2 = n * n * (1.0 - nd * nd);
            bool has = sin2 < 1.0;
            V3 refract = v3(0.0, 0.0, 0.0);
            if (has) {
                double cs = sqrt(1.0 - sin2);
                refract = d * n - s.normal * (n * fabs(nd) + cs);
            }
            double omcos = nd > 0.0 ? (has ? 1.0 - dot(s.normal, refract) : 0.0) : 1.0 - fabs(nd);
            double omcos2 = omcos * omcos;
            s.fresnel = has ? clamp_one(r0 + (1.0 - r0) * omcos2 * omcos2 * omcos) : 1.0;
            s.specular = s.ssig * s.fresnel * sig > fp.min_significance;
            s.lit = s.specular;
            if (s.fresnel < 1.0 && has) {
                s.has_refr = true;
                s.omf = clamp_one(1.0 - s.fresnel);
                s.refr = normalize(refract);
            }
        }
    }
    return add;
}

// one iteration of `for light in &scene.lights` (src/raytrace.rs:39-58 and its three siblings).
// Returns false when this hit casts no shadow ray towards light li.
RT_HD bool shade_light(const DeviceScene& sc, const ShadeCtx& s, uint32_t li, ShadowRay& out) {
    if (!s.alive || !s.lit) return false;
    const LightRec* l = sc.lights + li;
    const int32_t lk = RT_LDG(&l->kind);
    V3 la = v3(RT_LDG(&l->a[0]), RT_LDG(&l->a[1]), RT_LDG(&l->a[2]));
    V3 ldir;
    double range2;
    if (lk == LIGHT_DIRECTIONAL) {  // src/scene.rs:136-138
        ldir = -la;
        range2 = -1.0;
    } else {
        V3 loc = la;
        if (lk == LIGHT_AREA) {  // src/scene.rs:151-155
            double u1 = draw_f64(s.key, 2 + 2 * (uint64_t)li);
            double u2 = draw_f64(s.key, 3 + 2 * (uint64_t)li);
            V3 s1 = v3(RT_LDG(&l->s1[0]), RT_LDG(&l->s1[1]), RT_LDG(&l->s1[2]));
            V3 s2 = v3(RT_LDG(&l->s2[0]), RT_LDG(&l->s2[1]), RT_LDG(&l->s2[2]));
            loc = la + s1 * u1 + s2 * u2;
        }
        V3 diff = v3(loc.x - s.pt.x, loc.y - s.pt.y, loc.z - s.pt.z);  // src/scene.rs:124-126
        ldir = normalize(diff);
        range2 = sqnorm(diff);
    }
    C3 lcol = c3(RT_LDG(&l->color[0]), RT_LDG(&l->color[1]), RT_LDG(&l->color[2]));
    C3 c = c3(0.0, 0.0, 0.0);
    if (s.diffuse) c = c + s.T * (s.mdiff * lcol * clamp_zero(dot(ldir, s.normal)) * RT_FRAC_1_PI);
    if (s.specular) {
        double pw = pow(clamp_zero(dot(s.normal, normalize(ldir - s.d))), s.exponent);
        C3 sl = s.mspec * lcol;
        if (s.kind == MAT_FRESNEL || s.kind == MAT_TRANSPARENT) sl = sl * s.fresnel;
        c = c + s.T * (sl * pw);
    }
    out.o = s.pt + ldir * 0.00001;
    out.d = ldir;
    out.range2 = range2;
    out.c = c;
    return true;
}

// does child slot k of this hit spawn a ray?  (the cheap part of shade_child: lets the kernel reserve queue slots for
// all children of a warp with one atomic before doing the arithmetic)
RT_HD bool shade_child_wanted(const ShadeCtx& s, uint32_t k) {
    if (!s.alive) return false;
    if (s.kind == MAT_INDIRECT) return s.lit && k < s.samples;
    if (k == 0) return s.specular;
    return k == 1 && s.kind == MAT_TRANSPARENT && s.has_refr;
}

// child slot k of this hit: 0 = reflection (Phong/Fresnel/Transparent), 1 = refraction (Transparent),
// k < samples = hemisphere sample k (IndirectPhong).  Returns false when the slot spawns nothing.
RT_HD bool shade_child(const DeviceScene& sc, const ShadeCtx& s, uint32_t k, ChildRay& out) {
    if (!s.alive) return false;
    if (s.kind == MAT_INDIRECT) {  // src/raytrace.rs:98-117
        if (!s.lit || k >= s.samples) return false;
        const uint64_t dim0 = 2 + 2 * (uint64_t)sc.n_lights + 2 * (uint64_t)k;
        double r1 = draw_f64(s.key, dim0) * 2.0 - 1.0;
        double r2 = draw_f64(s.key, dim0 + 1) * (2.0 * RT_PI);
        double sin_theta = 1.0 - r1 * r1;  // sic (src/raytrace.rs:103)
        double x = sin_theta * cos(r2);
        double z = sin_theta * sin(r2);
        V3 d0 = v3(x, r1, z);
        V3 dir = dot(d0, s.normal) >= 0.0 ? d0 : -d0;
        double fac = (double)s.samples * 0.5;
        C3 w = c3(0.0, 0.0, 0.0);
        if (s.diffuse) w = w + s.mdiff * dot(s.normal, dir) / fac;
        if (s.specular) {
            // upstream evaluates normalize(dir - ray.direction) with `ray` shadowed by the new ray, i.e. the
            // zero vector: 0/0 = NaN (src/raytrace.rs:108,115).  Kept: it defines the output.
            V3 h = normalize(dir - dir);
            w = w + s.mspec * pow(clamp_zero(dot(s.normal, h)), s.exponent) / fac;
        }
        out.o = s.pt + dir * 0.00001;
        out.d = dir;
        out.T = s.T * w;
        out.sig = s.sig;
        out.key = child_key(s.key, CODE_INDIRECT0 + k);
        return true;
    }
    if (k == 0) {
        if (!s.specular) return false;
        V3 rd;
        if (s.kind == MAT_TRANSPARENT) rd = s.d - s.normal * (2.0 * s.ndv);       // src/raytrace.rs:210
        else rd = s.d - s.normal * (2.0 * dot(s.d, s.normal));                    // src/raytrace.rs:61,161
        out.o = s.pt + rd * 0.00001;
        out.d = rd;
        if (s.kind == MAT_PHONG) {
            out.T = s.T * s.mspec;
            out.sig = s.sig * s.ssig;
        } else {
            out.T = s.T * s.mspec * s.fresnel;
            out.sig = s.fresnel * s.sig * s.ssig;
        }
        out.key = child_key(s.key, CODE_REFLECT);
        return true;
    }
    if (k == 1 && s.kind == MAT_TRANSPARENT && s.has_refr) {  // src/raytrace.rs:214-223
        out.o = s.pt + s.refr * 0.00001;
        out.d = s.refr;
        out.T = s.T * s.omf;
        out.sig = s.omf * s.sig;
        out.key = child_key(s.key, CODE_REFRACT);
        return true;
    }
    return false;
}

// ---------------------------------------------------------------------------------------------------------
// image-space shard mapping: tile-ordered pixel slot -> (x, global y, local pixel index)
// A sample plane is npix_padded slots; slot q covers tile q/32 (8 wide x 4 high), lane q%32.
// ---------------------------------------------------------------------------------------------------------
RT_HD bool slot_to_pixel(const FrameParams& fp, uint64_t q, uint32_t& x, uint32_t& y, uint32_t& local_pixel) {
    uint32_t tile = (uint32_t)(q >> 5), lane = (uint32_t)(q & 31);
    uint32_t tx = tile % fp.tiles_x, ty = tile / fp.tiles_x;
    x = tx * 8 + (lane & 7);
    uint32_t yl = ty * 4 + (lane >> 3);
    uint32_t band = yl / fp.band_rows;
    y = (band * fp.world + fp.rank) * fp.band_rows + yl % fp.band_rows;
    local_pixel = yl * fp.width + x;
    return x < fp.width && yl < fp.local_rows && y < fp.height;
}

// final-image word w (4 bytes) -> index of the same word in the rank-major all-gathered buffer
RT_HD uint64_t untile_source_word(uint64_t w, uint32_t words_per_row, uint32_t band_rows, uint32_t world, uint32_t local_rows) {
    const uint32_t y = (uint32_t)(w / words_per_row);
    const uint32_t c = (uint32_t)(w - (uint64_t)y * words_per_row);
    const uint32_t band = y / band_rows;
    const uint32_t rank = band % world;
    const uint32_t yl = (band / world) * band_rows + y % band_rows;
    return ((uint64_t)rank * local_rows + yl) * words_per_row + c;
}

// ---------------------------------------------------------------------------------------------------------
// LBVH construction helpers (Morton codes, Karras 2012 hierarchy); the kernels in bvh_build.cu call these
// ---------------------------------------------------------------------------------------------------------
RT_HD uint64_t spread21(uint64_t v) {
    v &= 0x1FFFFFULL;
    v = (v | (v << 32)) & 0x001F00000000FFFFULL;
    v = (v | (v << 16)) & 0x001F0000FF0000FFULL;
    v = (v | (v << 8)) & 0x100F00F00F00F00FULL;
    v = (v | (v << 4)) & 0x10C30C30C30C30C3ULL;
    v = (v | (v << 2)) & 0x1249249249249249ULL;
    return v;
}
// Scale of the Morton grid: the SAME step on all three axes (1 / longest centroid extent), so that the tree splits the
// long axes of a flat scene first and its boxes stay near-cubical; per-axis scaling (RT_TUNE_MORTON_AXIS, kept for
// A/B runs) splits a 120 x 7 x 120 field as often in y as in x and z and yields pancake boxes.
inline void morton_scale(const double* lo, const double* hi, double* inv_ext, bool per_axis = false) {
    double longest = 0.0;
    for (int a = 0; a < 3; a++) {
        const double ext = hi[a] - lo[a];
        inv_ext[a] = ext > 0.0 ? 1.0 / ext : 0.0;
        if (ext > longest) longest = ext;   // NaN / inf extents never win; their axis keeps inv = 0 or 1/inf = 0
    }
    if (per_axis) return;
    if (longest > 0.0 && longest < (double)INFINITY)
        for (int a = 0; a < 3; a++)
            if (inv_ext[a] > 0.0) inv_ext[a] = 1.0 / longest;
}
RT_HD uint64_t morton63(double cx, double cy, double cz, const double* lo, const double* inv_ext) {
    double fx = (cx - lo[0]) * inv_ext[0], fy = (cy - lo[1]) * inv_ext[1], fz = (cz - lo[2]) * inv_ext[2];
    fx = fmin(fmax(fx * 2097152.0, 0.0), 2097151.0);
    fy = fmin(fmax(fy * 2097152.0, 0.0), 2097151.0);
    fz = fmin(fmax(fz * 2097152.0, 0.0), 2097151.0);
    if (!(fx == fx)) fx = 0.0;
    if (!(fy == fy)) fy = 0.0;
    if (!(fz == fz)) fz = 0.0;
    return (spread21((uint64_t)fx) << 2) | (spread21((uint64_t)fy) << 1) | spread21((uint64_t)fz);
}

RT_HD int clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __clzll((long long)x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}
// common-prefix length of sorted keys i and j (position as tie-break); -1 outside [0, n)
RT_HD int lbvh_delta(const uint64_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = RT_LDG(keys + i), b = RT_LDG(keys + j);
    if (a != b) return clz64(a ^ b);
    return 64 + clz64((uint64_t)(uint32_t)i ^ (uint64_t)(uint32_t)j);
}
// internal node i of a tree over n >= 2 sorted leaves: children (internal index, or ~leaf)
RT_HD void lbvh_node(const uint64_t* keys, int n, int i, int32_t& left, int32_t& right) {
    int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = lbvh_delta(keys, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = lbvh_delta(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + (d < 0 ? -1 : 0);
    int lo = i < j ? i : j, hi = i < j ? j : i;
    left = (lo == gamma) ? ~gamma : gamma;
    right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
}

}  // namespace RT_NS
}  // namespace rt
