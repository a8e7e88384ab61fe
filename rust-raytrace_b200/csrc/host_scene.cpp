// host_scene.cpp -- scene builder API, camera constructors, BMP sink, texture file loaders.
// Host-only; compiled with -ffp-contract=off so the camera matrix is bit-identical to the reference's
// unfused arithmetic (src/camera.rs:51-63).
#include "host_scene.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <thread>
#include <vector>
#include <algorithm>

#include "../../include/raytrace_b200.h"
#include "../../include/rt_srgb_tables.inc"

namespace rt {

const double SRGB_VALUES[256] = RT_SRGB_VALUES_INIT;
const double SRGB_AVERAGE[255] = RT_SRGB_AVERAGE_INIT;

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
const char* get_error() { return g_error.c_str(); }

namespace {
struct D3 {
    double x, y, z;
};
inline D3 d3(const double* p) { return {p[0], p[1], p[2]}; }
inline double dot3(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline D3 cross3(D3 a, D3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline D3 normalize3(D3 a) {
    double n = std::sqrt(dot3(a, a));
    return {a.x / n, a.y / n, a.z / n};
}
}  // namespace

// src/camera.rs:51-63
void camera_spc_new(CameraRec* cam, const double* position, const double* look_, const double* up_, double im_dist) {
    D3 look = d3(look_), up = d3(up_);
    D3 u = normalize3(cross3(look, up));
    D3 v = normalize3(cross3(u, look));
    D3 wn = normalize3(look);
    D3 w = {wn.x * im_dist, wn.y * im_dist, wn.z * im_dist};
    cam->kind = CAM_SPC;
    cam->samples = 1;
    cam->pos[0] = position[0]; cam->pos[1] = position[1]; cam->pos[2] = position[2];
    const double m[9] = {u.x, v.x, w.x, u.y, v.y, w.y, u.z, v.z, w.z};
    std::memcpy(cam->m, m, sizeof m);
    cam->focus = 0.0; cam->aperture = 0.0; cam->im_dist = 0.0;
}

// src/camera.rs:67-73
void camera_spc_look_at(CameraRec* cam, const double* focus, const double* look, const double* up, double pov, double h) {
    double cot = 1.0 / std::tan(pov / 2.0);
    double im_dist = cot;
    double d = h * cot;
    D3 ln = normalize3(d3(look));
    double position[3] = {focus[0] - ln.x * d, focus[1] - ln.y * d, focus[2] - ln.z * d};
    camera_spc_new(cam, position, look, up, im_dist);
}

// src/camera.rs:97-106: im_dist = |M * (0,0,1)|
void camera_make_dof(CameraRec* cam, double focus, double aperture, uint32_t samples) {
    const double* m = cam->m;
    D3 c = {m[0] * 0.0 + m[1] * 0.0 + m[2] * 1.0, m[3] * 0.0 + m[4] * 0.0 + m[5] * 1.0, m[6] * 0.0 + m[7] * 0.0 + m[8] * 1.0};
    cam->im_dist = std::sqrt(dot3(c, c));
    cam->kind = CAM_DOF;
    cam->focus = focus;
    cam->aperture = aperture;
    cam->samples = samples;
}

// ---------------------------------------------------------------------------------------------------------
// texture files.  Upstream decodes through the `image` crate (src/texture.rs:34-37), which is not available;
// 24-bit uncompressed BMP and binary PPM (P6, maxval 255) are decoded here.
// ---------------------------------------------------------------------------------------------------------
static bool read_file(const std::string& path, std::vector<uint8_t>* out, std::string* err) {
    std::ifstream f(path, std::ios::binary);
    if (!f) { *err = "cannot open file"; return false; }
    out->assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
    return true;
}
static uint32_t le32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
static uint16_t le16(const uint8_t* p) { return (uint16_t)(p[0] | (p[1] << 8)); }

bool load_texture_file(const std::string& path, HostTexture* out, std::string* err) {
    std::vector<uint8_t> b;
    if (!read_file(path, &b, err)) return false;
    if (b.size() >= 54 && b[0] == 'B' && b[1] == 'M') {
        uint32_t off = le32(&b[10]), dib = le32(&b[14]);
        if (dib < 40) { *err = "unsupported BMP header"; return false; }
        int32_t w = (int32_t)le32(&b[18]), h = (int32_t)le32(&b[22]);
        uint16_t bpp = le16(&b[28]);
        uint32_t comp = le32(&b[30]);
        if (bpp != 24 || comp != 0 || w <= 0 || h == 0) { *err = "unsupported BMP (need 24-bit uncompressed)"; return false; }
        bool bottom_up = h > 0;
        uint32_t hh = (uint32_t)(h > 0 ? h : -h);
        size_t stride = ((size_t)3 * w + 3) & ~(size_t)3;
        if (b.size() < off + stride * hh) { *err = "truncated BMP"; return false; }
        out->width = (uint32_t)w; out->height = hh;
        out->data.resize((size_t)3 * w * hh);
        for (uint32_t y = 0; y < hh; y++) {
            const uint8_t* row = &b[off + stride * (bottom_up ? hh - 1 - y : y)];
            uint8_t* dst = &out->data[(size_t)3 * w * y];
            for (int32_t x = 0; x < w; x++) { dst[3 * x] = row[3 * x + 2]; dst[3 * x + 1] = row[3 * x + 1]; dst[3 * x + 2] = row[3 * x]; }
        }
        return true;
    }
    if (b.size() >= 2 && b[0] == 'P' && b[1] == '6') {
        size_t i = 2;
        long vals[3];
        for (int k = 0; k < 3; k++) {
            while (i < b.size()) {
                if (b[i] == '#') { while (i < b.size() && b[i] != '\n') i++; }
                else if (b[i] == ' ' || b[i] == '\n' || b[i] == '\r' || b[i] == '\t') i++;
                else break;
            }
            long v = 0; bool any = false;
            while (i < b.size() && b[i] >= '0' && b[i] <= '9') { v = v * 10 + (b[i] - '0'); i++; any = true; }
            if (!any) { *err = "bad PPM header"; return false; }
            vals[k] = v;
        }
        i++;  // single whitespace after maxval
        if (vals[0] <= 0 || vals[1] <= 0 || vals[2] != 255) { *err = "unsupported PPM (need maxval 255)"; return false; }
        size_t need = (size_t)3 * vals[0] * vals[1];
        if (b.size() < i + need) { *err = "truncated PPM"; return false; }
        out->width = (uint32_t)vals[0]; out->height = (uint32_t)vals[1];
        out->data.assign(b.begin() + i, b.begin() + i + need);
        return true;
    }
    *err = "unsupported image format (24-bit BMP and binary PPM are decoded)";
    return false;
}

static void choose_quantisation(FlatScene* f, double rmean, uint32_t tuning);

// Everything of the flat form that depends on where the spheres are and how big: their scan-order records, the bounds,
// the padding scale terms, the grid of the 32-byte nodes.  (The structure -- which object is which shape, scan order,
// planes, materials -- is flatten_scene's.)  rt_scene_update_spheres calls it on the cached form in place.
struct SphereBounds {
    double centroid_lo[3], centroid_hi[3], box_lo[3], box_hi[3];
    double cmax = 0.0, rmin = INFINITY, rsum = 0.0;
    bool finite = true;
    SphereBounds() { for (int a = 0; a < 3; a++) { centroid_lo[a] = INFINITY; centroid_hi[a] = -INFINITY; box_lo[a] = INFINITY; box_hi[a] = -INFINITY; } }
    void merge(const SphereBounds& o) {
        for (int a = 0; a < 3; a++) {
            centroid_lo[a] = std::fmin(centroid_lo[a], o.centroid_lo[a]); centroid_hi[a] = std::fmax(centroid_hi[a], o.centroid_hi[a]);
            box_lo[a] = std::fmin(box_lo[a], o.box_lo[a]); box_hi[a] = std::fmax(box_hi[a], o.box_hi[a]);
        }
        cmax = std::fmax(cmax, o.cmax); rmin = std::fmin(rmin, o.rmin); rsum += o.rsum; finite = finite && o.finite;
    }
};

// runs body(first, last, chunk index) over [0, n) on up to 16 host threads (big scenes: a million objects per frame)
template <typename F>
static void parallel_chunks(size_t n, size_t max_chunks, F body) {
    size_t chunks = n >= (1u << 16) ? std::min<size_t>(std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), 16), max_chunks) : 1;
    if (chunks <= 1) { body((size_t)0, n, (size_t)0); return; }
    std::vector<std::thread> th;
    const size_t per = (n + chunks - 1) / chunks;
    for (size_t c = 0; c < chunks; c++) {
        const size_t lo = c * per, hi = std::min(n, lo + per);
        if (lo >= hi) break;
        th.emplace_back([=] { body(lo, hi, c); });
    }
    for (std::thread& t : th) t.join();
}

// Everything of the flat form that depends on where the spheres are and how big: their scan-order records, the bounds,
// the padding scale terms, the grid of the 32-byte nodes.  (The structure -- which object is which shape, scan order,
// planes, materials -- is flatten_scene's.)  rt_scene_update_spheres calls it on the cached form in place.
static void refresh_sphere_data(const rt_scene* s, FlatScene* f, uint32_t tuning) {
    const size_t n = f->sph.size();
    SphereBounds part[16];
    parallel_chunks(n, 16, [&](size_t first, size_t last, size_t c) {
        SphereBounds b;
        for (size_t k = first; k < last; k++) {
            const HostObject& o = s->objects[(size_t)f->sph_obj[k]];
            SphereRec& r = f->sph[k];
            r.cx = o.p[0]; r.cy = o.p[1]; r.cz = o.p[2];
            r.r2 = o.radius * o.radius;  // src/shapes.rs:63
            const double ar = std::fabs(o.radius);
            b.rsum += ar;
            for (int a = 0; a < 3; a++) {
                if (o.p[a] < b.centroid_lo[a]) b.centroid_lo[a] = o.p[a];
                if (o.p[a] > b.centroid_hi[a]) b.centroid_hi[a] = o.p[a];
                const double m = std::fabs(o.p[a]) + ar;
                if (m > b.cmax) b.cmax = m;
                if (o.p[a] - ar < b.box_lo[a]) b.box_lo[a] = o.p[a] - ar;
                if (o.p[a] + ar > b.box_hi[a]) b.box_hi[a] = o.p[a] + ar;
                if (!std::isfinite(m)) b.finite = false;
            }
            if (ar < b.rmin) b.rmin = ar;
        }
        part[c] = b;
    });
    SphereBounds all;
    for (int c = 0; c < 16; c++) all.merge(part[c]);
    for (int a = 0; a < 3; a++) { f->centroid_lo[a] = all.centroid_lo[a]; f->centroid_hi[a] = all.centroid_hi[a]; f->box_lo[a] = all.box_lo[a]; f->box_hi[a] = all.box_hi[a]; }
    if (!all.finite || f->sph.empty())  // no usable bounds: rays are never advanced (t0 = 0)
        for (int a = 0; a < 3; a++) { f->box_lo[a] = -INFINITY; f->box_hi[a] = INFINITY; }
    f->cmax = all.cmax;
    f->rmin = (all.rmin > 0.0 && std::isfinite(all.rmin)) ? all.rmin : 1e-300;
    choose_quantisation(f, n ? all.rsum / (double)n : 0.0, tuning);
}

int flatten_scene(const rt_scene* s, FlatScene* f, uint32_t tuning) {
    if (!s->camera_set) { set_error("scene has no camera"); return RT_ERR_INVALID; }
    if (!s->background_set) { set_error("scene has no background"); return RT_ERR_INVALID; }
    if (s->bg_kind == BG_SKYBOX)
        for (int k = 0; k < 6; k++)
            if (s->faces[k].data.empty()) { set_error("skybox face missing"); return RT_ERR_INVALID; }
    f->sph.clear(); f->sph_obj.clear(); f->planes.clear();
    f->obj_shape.assign(s->objects.size(), 0u);
    f->obj_material.assign(s->objects.size(), 0u);
    f->sph_obj.reserve(s->n_spheres);
    for (size_t i = 0; i < s->objects.size(); i++) {
        const HostObject& o = s->objects[i];
        f->obj_material[i] = o.material;
        if (o.is_plane) {
            PlaneRec p{};
            std::memcpy(p.p, o.p, sizeof p.p);
            std::memcpy(p.n, o.n, sizeof p.n);
            p.obj = (int32_t)i;
            f->obj_shape[i] = 0x80000000u | (uint32_t)f->planes.size();
            f->planes.push_back(p);
        } else {
            f->obj_shape[i] = (uint32_t)f->sph_obj.size();
            f->sph_obj.push_back((int32_t)i);
        }
    }
    f->sph.assign(f->sph_obj.size(), SphereRec{});
    refresh_sphere_data(s, f, tuning);
    int mc = 0;
    for (const MaterialRec& m : s->materials) {
        int c = m.kind == MAT_INDIRECT ? (int)m.samples : (m.kind == MAT_TRANSPARENT ? 2 : 1);
        if (c > mc) mc = c;
    }
    f->max_children = mc;
    return RT_OK;
}

void (*g_scene_free_hook)(const rt_scene*) = nullptr;

void scene_touch(rt_scene* s) {
    {
        std::lock_guard<std::mutex> g(s->flat_cache.m);
        s->flat_cache.p.reset();
        s->revision++;
    }
    std::lock_guard<std::mutex> g(s->device_cache.m);
    s->device_cache.entries.clear();
}

// rt_scene_update_spheres: the objects' positions / radii have just changed and nothing else.  The cached flat form (if
// it is current) is brought up to date IN PLACE -- its vectors keep their addresses, so the page-lock on them stays -- and
// moves to the new revision; the device copies are dropped as for any other edit.
void scene_spheres_moved(rt_scene* s) {
    {
        std::lock_guard<std::mutex> g(s->flat_cache.m);
        std::shared_ptr<FlatCache>& c = s->flat_cache.p;
        const bool current = c && c->revision == s->revision && c->flat.obj_shape.size() == s->objects.size();
        s->revision++;
        if (current) {
            refresh_sphere_data(s, &c->flat, c->node_tuning);
            c->revision = s->revision;
        } else {
            c.reset();
        }
    }
    std::lock_guard<std::mutex> g(s->device_cache.m);
    s->device_cache.entries.clear();
}

int flatten_scene_cached(const rt_scene* s, uint32_t tuning, std::shared_ptr<FlatCache>* out) {
    const uint32_t node_tuning = tuning & (RT_TUNE_NODES_F32 | RT_TUNE_NODES_Q16);
    std::lock_guard<std::mutex> g(s->flat_cache.m);
    std::shared_ptr<FlatCache>& c = s->flat_cache.p;
    if (c && c->revision == s->revision && c->node_tuning == node_tuning && c->flat.obj_shape.size() == s->objects.size()) {
        // what flatten_scene checks without looking at the objects
        if (!s->camera_set) { set_error("scene has no camera"); return RT_ERR_INVALID; }
        if (!s->background_set) { set_error("scene has no background"); return RT_ERR_INVALID; }
        if (s->bg_kind == BG_SKYBOX)
            for (int k = 0; k < 6; k++)
                if (s->faces[k].data.empty()) { set_error("skybox face missing"); return RT_ERR_INVALID; }
        *out = c;
        return RT_OK;
    }
    std::shared_ptr<FlatCache> fresh = std::make_shared<FlatCache>();
    int rc = flatten_scene(s, &fresh->flat, node_tuning);
    if (rc != RT_OK) return rc;
    fresh->revision = s->revision;
    fresh->node_tuning = node_tuning;
    c = fresh;
    *out = c;
    return RT_OK;
}

// The 32-byte nodes store box planes on a 65535-step grid over the sphere bounds.  A box grows by up to one cell from
// the rounding and the filter pads it by another 1.25 cells (rt_device.cuh make_boxray32), so the grid is used only
// when 2.25 cells stay under a quarter of the mean radius on every axis and the bounds are small enough for the f32
// filter terms (2^23 * cell / |d|, |d| clamped at 1e-30) not to overflow.  RT_TUNE_NODES_F32 / RT_TUNE_NODES_Q16 override the
// size test (A-B measurements, parity tests); unusable bounds always mean f32 nodes.
static void choose_quantisation(FlatScene* f, double rmean, uint32_t tuning) {
    f->quant = false;
    for (int a = 0; a < 3; a++) { f->q_org[a] = 0.0; f->q_cell[a] = 0.0; }
    if (f->sph.empty()) return;
    double cell_max = 0.0, ext_max = 0.0;
    for (int a = 0; a < 3; a++) {
        const double lo = f->box_lo[a], hi = f->box_hi[a];
        if (!std::isfinite(lo) || !std::isfinite(hi) || !(hi >= lo)) return;
        // the node boxes are the f64 bounds rounded outward to f32: leave room for that
        const double m = (std::fabs(lo) + std::fabs(hi)) * 0x1.0p-21 + 1e-30;
        f->q_org[a] = lo - m;
        f->q_cell[a] = (hi + m - f->q_org[a]) / 65534.0;
        if (!(f->q_cell[a] > 0.0) || !std::isfinite(f->q_cell[a])) return;
        if (f->q_cell[a] > cell_max) cell_max = f->q_cell[a];
        if (hi - lo > ext_max) ext_max = hi - lo;
    }
    if (!std::isfinite(rmean) || !(ext_max <= 1e6) || !(f->cmax <= 1e6)) return;
    bool fine = 2.25 * cell_max <= 0.25 * rmean;
    if (tuning & RT_TUNE_NODES_F32) fine = false;
    else if (tuning & RT_TUNE_NODES_Q16) fine = true;
    f->quant = fine;
}

void fill_scene_header(const rt_scene* s, const FlatScene& f, DeviceScene* ds) {
    ds->n_objects = (uint32_t)s->objects.size();
    ds->n_spheres = (uint32_t)f.sph.size();
    ds->n_planes = (uint32_t)f.planes.size();
    ds->n_lights = (uint32_t)s->lights.size();
    ds->camera = s->camera;
    ds->bg_kind = s->bg_kind;
    std::memcpy(ds->bg_color, s->bg_color, sizeof ds->bg_color);
    for (int k = 0; k < 6; k++) {
        ds->faces[k].width = s->faces[k].width;
        ds->faces[k].height = s->faces[k].height;
        ds->faces[k].data = nullptr;
    }
    ds->max_children = f.max_children;
    ds->bvh_cmax = f.cmax;
    ds->bvh_rmin = f.rmin;
    ds->bvh_inv2rmin = 0.5 / f.rmin * (1.0 + 0x1.0p-50);
    for (int a = 0; a < 3; a++) {
        ds->bvh_lo[a] = f.box_lo[a]; ds->bvh_hi[a] = f.box_hi[a];
        ds->bvh_lo_f[a] = std::nextafterf((float)f.box_lo[a], -INFINITY);   // outward: at least one f32 step beyond the f64 bound
        ds->bvh_hi_f[a] = std::nextafterf((float)f.box_hi[a], INFINITY);
    }
    ds->bvh_qnodes = nullptr;  // set by the owner of the device buffers when f.quant
    ds->bvh_wnodes = nullptr;
    ds->_pad_bvh = 0;
    for (int a = 0; a < 3; a++) { ds->q_org[a] = f.q_org[a]; ds->q_cell[a] = f.q_cell[a]; }
}

bool accel_uses_bvh(int accel, uint32_t n_spheres) {
    if (n_spheres == 0) return false;
    return accel == RT_ACCEL_BVH || (accel == RT_ACCEL_AUTO && n_spheres >= 64);
}

int make_frame_params(const rt_scene* s, const rt_render_opts* o, FrameParams* fp) {
    if (o->world < 1 || o->rank < 0 || o->rank >= o->world) { set_error("bad rank/world"); return RT_ERR_INVALID; }
    const uint32_t band = o->band_rows ? o->band_rows : 4;
    if (band % 4) { set_error("band_rows must be a multiple of 4"); return RT_ERR_INVALID; }
    fp->width = o->width ? o->width : s->width;
    fp->height = o->height ? o->height : s->height;
    fp->antialias = o->antialias != 0xFFFFFFFFu ? o->antialias : s->antialias;
    fp->cam_samples = s->camera.samples;
    if (fp->width == 0 || fp->height == 0) { set_error("image size is zero (set Options or opts.width/height)"); return RT_ERR_INVALID; }
    if (o->max_depth > 60) { set_error("max_depth > 60 unsupported"); return RT_ERR_INVALID; }
    fp->rank = (uint32_t)o->rank; fp->world = (uint32_t)o->world; fp->band_rows = band;
    const uint32_t n_bands = (fp->height + band - 1) / band;
    fp->local_rows = ((n_bands + fp->world - 1) / fp->world) * band;
    fp->tiles_x = (fp->width + 7) / 8;
    fp->max_depth = o->max_depth;
    fp->min_significance = o->min_significance;
    fp->seed = o->seed;
    fp->npix_padded = (uint64_t)fp->tiles_x * 8 * fp->local_rows;
    return RT_OK;
}

}  // namespace rt

// =========================================================================================================
// C ABI: scene builder
// =========================================================================================================
using namespace rt;

extern "C" {

const char* rt_last_error(void) { return get_error(); }
const char* rt_version(void) { return "raytrace_b200 0.1 (sm_100a)"; }

rt_scene* rt_scene_new(void) { return new rt_scene(); }
void rt_scene_free(rt_scene* s) {
    if (s && rt::g_scene_free_hook) rt::g_scene_free_hook(s);  // context.cu: nothing cached may keep pointing at the scene
    delete s;
}

static int add_material(rt_scene* s, int32_t kind, const double* diffuse, const double* specular, double exponent, const double* ambient, double ior,
                        uint32_t samples) {
    if (!s) { set_error("null scene"); return RT_ERR_INVALID; }
    MaterialRec m{};
    m.kind = kind;
    m.samples = samples;
    for (int i = 0; i < 3; i++) {
        m.diffuse[i] = diffuse ? diffuse[i] : 0.0;
        m.specular[i] = specular ? specular[i] : 0.0;
        m.ambient[i] = ambient ? ambient[i] : 0.0;
    }
    m.exponent = exponent;
    m.ior = ior;
    scene_touch(s);
    s->materials.push_back(m);
    return (int)s->materials.size() - 1;
}
int rt_scene_add_material_phong(rt_scene* s, const double* d, const double* sp, double e, const double* a) {
    return add_material(s, MAT_PHONG, d, sp, e, a, 1.0, 0);
}
int rt_scene_add_material_indirect(rt_scene* s, const double* d, const double* sp, double e, const double* a, uint32_t samples) {
    return add_material(s, MAT_INDIRECT, d, sp, e, a, 1.0, samples);
}
int rt_scene_add_material_fresnel(rt_scene* s, const double* d, const double* sp, double e, const double* a, double ior) {
    return add_material(s, MAT_FRESNEL, d, sp, e, a, ior, 0);
}
int rt_scene_add_material_transparent(rt_scene* s, const double* sp, double e, double ior) {
    return add_material(s, MAT_TRANSPARENT, nullptr, sp, e, nullptr, ior, 0);
}

int rt_scene_add_spheres(rt_scene* s, uint64_t n, const double* cr, const uint32_t* material) {
    if (!s || (n && (!cr || !material))) { set_error("rt_scene_add_spheres: null argument"); return RT_ERR_INVALID; }
    for (uint64_t i = 0; i < n; i++)
        if (material[i] >= s->materials.size()) { set_error("rt_scene_add_spheres: material index out of range"); return RT_ERR_INVALID; }
    if (s->objects.size() + n > 0x7FFFFFF0ull) { set_error("too many objects"); return RT_ERR_INVALID; }
    scene_touch(s);
    s->objects.reserve(s->objects.size() + n);
    for (uint64_t i = 0; i < n; i++) {
        HostObject o{};
        o.is_plane = false;
        o.p[0] = cr[4 * i]; o.p[1] = cr[4 * i + 1]; o.p[2] = cr[4 * i + 2];
        o.radius = cr[4 * i + 3];
        o.material = material[i];
        s->objects.push_back(o);
    }
    s->n_spheres += n;
    return RT_OK;
}
int rt_scene_update_spheres(rt_scene* s, uint64_t first_object, uint64_t n, const double* cr) {
    if (!s || (n && !cr)) { set_error("rt_scene_update_spheres: null argument"); return RT_ERR_INVALID; }
    if (first_object > s->objects.size() || n > s->objects.size() - first_object) { set_error("rt_scene_update_spheres: object range out of bounds"); return RT_ERR_INVALID; }
    // the range must hold spheres only: checked through the flat form's structure when it is current (4 bytes per object
    // instead of the 72-byte HostObject), else on the objects
    uint64_t bad = n;
    {
        std::lock_guard<std::mutex> g(s->flat_cache.m);
        const std::shared_ptr<FlatCache>& c = s->flat_cache.p;
        const bool current = c && c->revision == s->revision && c->flat.obj_shape.size() == s->objects.size();
        for (uint64_t i = 0; i < n && bad == n; i++)
            if (current ? (c->flat.obj_shape[first_object + i] & 0x80000000u) != 0u : s->objects[first_object + i].is_plane) bad = i;
    }
    if (bad != n) { set_error("rt_scene_update_spheres: object " + std::to_string(first_object + bad) + " is a plane"); return RT_ERR_INVALID; }
    parallel_chunks((size_t)n, 16, [&](size_t first, size_t last, size_t) {
        for (size_t i = first; i < last; i++) {
            HostObject& o = s->objects[first_object + i];
            o.p[0] = cr[4 * i]; o.p[1] = cr[4 * i + 1]; o.p[2] = cr[4 * i + 2];
            o.radius = cr[4 * i + 3];
        }
    });
    scene_spheres_moved(s);
    return RT_OK;
}
int rt_scene_add_plane(rt_scene* s, const double* point, const double* normal, uint32_t material) {
    if (!s || !point || !normal) { set_error("rt_scene_add_plane: null argument"); return RT_ERR_INVALID; }
    if (material >= s->materials.size()) { set_error("rt_scene_add_plane: material index out of range"); return RT_ERR_INVALID; }
    HostObject o{};
    o.is_plane = true;
    for (int i = 0; i < 3; i++) { o.p[i] = point[i]; o.n[i] = normal[i]; }
    o.material = material;
    scene_touch(s);
    s->objects.push_back(o);
    s->n_planes++;
    return RT_OK;
}

static int add_light(rt_scene* s, int32_t kind, const double* a, const double* s1, const double* s2, const double* color) {
    if (!s || !a || !color) { set_error("add_light: null argument"); return RT_ERR_INVALID; }
    LightRec l{};
    l.kind = kind;
    for (int i = 0; i < 3; i++) { l.a[i] = a[i]; l.s1[i] = s1 ? s1[i] : 0.0; l.s2[i] = s2 ? s2[i] : 0.0; l.color[i] = color[i]; }
    scene_touch(s);  // the device copy kept with the scene (DeviceSceneSlot) holds the light records
    s->lights.push_back(l);
    return RT_OK;
}
int rt_scene_add_light_point(rt_scene* s, const double* loc, const double* color) { return add_light(s, LIGHT_POINT, loc, nullptr, nullptr, color); }
int rt_scene_add_light_directional(rt_scene* s, const double* dir, const double* color) {
    return add_light(s, LIGHT_DIRECTIONAL, dir, nullptr, nullptr, color);
}
int rt_scene_add_light_area(rt_scene* s, const double* origin, const double* s1, const double* s2, const double* color) {
    if (!s1 || !s2) { set_error("add_light_area: null side"); return RT_ERR_INVALID; }
    return add_light(s, LIGHT_AREA, origin, s1, s2, color);
}

int rt_scene_set_camera_spc_new(rt_scene* s, const double* position, const double* look, const double* up, double im_dist) {
    if (!s || !position || !look || !up) { set_error("set_camera: null argument"); return RT_ERR_INVALID; }
    camera_spc_new(&s->camera, position, look, up, im_dist);
    s->camera_set = true;
    return RT_OK;
}
int rt_scene_set_camera_spc_look_at(rt_scene* s, const double* focus, const double* look, const double* up, double pov, double h) {
    if (!s || !focus || !look || !up) { set_error("set_camera: null argument"); return RT_ERR_INVALID; }
    camera_spc_look_at(&s->camera, focus, look, up, pov, h);
    s->camera_set = true;
    return RT_OK;
}
int rt_scene_set_camera_dof(rt_scene* s, double focus, double aperture, uint32_t samples) {
    if (!s || !s->camera_set) { set_error("set_camera_dof: set a perspective camera first"); return RT_ERR_INVALID; }
    camera_make_dof(&s->camera, focus, aperture, samples);
    return RT_OK;
}
int rt_scene_get_camera(const rt_scene* s, double* position, double* m) {
    if (!s || !s->camera_set) { set_error("get_camera: no camera"); return RT_ERR_INVALID; }
    if (position) std::memcpy(position, s->camera.pos, sizeof s->camera.pos);
    if (m) std::memcpy(m, s->camera.m, sizeof s->camera.m);
    return RT_OK;
}

int rt_scene_set_background_solid(rt_scene* s, const double* color) {
    if (!s || !color) { set_error("set_background: null argument"); return RT_ERR_INVALID; }
    s->bg_kind = BG_SOLID;
    std::memcpy(s->bg_color, color, sizeof s->bg_color);
    s->background_set = true;
    return RT_OK;
}
int rt_scene_set_skybox_face(rt_scene* s, int face, uint32_t w, uint32_t h, const uint8_t* rgb) {
    if (!s || !rgb || face < 0 || face > 5 || w == 0 || h == 0) { set_error("set_skybox_face: bad argument"); return RT_ERR_INVALID; }
    s->bg_kind = BG_SKYBOX;
    scene_touch(s);
    s->faces[face].width = w; s->faces[face].height = h;
    s->faces[face].data.assign(rgb, rgb + (size_t)3 * w * h);
    s->background_set = true;
    return RT_OK;
}
int rt_scene_load_skybox_face(rt_scene* s, int face, const char* path) {
    if (!s || !path || face < 0 || face > 5) { set_error("load_skybox_face: bad argument"); return RT_ERR_INVALID; }
    std::string err;
    HostTexture t;
    if (!load_texture_file(path, &t, &err)) { set_error(std::string("error loading \"") + path + "\": " + err); return RT_ERR_IO; }
    s->bg_kind = BG_SKYBOX;
    scene_touch(s);
    s->faces[face] = std::move(t);
    s->background_set = true;
    return RT_OK;
}

int rt_scene_set_options(rt_scene* s, uint32_t w, uint32_t h, uint32_t aa) {
    if (!s) { set_error("null scene"); return RT_ERR_INVALID; }
    s->width = w; s->height = h; s->antialias = aa;
    s->options_set = true;
    return RT_OK;
}
int rt_scene_get_options(const rt_scene* s, uint32_t* w, uint32_t* h, uint32_t* aa) {
    if (!s) { set_error("null scene"); return RT_ERR_INVALID; }
    if (w) *w = s->width;
    if (h) *h = s->height;
    if (aa) *aa = s->antialias;
    return RT_OK;
}
const char* rt_scene_warnings(const rt_scene* s) { return s ? s->warnings.c_str() : ""; }

int rt_scene_get_info(const rt_scene* s, rt_scene_info* info) {
    if (!s || !info) { set_error("null argument"); return RT_ERR_INVALID; }
    info->n_objects = s->objects.size();
    info->n_spheres = s->n_spheres;
    info->n_planes = s->n_planes;
    info->n_materials = s->materials.size();
    info->n_lights = s->lights.size();
    info->camera_kind = s->camera.kind;
    info->background_kind = s->bg_kind;
    return RT_OK;
}

static void put3(std::ostringstream& o, const char* name, const double* v) {
    char b[128];
    std::snprintf(b, sizeof b, " %s=(%a,%a,%a)", name, v[0], v[1], v[2]);
    o << b;
}
static void put1(std::ostringstream& o, const char* name, double v) {
    char b[64];
    std::snprintf(b, sizeof b, " %s=%a", name, v);
    o << b;
}

size_t rt_scene_dump(const rt_scene* s, char* buf, size_t cap) {
    std::ostringstream o;
    static const char* MK[] = {"phong", "indirect", "fresnel", "transparent"};
    static const char* LK[] = {"point", "directional", "area"};
    for (size_t i = 0; i < s->objects.size(); i++) {
        const HostObject& ob = s->objects[i];
        const MaterialRec& m = s->materials[ob.material];
        o << "object " << i;
        if (ob.is_plane) { o << " plane"; put3(o, "point", ob.p); put3(o, "normal", ob.n); }
        else { o << " sphere"; put3(o, "center", ob.p); put1(o, "radius", ob.radius); }
        o << " material=" << MK[m.kind];
        if (m.kind != MAT_TRANSPARENT) put3(o, "diffuse", m.diffuse);
        put3(o, "specular", m.specular);
        put1(o, "exponent", m.exponent);
        if (m.kind != MAT_TRANSPARENT) put3(o, "ambient", m.ambient);
        if (m.kind == MAT_INDIRECT) o << " samples=" << m.samples;
        if (m.kind == MAT_FRESNEL || m.kind == MAT_TRANSPARENT) put1(o, "ior", m.ior);
        o << "\n";
    }
    for (size_t i = 0; i < s->lights.size(); i++) {
        const LightRec& l = s->lights[i];
        o << "light " << i << " " << LK[l.kind];
        put3(o, l.kind == LIGHT_POINT ? "location" : (l.kind == LIGHT_DIRECTIONAL ? "direction" : "origin"), l.a);
        if (l.kind == LIGHT_AREA) { put3(o, "side1", l.s1); put3(o, "side2", l.s2); }
        put3(o, "color", l.color);
        o << "\n";
    }
    o << "camera " << (s->camera.kind == CAM_SPC ? "spc" : "dof");
    put3(o, "position", s->camera.pos);
    put3(o, "row0", s->camera.m); put3(o, "row1", s->camera.m + 3); put3(o, "row2", s->camera.m + 6);
    if (s->camera.kind == CAM_DOF) {
        put1(o, "focus", s->camera.focus); put1(o, "aperture", s->camera.aperture); put1(o, "im_dist", s->camera.im_dist);
        o << " samples=" << s->camera.samples;
    }
    o << "\n";
    if (s->bg_kind == BG_SOLID) { o << "background solid"; put3(o, "color", s->bg_color); o << "\n"; }
    else {
        o << "background skybox";
        for (int f = 0; f < 6; f++) {
            uint64_t h = 1469598103934665603ull;
            for (uint8_t b : s->faces[f].data) { h ^= b; h *= 1099511628211ull; }
            o << " face" << f << "=" << s->faces[f].width << "x" << s->faces[f].height << ":" << std::hex << h << std::dec;
        }
        o << "\n";
    }
    o << "options width=" << s->width << " height=" << s->height << " antialias=" << s->antialias << "\n";
    std::string str = o.str();
    if (buf && cap) {
        size_t n = str.size() < cap - 1 ? str.size() : cap - 1;
        std::memcpy(buf, str.data(), n);
        buf[n] = 0;
    }
    return str.size();
}

// ---------------------------------------------------------------------------------------------------------
// BMP sink (src/bmp.rs:10-61)
// ---------------------------------------------------------------------------------------------------------
uint32_t rt_bmp_bytewidth(uint32_t width) { return (3u * width + 3u) & 0xFFFFFFFCu; }

void rt_bmp_header(uint32_t width, uint32_t height, uint8_t* h) {
    const uint32_t bytewidth = rt_bmp_bytewidth(width);
    const uint32_t pasize = bytewidth * height;
    const uint32_t fsize = 14u + 108u + pasize;
    std::memset(h, 0, 122);
    auto put32 = [h](int off, uint32_t v) { h[off] = (uint8_t)v; h[off + 1] = (uint8_t)(v >> 8); h[off + 2] = (uint8_t)(v >> 16); h[off + 3] = (uint8_t)(v >> 24); };
    h[0] = 'B'; h[1] = 'M';
    put32(2, fsize);       // file size
    put32(10, 122);        // offset of the pixel array
    put32(14, 108);        // BITMAPV4HEADER
    put32(18, width);
    put32(22, height);     // positive: bottom-up rows
    h[26] = 1;             // planes
    h[28] = 24;            // bits per pixel
    put32(34, pasize);
    put32(38, 2835);       // 72 dpi
    put32(42, 2835);
    h[70] = 'B'; h[71] = 'G'; h[72] = 'R'; h[73] = 's';  // LCS_sRGB, little endian
}

int rt_bmp_write(const char* path, uint32_t width, uint32_t height, const uint8_t* rows) {
    if (!path || !rows) { set_error("rt_bmp_write: null argument"); return RT_ERR_INVALID; }
    FILE* f = std::fopen(path, "wb");
    if (!f) { set_error(std::string("error creating ") + path); return RT_ERR_IO; }
    uint8_t hdr[122];
    rt_bmp_header(width, height, hdr);
    size_t n = (size_t)rt_bmp_bytewidth(width) * height;
    bool ok = std::fwrite(hdr, 1, 122, f) == 122 && std::fwrite(rows, 1, n, f) == n;
    ok = (std::fclose(f) == 0) && ok;
    if (!ok) { set_error(std::string("error writing ") + path); return RT_ERR_IO; }
    return RT_OK;
}

void rt_render_opts_default(rt_render_opts* o) {
    if (!o) return;
    std::memset(o, 0, sizeof *o);
    o->antialias = 0xFFFFFFFFu;
    o->max_depth = 4;
    o->min_significance = 1.0 / 256.0 / 2.0;
    o->seed = 1;
    o->accel = RT_ACCEL_AUTO;
    o->fp_mode = RT_FP_STRICT;
    o->world = 1;
    o->band_rows = 4;
}

}  // extern "C"
