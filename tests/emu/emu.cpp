// emu.cpp -- TEST-ONLY serial emulation of the wavefront, compiled with g++ from the product's own headers.
//
// Purpose: exercise rust-raytrace_b200/csrc/rt_device.cuh (shading in path-throughput form, counter RNG keys,
// BVH traversal + Karras hierarchy helpers, shard mapping, quantiser) on a machine without a GPU, against the
// oracle.  It is NOT a CPU fallback: it lives under tests/, is never built into or loaded by
// libraytrace_b200.so, and nothing under rust-raytrace_b200/ refers to it.
//
// What it does not cover (GPU-only, tested by `pytest -m gpu`): the kernels themselves -- TMA staging, warp
// aggregation, atomics, the device BVH build kernels (sort/refit), persistent scheduling.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/raytrace_b200.h"
#include "../../rust-raytrace_b200/csrc/host_scene.h"
#define RT_NS strict
#include "../../rust-raytrace_b200/csrc/rt_device.cuh"

using namespace rt;
using namespace rt::strict;

namespace {

struct EmuScene {
    FlatScene flat;
    DeviceScene ds{};
    std::vector<BvhNode> nodes;
    std::vector<BvhQNode> qnodes;  // grid form, when the scene's grid is fine enough (host_scene.cpp choose_quantisation)
    std::vector<BvhWNode> wnodes;  // wide form for shadow queries, unless RT_TUNE_BINARY_SHADOW (as in context.cu)
    std::vector<SphereRec> bvh_sph;
    std::vector<int32_t> bvh_obj;
    bool use_bvh = false;
};

struct Box {
    float lo[3], hi[3];
};
float round_down(double x) { float f = (float)x; return ((double)f > x) ? nextafterf(f, -INFINITY) : f; }
float round_up(double x) { float f = (float)x; return ((double)f < x) ? nextafterf(f, INFINITY) : f; }
Box leaf_box(const SphereRec& s) {
    const double r = std::sqrt(s.r2) * (1.0 + 0x1.0p-50);
    Box b;
    b.lo[0] = round_down(s.cx - r); b.lo[1] = round_down(s.cy - r); b.lo[2] = round_down(s.cz - r);
    b.hi[0] = round_up(s.cx + r); b.hi[1] = round_up(s.cy + r); b.hi[2] = round_up(s.cz + r);
    return b;
}

// child boxes of every node from its links, bottom-up (iterative post-order from the root)
void fill_boxes(EmuScene& e, const std::vector<int32_t>& c0, const std::vector<int32_t>& c1) {
    const size_t n_nodes = c0.size();
    std::vector<Box> nb(n_nodes);
    std::vector<std::pair<int32_t, int>> st;
    st.push_back({0, 0});
    while (!st.empty()) {
        auto [node, phase] = st.back();
        if (phase == 0) {
            st.back().second = 1;
            if (c0[node] >= 0) st.push_back({c0[node], 0});
            if (c1[node] >= 0) st.push_back({c1[node], 0});
        } else {
            st.pop_back();
            Box b0 = c0[node] < 0 ? leaf_box(e.bvh_sph[~c0[node]]) : nb[c0[node]];
            Box b1 = c1[node] < 0 ? leaf_box(e.bvh_sph[~c1[node]]) : nb[c1[node]];
            BvhNode& nd = e.nodes[node];
            for (int a = 0; a < 3; a++) {
                nd.lo0[a] = b0.lo[a]; nd.hi0[a] = b0.hi[a];
                nd.lo1[a] = b1.lo[a]; nd.hi1[a] = b1.hi[a];
                nb[node].lo[a] = std::fmin(b0.lo[a], b1.lo[a]);
                nb[node].hi[a] = std::fmax(b0.hi[a], b1.hi[a]);
            }
            nd.c0 = c0[node]; nd.c1 = c1[node];
        }
    }
}

// EXPERIMENT (EMU_BVH_BUILDER=sah): full-sweep SAH, top-down, one sphere per leaf -- an upper bound on what a better tree than
// the LBVH could save in node visits.  Same node format, so every walk of rt_device.cuh runs on it unchanged.
void build_bvh_sah(EmuScene& e) {
    const uint32_t n = (uint32_t)e.flat.sph.size();
    std::vector<Box> pb(n);
    for (uint32_t i = 0; i < n; i++) pb[i] = leaf_box(e.flat.sph[i]);
    std::vector<uint32_t> idx(n);
    for (uint32_t i = 0; i < n; i++) idx[i] = i;
    e.nodes.assign(n - 1, BvhNode{});
    std::vector<int32_t> c0(n - 1), c1(n - 1);
    std::vector<uint32_t> leaf_order; leaf_order.reserve(n);
    auto area = [](const Box& b) {
        const double x = (double)b.hi[0] - b.lo[0], y = (double)b.hi[1] - b.lo[1], z = (double)b.hi[2] - b.lo[2];
        return x * y + y * z + z * x;
    };
    auto grow = [](Box& a, const Box& b) { for (int k = 0; k < 3; k++) { a.lo[k] = std::fmin(a.lo[k], b.lo[k]); a.hi[k] = std::fmax(a.hi[k], b.hi[k]); } };
    const Box empty{{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}};
    uint32_t next_node = 0;
    struct Job { uint32_t lo, hi; int32_t* link; };
    std::vector<Job> jobs;
    int32_t root_link = 0;
    jobs.push_back({0, n, &root_link});
    std::vector<double> right_area;
    while (!jobs.empty()) {
        const Job j = jobs.back(); jobs.pop_back();
        const uint32_t cnt = j.hi - j.lo;
        if (cnt == 1) { *j.link = ~(int32_t)leaf_order.size(); leaf_order.push_back(idx[j.lo]); continue; }
        double best = INFINITY; int best_axis = 0; uint32_t best_split = j.lo + cnt / 2;
        for (int a = 0; a < 3; a++) {
            std::sort(idx.begin() + j.lo, idx.begin() + j.hi, [&](uint32_t x, uint32_t y) { return pb[x].lo[a] + pb[x].hi[a] < pb[y].lo[a] + pb[y].hi[a]; });
            right_area.assign(cnt, 0.0);
            Box r = empty;
            for (uint32_t k = cnt; k-- > 1;) { grow(r, pb[idx[j.lo + k]]); right_area[k] = area(r); }
            Box l = empty;
            for (uint32_t k = 1; k < cnt; k++) {
                grow(l, pb[idx[j.lo + k - 1]]);
                const double cost = area(l) * k + right_area[k] * (cnt - k);
                if (cost < best) { best = cost; best_axis = a; best_split = j.lo + k; }
            }
        }
        const int a = best_axis;
        std::sort(idx.begin() + j.lo, idx.begin() + j.hi, [&](uint32_t x, uint32_t y) { return pb[x].lo[a] + pb[x].hi[a] < pb[y].lo[a] + pb[y].hi[a]; });
        const uint32_t node = next_node++;
        *j.link = (int32_t)node;
        jobs.push_back({best_split, j.hi, &c1[node]});   // child 0 is popped first: leaves come out in depth-first order
        jobs.push_back({j.lo, best_split, &c0[node]});
    }
    e.bvh_sph.resize(n); e.bvh_obj.resize(n);
    for (uint32_t i = 0; i < n; i++) { e.bvh_sph[i] = e.flat.sph[leaf_order[i]]; e.bvh_obj[i] = e.flat.sph_obj[leaf_order[i]]; }
    fill_boxes(e, c0, c1);
}

// the device build's algorithm, serially: same Morton codes, same (code, index) order, same Karras nodes
void build_bvh(EmuScene& e, bool morton_per_axis) {
    const uint32_t n = (uint32_t)e.flat.sph.size();
    if (n == 0) return;
    if (n > 1 && getenv("EMU_BVH_BUILDER") && std::string(getenv("EMU_BVH_BUILDER")) == "sah") { build_bvh_sah(e); return; }
    double inv[3];
    morton_scale(e.flat.centroid_lo, e.flat.centroid_hi, inv, morton_per_axis);
    std::vector<std::pair<uint64_t, uint32_t>> keyed(n);
    for (uint32_t i = 0; i < n; i++) keyed[i] = {morton63(e.flat.sph[i].cx, e.flat.sph[i].cy, e.flat.sph[i].cz, e.flat.centroid_lo, inv), i};
    std::sort(keyed.begin(), keyed.end());
    std::vector<uint64_t> keys(n);
    e.bvh_sph.resize(n); e.bvh_obj.resize(n);
    for (uint32_t i = 0; i < n; i++) {
        keys[i] = keyed[i].first;
        e.bvh_sph[i] = e.flat.sph[keyed[i].second];
        e.bvh_obj[i] = e.flat.sph_obj[keyed[i].second];
    }
    if (n == 1) {
        Box b = leaf_box(e.bvh_sph[0]);
        BvhNode nd{};
        for (int a = 0; a < 3; a++) { nd.lo0[a] = b.lo[a]; nd.hi0[a] = b.hi[a]; nd.lo1[a] = b.lo[a]; nd.hi1[a] = b.hi[a]; }
        nd.c0 = ~0; nd.c1 = ~0;
        e.nodes.assign(1, nd);
        return;
    }
    e.nodes.assign(n - 1, BvhNode{});
    std::vector<int32_t> c0(n - 1), c1(n - 1);
    for (uint32_t i = 0; i + 1 < n; i++) lbvh_node(keys.data(), (int)n, (int)i, c0[i], c1[i]);
    fill_boxes(e, c0, c1);
}

int make_scene(const rt_scene* s, int accel, uint32_t tuning, EmuScene* e) {
    const rt_scene* use = s;
    rt_scene patched;
    if (!s->camera_set || !s->background_set) {
        patched = *s;
        patched.camera_set = true; patched.background_set = true; patched.bg_kind = BG_SOLID;
        use = &patched;
    }
    int rc = flatten_scene(use, &e->flat, tuning);
    if (rc != RT_OK) return rc;
    fill_scene_header(use, e->flat, &e->ds);
    e->ds.sph = e->flat.sph.data(); e->ds.sph_obj = e->flat.sph_obj.data(); e->ds.planes = e->flat.planes.data();
    e->ds.obj_shape = e->flat.obj_shape.data(); e->ds.obj_material = e->flat.obj_material.data();
    e->ds.materials = s->materials.data(); e->ds.lights = s->lights.data();
    for (int k = 0; k < 6; k++) e->ds.faces[k].data = s->faces[k].data.data();
    e->ds.srgb_values = SRGB_VALUES; e->ds.srgb_average = SRGB_AVERAGE;
    e->use_bvh = accel_uses_bvh(accel, e->ds.n_spheres);
    if (e->use_bvh) {
        build_bvh(*e, (tuning & RT_TUNE_MORTON_AXIS) != 0);
        e->ds.bvh_nodes = e->nodes.data(); e->ds.bvh_sph = e->bvh_sph.data(); e->ds.bvh_obj = e->bvh_obj.data();
        e->ds.n_bvh_nodes = (uint32_t)e->nodes.size();
        if (e->flat.quant) {
            bool ok = true;
            e->qnodes.resize(e->nodes.size());
            for (size_t i = 0; i < e->nodes.size(); i++) e->qnodes[i] = quantise_node(e->nodes[i], e->flat.q_org, e->flat.q_cell, ok);
            if (ok) e->ds.bvh_qnodes = e->qnodes.data();
            if (ok && !(tuning & RT_TUNE_BINARY_SHADOW)) {  // as context.cu: shadow queries walk the four-wide nodes
                e->wnodes.resize(e->qnodes.size());
                for (size_t i = 0; i < e->qnodes.size(); i++) e->wnodes[i] = widen_node(e->qnodes.data(), (uint32_t)i);
                e->ds.bvh_wnodes = e->wnodes.data();
            }
        }
    }
    return RT_OK;
}

TraverseCount g_tc{0, 0};

struct QRay {
    V3 o, d;
    C3 T;
    double sig;
    uint64_t key;
    uint32_t pixel, meta;
};

}  // namespace

extern "C" {

// same contract as rt_context_render + rt_context_download for one shard (local row order)
RT_API int emu_render_shard(const rt_scene* s, const rt_render_opts* o, uint8_t* bgr_rows, int32_t* ids, double* accum_out, rt_stats* stats) {
    EmuScene e;
    int rc = make_scene(s, o->accel, o->tuning, &e);
    if (rc != RT_OK) return rc;
    FrameParams fp{};
    if ((rc = make_frame_params(s, o, &fp)) != RT_OK) return rc;
    const uint32_t bytewidth = rt_bmp_bytewidth(fp.width);
    const size_t npix = (size_t)fp.local_rows * fp.width;
    std::vector<double> accum(npix * 3, 0.0);
    std::vector<int32_t> idbuf(npix, -1);
    rt_stats st{};
    TraverseCount tc_closest{0, 0}, tc_shadow{0, 0};
    const uint64_t total_paths = fp.npix_padded * fp.antialias * fp.cam_samples;
    std::vector<QRay> cur, next;
    struct SRay { ShadowRay r; uint32_t pixel; };
    std::vector<SRay> shadows;
    auto add = [&](uint32_t pixel, C3 c) {
        if (c.r == 0.0 && c.g == 0.0 && c.b == 0.0) return;
        accum[(size_t)pixel * 3] += c.r; accum[(size_t)pixel * 3 + 1] += c.g; accum[(size_t)pixel * 3 + 2] += c.b;
    };
    // primary rays, sample plane by sample plane
    const uint64_t batch = 1u << 20;
    for (uint64_t base = 0; base < total_paths; base += batch) {
        cur.clear();
        const uint64_t end = std::min(total_paths, base + batch);
        for (uint64_t g = base; g < end; g++) {
            const uint64_t plane = g / fp.npix_padded, slot = g - plane * fp.npix_padded;
            const uint32_t aa = (uint32_t)(plane / fp.cam_samples), cs = (uint32_t)(plane - (uint64_t)aa * fp.cam_samples);
            uint32_t x, y, lp;
            if (!slot_to_pixel(fp, slot, x, y, lp)) continue;
            const uint64_t sk = key_sample(fp.seed, (uint64_t)y * fp.width + x, aa);
            double px, py;
            pixel_position(fp, x, y, sk, px, py);
            QRay r;
            r.key = child_key(sk, CODE_CAMERA0 + cs);
            camera_project(e.ds.camera, px, py, r.key, r.o, r.d);
            r.T = c3(1.0, 1.0, 1.0); r.sig = 1.0; r.pixel = lp;
            r.meta = (aa == 0 && cs == 0) ? META_RECORD_ID : 0u;
            cur.push_back(r);
        }
        st.paths += cur.size();
        for (uint32_t depth = 0; !cur.empty(); depth++) {
            if (depth == 0) st.rays_primary += cur.size(); else st.rays_secondary += cur.size();
            next.clear(); shadows.clear();
            for (const QRay& r : cur) {
                RayQ rq = make_rayq(r.o, r.d);
                int32_t id; double t;
                if (e.use_bvh) closest_bvh(e.ds, rq, id, t, tc_closest); else closest_linear(e.ds, rq, id, t);
                if (r.meta & META_RECORD_ID) idbuf[r.pixel] = id;
                ShadeCtx sc;
                C3 a = shade_begin(e.ds, fp, r.o, r.d, r.T, r.sig, r.key, r.meta & META_DEPTH_MASK, id, t, sc);
                add(r.pixel, a);
                for (uint32_t li = 0; li < e.ds.n_lights; li++) {
                    SRay sr;
                    if (shade_light(e.ds, sc, li, sr.r)) { sr.pixel = r.pixel; shadows.push_back(sr); }
                }
                for (uint32_t k = 0; k < (uint32_t)e.ds.max_children; k++) {
                    ChildRay cr;
                    if (shade_child(e.ds, sc, k, cr)) {
                        QRay q;
                        q.o = cr.o; q.d = cr.d; q.T = cr.T; q.sig = cr.sig; q.key = cr.key; q.pixel = r.pixel;
                        q.meta = (r.meta & META_DEPTH_MASK) + 1u;
                        next.push_back(q);
                    }
                }
            }
            st.rays_shadow += shadows.size();
            for (const SRay& sr : shadows) {
                RayQ rq = make_rayq(sr.r.o, sr.r.d);
                const bool occ = e.use_bvh ? shadow_bvh(e.ds, rq, sr.r.range2, tc_shadow) : shadow_linear(e.ds, rq, sr.r.range2);
                if (!occ) add(sr.pixel, sr.r.c);
            }
            cur.swap(next);
        }
    }
    st.rays_total = st.rays_primary + st.rays_secondary + st.rays_shadow;
    st.bvh_nodes_visited[0] = tc_closest.nodes; st.bvh_nodes_visited[1] = tc_shadow.nodes;
    st.bvh_prims_tested[0] = tc_closest.prims; st.bvh_prims_tested[1] = tc_shadow.prims;
    st.local_rows = fp.local_rows; st.bytewidth = bytewidth;
    st.accel_used = e.use_bvh ? RT_ACCEL_BVH : RT_ACCEL_LINEAR;
    st.n_bvh_nodes = e.nodes.size();
    st.bvh_node_bytes = !e.use_bvh ? 0 : (e.ds.bvh_qnodes ? sizeof(BvhQNode) : sizeof(BvhNode));
    if (bgr_rows) {
        std::memset(bgr_rows, 0, (size_t)fp.local_rows * bytewidth);
        for (uint32_t y = 0; y < fp.local_rows; y++)
            for (uint32_t x = 0; x < fp.width; x++)
                for (int ch = 0; ch < 3; ch++) {
                    const double a = accum[((size_t)y * fp.width + x) * 3 + (2 - ch)];
                    bgr_rows[(size_t)y * bytewidth + 3 * x + ch] = to_srgb(SRGB_AVERAGE, (a / (double)fp.cam_samples) / (double)fp.antialias);
                }
    }
    if (ids) std::memcpy(ids, idbuf.data(), npix * sizeof(int32_t));
    if (accum_out) std::memcpy(accum_out, accum.data(), npix * 3 * sizeof(double));
    if (stats) *stats = st;
    return RT_OK;
}

RT_API int emu_trace_rays(const rt_scene* s, const rt_render_opts* o, uint64_t n, const double* od, int32_t* ids, double* t_out, double* normal) {
    EmuScene e;
    int rc = make_scene(s, o->accel, o->tuning, &e);
    if (rc != RT_OK) return rc;
    for (uint64_t i = 0; i < n; i++) {
        V3 org = v3(od + 6 * i), dir = v3(od + 6 * i + 3);
        RayQ rq = make_rayq(org, dir);
        int32_t id; double t;
        if (e.use_bvh) closest_bvh(e.ds, rq, id, t, g_tc); else closest_linear(e.ds, rq, id, t);
        ids[i] = id;
        if (t_out) t_out[i] = id >= 0 ? t : (double)INFINITY;
        if (normal) {
            V3 nn = id >= 0 ? shape_normal(e.ds, id, org, dir, t) : v3(0, 0, 0);
            normal[3 * i] = nn.x; normal[3 * i + 1] = nn.y; normal[3 * i + 2] = nn.z;
        }
    }
    return RT_OK;
}

// shadow queries: occluded[i] for ray i with squared range r2[i] (< 0 = none)
RT_API int emu_shadow_rays(const rt_scene* s, const rt_render_opts* o, uint64_t n, const double* od, const double* r2, uint8_t* occluded) {
    EmuScene e;
    int rc = make_scene(s, o->accel, o->tuning, &e);
    if (rc != RT_OK) return rc;
    for (uint64_t i = 0; i < n; i++) {
        RayQ rq = make_rayq(v3(od + 6 * i), v3(od + 6 * i + 3));
        occluded[i] = (e.use_bvh ? shadow_bvh(e.ds, rq, r2[i], g_tc) : shadow_linear(e.ds, rq, r2[i])) ? 1 : 0;
    }
    return RT_OK;
}

RT_API void emu_untile(const uint32_t* gathered, uint32_t* out, uint32_t height, uint32_t words_per_row, uint32_t band_rows, uint32_t world,
                       uint32_t local_rows) {
    const uint64_t total = (uint64_t)height * words_per_row;
    for (uint64_t w = 0; w < total; w++) out[w] = gathered[untile_source_word(w, words_per_row, band_rows, world, local_rows)];
}

RT_API void emu_quantize_bgr(const double* rgb, uint64_t n, uint8_t* bgr) {
    for (uint64_t i = 0; i < n; i++) {
        bgr[3 * i] = to_srgb(SRGB_AVERAGE, rgb[3 * i + 2]);
        bgr[3 * i + 1] = to_srgb(SRGB_AVERAGE, rgb[3 * i + 1]);
        bgr[3 * i + 2] = to_srgb(SRGB_AVERAGE, rgb[3 * i]);
    }
}

RT_API int emu_bvh_stats(const rt_scene* s, uint64_t* n_nodes, uint32_t* max_depth) {
    EmuScene e;
    int rc = make_scene(s, RT_ACCEL_BVH, 0, &e);
    if (rc != RT_OK) return rc;
    *n_nodes = e.nodes.size();
    uint32_t md = 0;
    if (e.nodes.size() > 0 && e.flat.sph.size() > 1) {
        std::vector<std::pair<int32_t, uint32_t>> st{{0, 1}};
        while (!st.empty()) {
            auto [nd, d] = st.back();
            st.pop_back();
            md = std::max(md, d);
            if (e.nodes[nd].c0 >= 0) st.push_back({e.nodes[nd].c0, d + 1});
            if (e.nodes[nd].c1 >= 0) st.push_back({e.nodes[nd].c1, d + 1});
        }
    }
    *max_depth = md;
    return RT_OK;
}

// the scene's flatten cache (host_scene.h FlatCache), as rt_context_create sees it: out = {entry identity, revision,
// spheres, planes, objects, max_children}
RT_API int emu_flat_cache_probe(const rt_scene* s, uint32_t tuning, uint64_t* out) {
    std::shared_ptr<FlatCache> fc;
    int rc = flatten_scene_cached(s, tuning, &fc);
    if (rc != RT_OK) return rc;
    out[0] = (uint64_t)(uintptr_t)fc.get();
    out[1] = fc->revision;
    out[2] = fc->flat.sph.size();
    out[3] = fc->flat.planes.size();
    out[4] = fc->flat.obj_shape.size();
    out[5] = (uint64_t)fc->flat.max_children;
    return RT_OK;
}

RT_API int emu_bvh_node_bytes(const rt_scene* s, uint32_t tuning, uint32_t* bytes) {
    EmuScene e;
    int rc = make_scene(s, RT_ACCEL_BVH, tuning, &e);
    if (rc != RT_OK) return rc;
    *bytes = !e.use_bvh ? 0u : (e.ds.bvh_qnodes ? (uint32_t)sizeof(BvhQNode) : (uint32_t)sizeof(BvhNode));
    return RT_OK;
}

}  // extern "C"
