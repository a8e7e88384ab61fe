// context.cu -- device-resident scene + the host driver of the wavefront (replaces the frame loop of
// src/main.rs:39-59 and the recursion of src/raytrace.rs:261-276).
//
// The recursion `ray_color -> material.color -> ray_color` becomes a stack of ray queues, one per depth.
// A level is processed in slices small enough that whatever the slice spawns fits the next level's queue and
// the shadow queue, so memory is bounded for any branching factor (TransparentMaterial doubles, IndirectPhong
// multiplies by `samples`):
//
//   process(level d, n rays):  for each slice:  closest-hit -> shade (emits shadow rays + level d+1 rays)
//                                               -> shadow kernel -> process(level d+1)
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstddef>
#include <functional>
#include <memory>
#include <thread>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/raytrace_b200.h"
#include "host_scene.h"
#include "kernels.h"

using namespace rt;

#define CUDA_TRY(expr)                                                                                     \
    do {                                                                                                   \
        cudaError_t e_ = (expr);                                                                           \
        if (e_ != cudaSuccess) {                                                                           \
            set_error(std::string(#expr) + ": " + cudaGetErrorString(e_));                                 \
            return RT_ERR_CUDA;                                                                            \
        }                                                                                                  \
    } while (0)

namespace {

// Process-wide cache of device allocations.  The one-call entry point rt_render builds a context per frame (upload, BVH,
// ray queues of several GB); paying cudaMalloc/cudaFree for those every frame costs more than the trace itself, so blocks
// released by a context are kept (per device, by size) and handed to the next one.  rt_release_cached_memory() returns
// them to the driver.
void (*g_drop_cached_multi)() = nullptr;  // drops the contexts / rt_multi objects kept between one-shot calls (set at the end of this file)

class DeviceCache {
public:
    cudaError_t alloc(void** p, size_t bytes) {
        *p = nullptr;
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        const size_t sz = (bytes + kGranule - 1) / kGranule * kGranule;
        {
            std::lock_guard<std::mutex> g(m_);
            auto it = free_.find({dev, sz});
            if (it != free_.end()) {
                *p = it->second;
                free_.erase(it);
                cached_ -= sz;
                live_[*p] = {dev, sz};
                return cudaSuccess;
            }
        }
        e = cudaMalloc(p, sz);
        if (e != cudaSuccess) {  // out of memory with blocks parked in the cache (or held by contexts kept between one-shot calls): give them back and retry once
            cudaGetLastError();
            if (g_drop_cached_multi) g_drop_cached_multi();
            trim();
            e = cudaMalloc(p, sz);
            if (e != cudaSuccess) return e;
        }
        std::lock_guard<std::mutex> g(m_);
        live_[*p] = {dev, sz};
        return cudaSuccess;
    }
    void release(void* p) {
        if (!p) return;
        std::lock_guard<std::mutex> g(m_);
        auto it = live_.find(p);
        if (it == live_.end()) { cudaFree(p); return; }
        free_.insert({{it->second.first, it->second.second}, p});
        cached_ += it->second.second;
        live_.erase(it);
    }
    void trim() {
        std::lock_guard<std::mutex> g(m_);
        int cur = 0;
        cudaGetDevice(&cur);
        for (auto& kv : free_) { cudaSetDevice(kv.first.first); cudaFree(kv.second); }
        cudaSetDevice(cur);
        free_.clear();
        cached_ = 0;
    }
    size_t cached_bytes() { std::lock_guard<std::mutex> g(m_); return cached_; }

private:
    static constexpr size_t kGranule = 1u << 16;
    std::mutex m_;
    std::multimap<std::pair<int, size_t>, void*> free_;
    std::unordered_map<void*, std::pair<int, size_t>> live_;
    size_t cached_ = 0;
};
DeviceCache g_cache;

// pinned read-back blocks (Counters records of a context), recycled for the same reason
class PinnedPool {
public:
    cudaError_t get(Counters** p, size_t* slots, size_t want) {
        {
            std::lock_guard<std::mutex> g(m_);
            for (size_t i = 0; i < free_.size(); i++)
                if (free_[i].second >= want) {
                    *p = free_[i].first; *slots = free_[i].second;
                    free_.erase(free_.begin() + i);
                    return cudaSuccess;
                }
        }
        *slots = want < 4 ? 4 : want;
        return cudaMallocHost((void**)p, *slots * sizeof(Counters));
    }
    void put(Counters* p, size_t slots) {
        if (!p) return;
        std::lock_guard<std::mutex> g(m_);
        free_.push_back({p, slots});
    }
    void trim() {
        std::lock_guard<std::mutex> g(m_);
        for (auto& b : free_) cudaFreeHost(b.first);
        free_.clear();
    }
private:
    std::mutex m_;
    std::vector<std::pair<Counters*, size_t>> free_;
};
PinnedPool g_pinned;
const DeviceAllocator g_cache_allocator = {[](void** p, size_t bytes) { return g_cache.alloc(p, bytes); }, [](void* p) { g_cache.release(p); }};

template <typename T>
int dev_alloc(T** p, size_t count) {
    *p = nullptr;
    if (count == 0) return RT_OK;
    CUDA_TRY(g_cache.alloc((void**)p, count * sizeof(T)));
    return RT_OK;
}
// Page-locks the big vectors of a cache entry, once (portable: every device of the process copies from them).  Small
// scenes stay pageable: the registration would cost more than it saves.
struct HostPins {
    std::vector<void*> regions;
    ~HostPins() {
        for (void* p : regions) cudaHostUnregister(p);
        cudaGetLastError();
    }
};
void pin_flat_cache(const rt_scene* s, FlatCache* fc) {
    std::lock_guard<std::mutex> g(s->flat_cache.m);
    if (fc->pin) return;
    const FlatScene& f = fc->flat;
    size_t total = f.sph.size() * sizeof(SphereRec) + (f.sph_obj.size() + f.obj_shape.size() + f.obj_material.size()) * 4 +
                   s->materials.size() * sizeof(MaterialRec);
    if (s->bg_kind == BG_SKYBOX)
        for (int k = 0; k < 6; k++) total += s->faces[k].data.size();
    std::shared_ptr<HostPins> pins = std::make_shared<HostPins>();
    if (total >= (4u << 20)) {
        auto reg = [&](const void* p, size_t bytes) {
            if (!bytes) return;
            if (cudaHostRegister(const_cast<void*>(p), bytes, cudaHostRegisterPortable) == cudaSuccess) pins->regions.push_back(const_cast<void*>(p));
            else cudaGetLastError();  // not fatal: that buffer is copied from pageable memory
        };
        reg(f.sph.data(), f.sph.size() * sizeof(SphereRec));
        reg(f.sph_obj.data(), f.sph_obj.size() * sizeof(int32_t));
        reg(f.obj_shape.data(), f.obj_shape.size() * sizeof(uint32_t));
        reg(f.obj_material.data(), f.obj_material.size() * sizeof(uint32_t));
        // the scene's own vectors: rt::scene_touch drops this entry before anything reallocates them
        reg(s->materials.data(), s->materials.size() * sizeof(MaterialRec));
        if (s->bg_kind == BG_SKYBOX)
            for (int k = 0; k < 6; k++) reg(s->faces[k].data.data(), s->faces[k].data.size());
    }
    fc->pin = pins;
}

// device copy of a scene: flattened buffers + optional BVH.  Shared (read-only) by every context created from the same
// scene revision on the same device: rt_scene keeps the entries (host_scene.h DeviceSceneSlot).
struct DeviceSceneHolder {
    DeviceScene ds{};  // the device pointers and BVH fields; the by-value header fields are refreshed per context
    std::vector<void*> allocs;
    BvhBuildResult bvh{};
    bool use_bvh = false;
    double bvh_ms = 0.0;
    uint32_t build_launches = 0;
    int device = 0;
    // a device-driven frame of this scene overflowed a queue: later contexts (rt_render makes one per frame) go level by
    // level from the start instead of paying for the failed attempt again
    std::atomic<bool> dd_overflowed{false};

    ~DeviceSceneHolder() { release(); }
    void release() {
        for (void* p : allocs) g_cache.release(p);
        allocs.clear();
        bvh_free(&bvh, g_cache_allocator);
    }

    // the flat buffers come from the scene's cache (host_scene.h FlatCache) and, for big scenes, from page-locked memory:
    // copies on the context's stream, the BVH build queued right behind them
    // Arrays that are not page-locked where they lie (small scenes, the tables) go through a page-locked staging block of
    // the calling thread: a copy from pageable memory is staged by the driver under a lock that the upload threads of all
    // GPUs share (eight devices re-uploading a 10k-sphere scene spent more time there than in their BVH builds).
    struct Staging {
        unsigned char* p = nullptr;
        size_t cap = 0, used = 0;
        ~Staging() { if (p) cudaFreeHost(p); }
        void* take(size_t bytes) {  // nullptr: does not fit (the caller copies from where the data lies)
            const size_t at = (used + 255) & ~(size_t)255;
            if (at + bytes > cap) return nullptr;
            used = at + bytes;
            return p + at;
        }
    };
    static Staging& staging(size_t want) {
        static thread_local Staging s;
        s.used = 0;
        if (s.cap < want) {
            if (s.p) cudaFreeHost(s.p);
            s.p = nullptr; s.cap = 0;
            const size_t cap = std::max<size_t>(want + (want >> 2), 1u << 20);
            if (cudaMallocHost((void**)&s.p, cap) == cudaSuccess) s.cap = cap;
            else cudaGetLastError();
        }
        return s;
    }
    template <typename T>
    int up_flat(const std::vector<T>& v, const T** out, cudaStream_t st, Staging* stage = nullptr) {
        T* p = nullptr;
        int rc = dev_alloc(&p, v.size());
        if (rc != RT_OK) return rc;
        if (p) {
            allocs.push_back(p);
            const void* src = v.data();
            if (stage) {
                if (void* h = stage->take(v.size() * sizeof(T))) { std::memcpy(h, v.data(), v.size() * sizeof(T)); src = h; }
            }
            CUDA_TRY(cudaMemcpyAsync(p, src, v.size() * sizeof(T), cudaMemcpyHostToDevice, st));
        }
        *out = p;
        return RT_OK;
    }

    int init(const rt_scene* s, int accel, uint32_t tuning, cudaStream_t st) {
        // Upload and build run on a stream of their own at the device's highest priority.  The build waits for the host three
        // times (grid check, depth, end); when another caller's frame is being traced on this GPU (two frames in flight, each
        // with its own scene revision), each of those phases would otherwise queue behind that whole frame's launches -- its
        // resident grids leave nothing free -- instead of slipping in at the next kernel boundary (a cold two-caller frame at
        // 1M spheres took 73 ms instead of 42).
        // One such stream per host thread and device, made on first use and kept (the callers are long-lived: the worker threads
        // of an rt_multi, the caller's own thread for rt_render; creating and destroying one per call cost 0.5 ms per frame at
        // eight GPUs).
        static thread_local cudaStream_t tl_hp[64] = {};
        cudaStream_t hp = (device >= 0 && device < 64) ? tl_hp[device] : nullptr;
        if (!hp && device >= 0 && device < 64) {
            int lo = 0, hi = 0;
            if (cudaDeviceGetStreamPriorityRange(&lo, &hi) == cudaSuccess && cudaStreamCreateWithPriority(&hp, cudaStreamNonBlocking, hi) == cudaSuccess) tl_hp[device] = hp;
            else { cudaGetLastError(); hp = nullptr; }
        }
        const cudaStream_t use = hp ? hp : st;
        const int rc = init_queued(s, accel, tuning, use);
        // nothing of this holder may still be in flight when its blocks go back to the cache (early return), and the flat
        // cache entry the copies read from may go away (scene changed or freed) once this returns
        cudaStreamSynchronize(use);
        return rc;
    }

private:
    int init_queued(const rt_scene* s, int accel, uint32_t tuning, cudaStream_t st) {
        std::shared_ptr<FlatCache> fc;
        int rc = flatten_scene_cached(s, tuning, &fc);
        if (rc != RT_OK) return rc;
        pin_flat_cache(s, fc.get());
        const FlatScene& f = fc->flat;
        fill_scene_header(s, f, &ds);
        const double* clo = f.centroid_lo;
        const double* chi = f.centroid_hi;
        // big scenes are page-locked in place (pin_flat_cache): only their small arrays are staged
        const bool pinned = fc->pin && !std::static_pointer_cast<HostPins>(fc->pin)->regions.empty();
        size_t small = (f.planes.size() * sizeof(PlaneRec) + s->lights.size() * sizeof(LightRec) + 511 * sizeof(double)) + 16 * 256;
        if (!pinned) {
            small += f.sph.size() * sizeof(SphereRec) + (f.sph_obj.size() + f.obj_shape.size() + f.obj_material.size()) * 4 + s->materials.size() * sizeof(MaterialRec);
            if (s->bg_kind == BG_SKYBOX)
                for (int k = 0; k < 6; k++) small += s->faces[k].data.size();
        }
        Staging& stg = staging(small);
        Staging* big = pinned ? nullptr : &stg;
        if ((rc = up_flat(f.sph, &ds.sph, st, big)) != RT_OK) return rc;
        if ((rc = up_flat(f.sph_obj, &ds.sph_obj, st, big)) != RT_OK) return rc;
        if ((rc = up_flat(f.planes, &ds.planes, st, &stg)) != RT_OK) return rc;
        if ((rc = up_flat(f.obj_shape, &ds.obj_shape, st, big)) != RT_OK) return rc;
        if ((rc = up_flat(f.obj_material, &ds.obj_material, st, big)) != RT_OK) return rc;
        if ((rc = up_flat(s->materials, &ds.materials, st, big)) != RT_OK) return rc;
        if ((rc = up_flat(s->lights, &ds.lights, st, &stg)) != RT_OK) return rc;
        if (s->bg_kind == BG_SKYBOX)
            for (int k = 0; k < 6; k++) {
                const uint8_t* p = nullptr;
                if ((rc = up_flat(s->faces[k].data, &p, st, big)) != RT_OK) return rc;
                ds.faces[k].data = p;
            }
        std::vector<double> vals(SRGB_VALUES, SRGB_VALUES + 256), avg(SRGB_AVERAGE, SRGB_AVERAGE + 255);
        if ((rc = up_flat(vals, &ds.srgb_values, st, &stg)) != RT_OK) return rc;
        if ((rc = up_flat(avg, &ds.srgb_average, st, &stg)) != RT_OK) return rc;
        CUDA_TRY(cudaStreamSynchronize(st));  // vals / avg are locals; the staging block is this thread's next upload's too
        use_bvh = accel_uses_bvh(accel, ds.n_spheres);
        if (use_bvh) {
            cudaEvent_t e0 = nullptr, e1 = nullptr;
            cudaError_t e = cudaEventCreate(&e0);
            if (e == cudaSuccess) e = cudaEventCreate(&e1);
            if (e == cudaSuccess) e = cudaEventRecord(e0, st);
            // shadow queries walk four-wide nodes derived from the grid nodes (rt_types.h BvhWNode) unless tuned off
            const bool wide = !(tuning & RT_TUNE_BINARY_SHADOW);
            if (e == cudaSuccess)
                e = bvh_build(st, ds.sph, ds.sph_obj, ds.n_spheres, clo, chi, f.quant ? f.q_org : nullptr, f.quant ? f.q_cell : nullptr, g_cache_allocator, &bvh,
                              wide, (tuning & RT_TUNE_MORTON_AXIS) != 0);
            if (e == cudaSuccess) e = cudaEventRecord(e1, st);
            if (e == cudaSuccess) e = cudaEventSynchronize(e1);
            float ms = 0;
            if (e == cudaSuccess) cudaEventElapsedTime(&ms, e0, e1);
            if (e0) cudaEventDestroy(e0);
            if (e1) cudaEventDestroy(e1);
            if (e != cudaSuccess) { set_error(std::string("bvh_build: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
            bvh_ms = ms;
            ds.bvh_nodes = bvh.nodes; ds.bvh_qnodes = bvh.qnodes; ds.bvh_wnodes = bvh.wnodes; ds.bvh_sph = bvh.sph; ds.bvh_obj = bvh.obj; ds.n_bvh_nodes = bvh.n_nodes;
            build_launches = bvh.launches;
        }
        return RT_OK;
    }
};

constexpr uint32_t SCENE_TUNING = RT_TUNE_NODES_F32 | RT_TUNE_NODES_Q16 | RT_TUNE_MORTON_AXIS | RT_TUNE_BINARY_SHADOW;

// The device copy of `s` for (device, accel, tuning): the one kept with the scene if its revision still matches, else a
// fresh upload + BVH build (which then replaces it).  `fresh` tells the caller whether the build happened in this call.
int acquire_device_scene(const rt_scene* s, int accel, uint32_t tuning, int device, cudaStream_t st, std::shared_ptr<DeviceSceneHolder>* out, bool* fresh) {
    const uint32_t tn = tuning & SCENE_TUNING;
    {
        std::lock_guard<std::mutex> g(s->device_cache.m);
        for (const DeviceSceneSlot::Entry& e : s->device_cache.entries)
            if (e.device == device && e.accel == accel && e.tuning == tn && e.revision == s->revision) {
                *out = std::static_pointer_cast<DeviceSceneHolder>(e.holder);
                *fresh = false;
                return RT_OK;
            }
    }
    std::shared_ptr<DeviceSceneHolder> h = std::make_shared<DeviceSceneHolder>();
    h->device = device;
    const uint64_t rev = s->revision;
    int rc = h->init(s, accel, tn, st);
    if (rc != RT_OK) return rc;
    {
        std::lock_guard<std::mutex> g(s->device_cache.m);
        auto& v = s->device_cache.entries;
        for (size_t i = 0; i < v.size();)
            if (v[i].revision != s->revision || (v[i].device == device && v[i].accel == accel && v[i].tuning == tn)) v.erase(v.begin() + i);
            else i++;
        if (rev == s->revision) v.push_back({device, accel, tn, rev, h});
    }
    *out = h;
    *fresh = true;
    return RT_OK;
}

int alloc_ray_queue(RayQueue* q, uint32_t cap, std::vector<void*>* allocs, bool full) {
    std::memset(q, 0, sizeof *q);
    q->cap = cap;
    if (cap == 0) return RT_OK;
    const size_t c = cap;
    const size_t n_d = full ? 11 : 7;  // doubles per ray: o, d, (T, sig), hit_t
    const size_t bytes = c * (n_d * 8 + (full ? 8 + 4 + 4 : 0) + 4);
    unsigned char* base = nullptr;
    CUDA_TRY(g_cache.alloc((void**)&base, bytes));
    allocs->push_back(base);
    double* d = (double*)base;
    q->ox = d; q->oy = d + c; q->oz = d + 2 * c; q->dx = d + 3 * c; q->dy = d + 4 * c; q->dz = d + 5 * c;
    q->hit_t = d + 6 * c;
    unsigned char* p = (unsigned char*)(d + 7 * c);
    if (full) {
        q->tr = d + 7 * c; q->tg = d + 8 * c; q->tb = d + 9 * c; q->sig = d + 10 * c;
        q->key = (uint64_t*)(d + 11 * c);
        p = (unsigned char*)(q->key + c);
        q->pixel = (uint32_t*)p; p += 4 * c;
        q->meta = (uint32_t*)p; p += 4 * c;
    }
    q->hit_id = (int32_t*)p;
    return RT_OK;
}

int alloc_shadow_queue(ShadowQueue* q, uint32_t cap, std::vector<void*>* allocs) {
    std::memset(q, 0, sizeof *q);
    q->cap = cap;
    if (cap == 0) return RT_OK;
    const size_t c = cap;
    unsigned char* base = nullptr;
    CUDA_TRY(g_cache.alloc((void**)&base, c * (10 * 8 + 4)));
    allocs->push_back(base);
    double* d = (double*)base;
    q->ox = d; q->oy = d + c; q->oz = d + 2 * c; q->dx = d + 3 * c; q->dy = d + 4 * c; q->dz = d + 5 * c;
    q->r2 = d + 6 * c; q->cr = d + 7 * c; q->cg = d + 8 * c; q->cb = d + 9 * c;
    q->pixel = (uint32_t*)(d + 10 * c);
    return RT_OK;
}

}  // namespace

struct rt_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    int sm_count = 148;
    std::shared_ptr<DeviceSceneHolder> scene;  // shared, read-only
    DeviceScene ds{};                          // the holder's pointers + this scene's by-value header (camera, background)
    bool use_bvh = false;
    uint32_t* heavy = nullptr;  // list of rays handed from the BVH traversal to the block-wide scan (kernels.h HEAVY_CAP)
    const KernelTable* k = nullptr;
    FrameParams fp{};
    rt_render_opts opts{};
    uint32_t bytewidth = 0;
    std::vector<void*> allocs;
    std::vector<RayQueue> levels;
    ShadowQueue shadowq{};
    RayQueue null_queue{};
    Counters* counters = nullptr;       // device
    Counters* counters_host = nullptr;  // pinned: n_host_slots records (one per primary batch of a device-driven frame)
    size_t n_host_slots = 0;
    size_t final_slot = 0;        // counters_host[final_slot]: the record copied out behind the resolve kernel (frame-wide counters)
    size_t dd_batches = 0;        // primary batches of the device-driven frame in flight
    // RT_TUNE_FRAME_GRAPH: a device-driven BVH frame is the same ~70 launches, memsets and copies every time (counts live on
    // the device), so once two consecutive frames were queued from identical arguments the whole frame is captured into a CUDA
    // graph and replayed with one launch call (frame_graph_*).  A re-bound scene (new device pointers), another frame buffer
    // or profiling go back to plain launches.  Opt-in: the replay takes rt_context_render_begin from 0.30 to 0.01 ms of host
    // time, but the graph's fork / join per level runs slower on the device than the same launches on two streams (1/8 shard
    // 2.66 against 2.43 ms; whole frames equal: profiles/r02_v9_graph_ab.txt) -- for callers whose host thread is the bottleneck.
    cudaGraphExec_t graph_exec = nullptr;
    DeviceScene graph_ds{};           // what the graph was captured from / what the previous frame was queued from
    uint8_t* graph_out = nullptr;
    bool graph_key_valid = false;     // graph_ds / graph_out describe the previous frame
    bool graph_failed = false;        // capture or instantiation failed once: plain launches from then on
    uint64_t graph_kernel_launches = 0;
    bool in_flight = false;       // rt_context_render_begin has queued a frame that rt_context_render_end has not collected yet
    bool in_flight_dd = false;    // ... and it is a device-driven one (its counts are still to be read)
    double* accum = nullptr;
    int32_t* ids = nullptr;
    uint8_t* framebuffer = nullptr;
    size_t fb_bytes = 0;
    uint8_t* final_out = nullptr;  // non-null: k_resolve writes BMP-ordered rows of the whole frame here (rt_multi peer gather)
    uint8_t* frame_own = nullptr;  // rt_context_frame_create
    void* frame_ipc = nullptr;     // rt_context_frame_import
    uint32_t slice = 0;
    uint32_t shadow_queued = 0;  // shadow rays appended since the last trace of the shadow queue
    uint64_t total_paths = 0;
    bool host_levels = false;    // wavefront driven level by level from the host (RT_TUNE_HOST_LEVELS, or after an overflow)
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_counts = nullptr;
    rt_stats stats{};
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream2 = nullptr;      // device-driven frames: shadow segments run here beside the level loop (frame_device_driven)
    uint32_t* heavy2 = nullptr;          // the heavy-ray list of the launches on stream2
    std::vector<cudaEvent_t> ev_seg;     // [level]: that level's shade pass (and its mark) are done
    cudaEvent_t ev_join = nullptr;       // every shadow segment of the batch is done
    bool profiling = false;
    std::vector<cudaEvent_t> prof_events;       // pool, two per timed launch
    std::vector<int> prof_class;                // class of launch i (events 2i, 2i+1)

    // brackets one kernel-class launch with an event pair while profiling is on
    void prof_begin(int cls) {
        if (!profiling) return;
        const size_t i = prof_class.size();
        while (prof_events.size() < 2 * (i + 1)) { cudaEvent_t e; cudaEventCreate(&e); prof_events.push_back(e); }
        prof_class.push_back(cls);
        cudaEventRecord(prof_events[2 * i], stream);
    }
    void prof_end() {
        if (!profiling) return;
        cudaEventRecord(prof_events[2 * (prof_class.size() - 1) + 1], stream);
    }

    ~rt_context() {
        cudaSetDevice(device);
        if (stream) cudaStreamSynchronize(stream);  // blocks go back to the cache: nothing may still be using them
        if (stream2) cudaStreamSynchronize(stream2);
        for (void* p : allocs) g_cache.release(p);
        if (frame_own) cudaFree(frame_own);  // exported to other processes: not recycled through the cache
        if (frame_ipc) cudaIpcCloseMemHandle(frame_ipc);
        g_pinned.put(counters_host, n_host_slots);
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
        if (ev_counts) cudaEventDestroy(ev_counts);
        for (cudaEvent_t e : prof_events) cudaEventDestroy(e);
        if (graph_exec) cudaGraphExecDestroy(graph_exec);
        for (cudaEvent_t e : ev_seg) cudaEventDestroy(e);
        if (ev_join) cudaEventDestroy(ev_join);
        scene.reset();
        if (stream2) cudaStreamDestroy(stream2);
        if (own_stream) cudaStreamDestroy(own_stream);
    }
};

namespace {

// ---------------------------------------------------------------------------------------------------------
// host-driven wavefront: the host reads every level's ray count back, slices levels that would not fit the next
// queue, and flushes the shadow queue when it could overflow.  Never overflows by construction; it is the fallback of
// the device-driven frame below and what RT_TUNE_HOST_LEVELS selects.
// ---------------------------------------------------------------------------------------------------------

// Count read-back in two halves: the copy is queued (with an event) right behind the kernel that produced the counts, the
// host may queue more work behind it, and only then waits -- for the event, not for the stream.
int ctx_queue_counters(rt_context* c) {
    CUDA_TRY(cudaMemcpyAsync(c->counters_host, c->counters, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaEventRecord(c->ev_counts, c->stream));
    return RT_OK;
}
int ctx_wait_counters(rt_context* c) {
    CUDA_TRY(cudaEventSynchronize(c->ev_counts));
    if (c->counters_host->overflow) { set_error("internal error: ray queue or traversal stack overflow"); return RT_ERR_CUDA; }
    return RT_OK;
}

// The closest-hit pass over a whole level, queued BEFORE the host knows how many rays the level holds: the kernel reads the
// count from the device counter its producer has just finished (stream order).
bool trace_level_ahead(rt_context* c, uint32_t d) {
    if (!c->use_bvh || d >= c->levels.size() || c->levels[d].cap == 0) return false;
    c->prof_begin(1);
    c->k->closest(c->stream, c->ds, true, c->levels[d], 0, c->levels[d].cap, c->counters, c->sm_count, c->heavy, &c->counters->level_count[d],
                  &c->counters->lc[LC_SHARED], true);
    c->prof_end();
    return true;
}

// traces what the shadow queue holds and empties it
int flush_shadow(rt_context* c) {
    if (c->shadow_queued) {
        c->prof_begin(3);
        c->k->shadow(c->stream, c->ds, c->use_bvh, c->shadowq, c->shadow_queued, c->accum, c->counters, c->sm_count, c->heavy, nullptr,
                     &c->counters->lc[LC_SHARED], true, nullptr);
        c->prof_end();
        c->stats.kernel_launches += c->use_bvh ? 2 : 1;  // shadow (+ heavy scan)
        c->shadow_queued = 0;
    }
    CUDA_TRY(cudaMemsetAsync(&c->counters->shadow_count, 0, sizeof(uint32_t), c->stream));
    return RT_OK;
}

// `traced`: the closest-hit pass over ALL rays of this level has already been queued (trace_level_ahead)
int process_level(rt_context* c, uint32_t d, uint32_t n, bool traced) {
    const bool last = d + 1 >= c->levels.size();
    RayQueue& in = c->levels[d];
    RayQueue& out = last ? c->null_queue : c->levels[d + 1];
    uint32_t* next_count = &c->counters->level_count[last ? 63 : d + 1];
    const uint64_t lights = std::max(1u, c->ds.n_lights);
    for (uint32_t first = 0; first < n; first += c->slice) {
        const uint32_t cnt = std::min(c->slice, n - first);
        if (!traced) {
            c->prof_begin(1);
            c->k->closest(c->stream, c->ds, c->use_bvh, in, first, cnt, c->counters, c->sm_count, c->heavy, nullptr, &c->counters->lc[LC_SHARED], true);
            c->prof_end();
        }
        CUDA_TRY(cudaMemsetAsync(next_count, 0, sizeof(uint32_t), c->stream));
        if ((uint64_t)c->shadow_queued + (uint64_t)cnt * lights > c->shadowq.cap) {  // this slice might not fit behind what is queued
            int rc = flush_shadow(c);
            if (rc != RT_OK) return rc;
        }
        c->prof_begin(2);
        c->k->shade(c->stream, c->ds, c->fp, in, first, cnt, out, next_count, c->shadowq, &c->counters->shadow_count, &c->counters->overflow,
                    c->accum, c->ids, c->sm_count, nullptr, nullptr);
        c->prof_end();
        c->stats.kernel_launches += c->use_bvh ? 3 : 2;  // closest (+ heavy scan), shade
        int rc = ctx_queue_counters(c);
        if (rc != RT_OK) return rc;
        const bool next_traced = !last && trace_level_ahead(c, d + 1);
        rc = ctx_wait_counters(c);
        if (rc != RT_OK) return rc;
        const uint32_t n_total = c->counters_host->shadow_count;  // appended behind the rays of earlier slices / levels
        const uint32_t n_next = last ? 0 : c->counters_host->level_count[d + 1];
        if (d == 0) c->stats.rays_primary += cnt; else c->stats.rays_secondary += cnt;
        c->stats.rays_shadow += n_total - c->shadow_queued;
        c->shadow_queued = n_total;
        if (n_next) {
            rc = process_level(c, d + 1, n_next, next_traced);
            if (rc != RT_OK) return rc;
        } else if (next_traced) {
            c->stats.kernel_launches += 2;  // the traversal queued for a level that turned out empty (it exits at once)
        }
    }
    return RT_OK;
}

int frame_host_levels(rt_context* c) {
    c->shadow_queued = 0;
    const uint64_t batch = c->levels[0].cap;
    for (uint64_t base = 0; base < c->total_paths; base += batch) {
        const uint32_t n = (uint32_t)std::min<uint64_t>(batch, c->total_paths - base);
        CUDA_TRY(cudaMemsetAsync(&c->counters->level_count[0], 0, sizeof(uint32_t), c->stream));
        c->prof_begin(0);
        c->k->primary(c->stream, c->ds, c->fp, c->levels[0], &c->counters->level_count[0], &c->counters->overflow, base, n);
        c->prof_end();
        c->stats.kernel_launches += 1;
        int rc = ctx_queue_counters(c);
        if (rc != RT_OK) return rc;
        const bool traced = trace_level_ahead(c, 0);
        rc = ctx_wait_counters(c);
        if (rc != RT_OK) return rc;
        const uint32_t n0 = c->counters_host->level_count[0];
        c->stats.paths += n0;
        if (n0) {
            rc = process_level(c, 0, n0, traced);
            if (rc != RT_OK) return rc;
        }
    }
    return flush_shadow(c);
}

// ---------------------------------------------------------------------------------------------------------
// device-driven wavefront: every launch of the frame is queued without the host ever reading a count.  Each kernel takes
// its ray count from the device counter its producer filled (stream order) and its grid is sized for an upper bound
// (persistent kernels: the resident grid anyway).  A 4K frame's deep levels hold a few thousand rays; with a host round
// trip per level they cost launch floor + read-back latency each, which does not shrink when the frame is sharded over
// more GPUs (r01: 1 -> 8 GPU efficiency 0.65 on the default workload).
//
// The host cannot know in advance that a level fits its queue (a transparent hit spawns two rays), so the frame is
// OPTIMISTIC: appends past a queue's capacity are dropped and flagged (Counters.overflow), every consumer clamps its
// count to the capacity, and a flagged frame is simply traced again by the host-driven path above, which slices
// levels and cannot overflow.  The context then stays host-driven.
// Returns RT_OK with *overflowed set when the frame has to be redone.
// ---------------------------------------------------------------------------------------------------------
int frame_device_driven(rt_context* c) {
    const uint32_t n_levels = (uint32_t)c->levels.size();
    const uint64_t cap = c->levels[0].cap;
    const uint64_t branch = (uint64_t)std::max(1, c->ds.max_children);
    // primary batches of cap / branch paths: level 1 then fits whatever the materials do; deeper levels fit unless the
    // tree of rays keeps growing (it decays in every scene measured), in which case the frame is redone level by level
    const uint64_t batch = std::max<uint64_t>(cap / branch, 32);
    // what a batch resets: shadow_count, level_count[], lc[] (contiguous in Counters)
    char* batch_state = reinterpret_cast<char*>(&c->counters->shadow_count);
    const size_t batch_state_bytes = sizeof(Counters) - offsetof(Counters, shadow_count);
    size_t slot = 0;
    for (uint64_t base = 0; base < c->total_paths; base += batch, slot++) {
        const uint32_t n = (uint32_t)std::min<uint64_t>(batch, c->total_paths - base);
        if (base) CUDA_TRY(cudaMemsetAsync(batch_state, 0, batch_state_bytes, c->stream));
        c->prof_begin(0);
        c->k->primary(c->stream, c->ds, c->fp, c->levels[0], &c->counters->level_count[0], &c->counters->overflow, base, n);
        c->prof_end();
        c->stats.kernel_launches += 1;
        uint64_t bound = n;
        // The shadow queries of level d are traced on a second stream while the main stream goes on with level d + 1: the two
        // traversal launches have the same resident grid, so the later one's CTAs move in as the earlier one's run out of rays
        // -- the tail of every launch (a few long walks holding the GPU while the other lanes have nothing left to pull) is
        // filled with the other launch's rays.  Each segment [mark[d], mark[d + 1]) of the shadow queue is pinned by a
        // one-thread kernel between the two shade passes that append to it.  Not while profiling (the per-class times are
        // taken launch by launch on one stream) and not for the linear scan (no resident grid, no tail).
        const bool overlap = c->use_bvh && c->stream2 && !c->profiling && !(c->opts.tuning & RT_TUNE_SERIAL_SHADOW);
        for (uint32_t d = 0; d < n_levels; d++) {
            const bool last = d + 1 >= n_levels;
            RayQueue& in = c->levels[d];
            const uint32_t* n_dev = &c->counters->level_count[d];
            c->prof_begin(1);
            c->k->closest(c->stream, c->ds, c->use_bvh, in, 0, (uint32_t)bound, c->counters, c->sm_count, c->heavy, n_dev, &c->counters->lc[d], false);
            c->prof_end();
            c->prof_begin(2);
            c->k->shade(c->stream, c->ds, c->fp, in, 0, (uint32_t)bound, last ? c->null_queue : c->levels[d + 1], &c->counters->level_count[last ? 63 : d + 1],
                        c->shadowq, &c->counters->shadow_count, &c->counters->overflow, c->accum, c->ids, c->sm_count, n_dev,
                        d > 0 ? &c->counters->shade_cursor[d] : nullptr);
            c->prof_end();
            c->stats.kernel_launches += c->use_bvh ? 3 : 2;
            bound = std::min<uint64_t>(cap, bound * branch);
            if (overlap) {
                c->k->mark(c->stream, &c->counters->shadow_count, &c->counters->shadow_mark[d + 1]);
                CUDA_TRY(cudaEventRecord(c->ev_seg[d], c->stream));
                CUDA_TRY(cudaStreamWaitEvent(c->stream2, c->ev_seg[d], 0));
                c->k->shadow(c->stream2, c->ds, true, c->shadowq, c->shadowq.cap, c->accum, c->counters, c->sm_count, c->heavy2,
                             &c->counters->shadow_mark[d + 1], &c->counters->lc[LC_SHADOW_SEG + d], false, &c->counters->shadow_mark[d]);
                c->stats.kernel_launches += 3;
            }
        }
        if (overlap) {
            CUDA_TRY(cudaEventRecord(c->ev_join, c->stream2));
            CUDA_TRY(cudaStreamWaitEvent(c->stream, c->ev_join, 0));
        } else {
            c->prof_begin(3);
            c->k->shadow(c->stream, c->ds, c->use_bvh, c->shadowq, c->shadowq.cap, c->accum, c->counters, c->sm_count, c->heavy, &c->counters->shadow_count,
                         &c->counters->lc[LC_SHADOW], false, nullptr);
            c->prof_end();
            c->stats.kernel_launches += c->use_bvh ? 2 : 1;
        }
        CUDA_TRY(cudaMemcpyAsync(&c->counters_host[slot], c->counters, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
    }
    c->dd_batches = slot;
    return RT_OK;
}

// after the frame's last event: the per-batch records of a device-driven frame -> stats; true when a queue overflowed
bool dd_collect(rt_context* c) {
    const uint32_t n_levels = (uint32_t)c->levels.size();
    for (size_t b = 0; b < c->dd_batches; b++) {
        const Counters& h = c->counters_host[b];
        if (h.overflow) return true;
        c->stats.paths += h.level_count[0];
        c->stats.rays_primary += h.level_count[0];
        for (uint32_t d = 1; d < n_levels; d++) c->stats.rays_secondary += h.level_count[d];
        c->stats.rays_shadow += h.shadow_count;
    }
    return false;
}

int frame_begin(rt_context* c) {
    const size_t npix = (size_t)c->fp.local_rows * c->fp.width;
    c->stats.rays_total = c->stats.rays_primary = c->stats.rays_secondary = c->stats.rays_shadow = 0;
    c->stats.paths = 0;
    CUDA_TRY(cudaMemsetAsync(c->accum, 0, npix * 3 * sizeof(double), c->stream));
    CUDA_TRY(cudaMemsetAsync(c->ids, 0xFF, npix * sizeof(int32_t), c->stream));
    CUDA_TRY(cudaMemsetAsync(c->counters, 0, sizeof(Counters), c->stream));
    return RT_OK;
}

}  // namespace

extern "C" {

void rt_release_cached_memory(void) {
    if (g_drop_cached_multi) g_drop_cached_multi();  // the rt_multi objects kept by rt_render_multi hold queues on every device
    g_cache.trim();
    g_pinned.trim();
}

int rt_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int rt_context_create(const rt_scene* s, const rt_render_opts* o, rt_context** out) {
    if (!s || !o || !out) { set_error("rt_context_create: null argument"); return RT_ERR_INVALID; }
    *out = nullptr;
    if (o->_reserved != 0) { set_error("rt_render_opts._reserved must be 0 (initialise with rt_render_opts_default)"); return RT_ERR_INVALID; }
    if (rt_device_count() <= 0) { set_error("no CUDA device: this library has no CPU fallback"); return RT_ERR_CUDA; }
    rt_context* c = new rt_context();
    c->device = o->device;
    c->opts = *o;
    auto fail = [&](int rc) { delete c; return rc; };
    if (cudaSetDevice(o->device) != cudaSuccess) { set_error("cudaSetDevice failed"); return fail(RT_ERR_CUDA); }
    if (cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess) { set_error("cudaStreamCreate failed"); return fail(RT_ERR_CUDA); }
    c->stream = c->own_stream;
    cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, o->device);
    c->k = o->fp_mode == RT_FP_FAST ? kernels_fast() : kernels_strict();
    bool fresh = false;
    int rc = acquire_device_scene(s, o->accel, o->tuning, o->device, c->stream, &c->scene, &fresh);
    if (rc != RT_OK) return fail(rc);
    c->ds = c->scene->ds;
    c->ds.camera = s->camera;           // by-value header fields that do not bump the scene's revision
    c->ds.bg_kind = s->bg_kind;
    std::memcpy(c->ds.bg_color, s->bg_color, sizeof c->ds.bg_color);
    c->use_bvh = c->scene->use_bvh;
    c->host_levels = (o->tuning & RT_TUNE_HOST_LEVELS) != 0 || c->scene->dd_overflowed.load();

    FrameParams& fp = c->fp;
    if ((rc = make_frame_params(s, o, &fp)) != RT_OK) return fail(rc);
    c->bytewidth = rt_bmp_bytewidth(fp.width);
    c->total_paths = fp.npix_padded * fp.antialias * fp.cam_samples;

    // queue sizing
    const uint32_t branch = (uint32_t)std::max(1, c->ds.max_children);
    const uint32_t lights = std::max(1u, c->ds.n_lights);
    uint64_t cap = o->max_rays_in_flight ? o->max_rays_in_flight : (1ull << 24);
    const uint64_t want = std::max<uint64_t>(c->total_paths * branch, 1024);
    cap = std::min(cap, want);
    cap = std::max<uint64_t>(cap, (uint64_t)branch * 32);
    cap = (cap + 1023) & ~1023ull;
    // Shadow queries only feed the accumulator, so they are not traced level by level: the queue keeps filling across
    // levels and is traced when it could not take the next slice's worst case, or at the end of the frame -- one or two
    // large launches instead of one small one per level (a 4K frame's deep levels hold a few thousand rays each).
    const uint64_t shadow_cap = std::min<uint64_t>(cap * 8, (std::max<uint64_t>(c->total_paths * lights * 3, 1024) + 1023) & ~1023ull);
    c->slice = (uint32_t)std::max<uint64_t>(32, std::min<uint64_t>(cap / branch, shadow_cap / lights));
    const uint32_t n_levels = fp.max_depth + 2;
    c->levels.resize(n_levels);
    for (uint32_t d = 0; d < n_levels; d++)
        if ((rc = alloc_ray_queue(&c->levels[d], (uint32_t)cap, &c->allocs, true)) != RT_OK) return fail(rc);
    if ((rc = alloc_shadow_queue(&c->shadowq, (uint32_t)shadow_cap, &c->allocs)) != RT_OK) return fail(rc);
    std::memset(&c->null_queue, 0, sizeof c->null_queue);

    const size_t npix = (size_t)fp.local_rows * fp.width;
    c->fb_bytes = (size_t)fp.local_rows * c->bytewidth;
    if ((rc = dev_alloc(&c->accum, npix * 3)) != RT_OK) return fail(rc);
    c->allocs.push_back(c->accum);
    if ((rc = dev_alloc(&c->ids, npix)) != RT_OK) return fail(rc);
    c->allocs.push_back(c->ids);
    if ((rc = dev_alloc(&c->framebuffer, c->fb_bytes + 16)) != RT_OK) return fail(rc);
    c->allocs.push_back(c->framebuffer);
    if ((rc = dev_alloc(&c->counters, 1)) != RT_OK) return fail(rc);
    c->allocs.push_back(c->counters);
    if (c->use_bvh) {
        if ((rc = dev_alloc(&c->heavy, HEAVY_WORDS)) != RT_OK) return fail(rc);
        c->allocs.push_back(c->heavy);
        // the split-scan scratch behind the list must start zeroed; its users leave it zeroed (k_heavy_scan)
        if (cudaMemsetAsync(c->heavy + HEAVY_CAP, 0, sizeof(HeavyScratch), c->stream) != cudaSuccess) { set_error("cudaMemsetAsync failed"); return fail(RT_ERR_CUDA); }
        if (!(o->tuning & RT_TUNE_SERIAL_SHADOW)) {  // second stream, its heavy list and the events of frame_device_driven's overlap
            if ((rc = dev_alloc(&c->heavy2, HEAVY_WORDS)) != RT_OK) return fail(rc);
            c->allocs.push_back(c->heavy2);
            if (cudaMemsetAsync(c->heavy2 + HEAVY_CAP, 0, sizeof(HeavyScratch), c->stream) != cudaSuccess) { set_error("cudaMemsetAsync failed"); return fail(RT_ERR_CUDA); }
            if (cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking) != cudaSuccess) { set_error("cudaStreamCreate failed"); return fail(RT_ERR_CUDA); }
            c->ev_seg.reserve(fp.max_depth + 2);
            for (uint32_t d = 0; d < fp.max_depth + 2; d++) {
                cudaEvent_t e = nullptr;
                if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) { set_error("cudaEventCreate failed"); return fail(RT_ERR_CUDA); }
                c->ev_seg.push_back(e);
            }
            if (cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming) != cudaSuccess) { set_error("cudaEventCreate failed"); return fail(RT_ERR_CUDA); }
        }
    }
    const uint64_t dd_batch = std::max<uint64_t>(cap / branch, 32);
    c->final_slot = (size_t)((c->total_paths + dd_batch - 1) / dd_batch);
    if (g_pinned.get(&c->counters_host, &c->n_host_slots, c->final_slot + 1) != cudaSuccess) { set_error("cudaMallocHost failed"); return fail(RT_ERR_CUDA); }
    if (cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->ev_counts, cudaEventDisableTiming) != cudaSuccess) { set_error("cudaEventCreate failed"); return fail(RT_ERR_CUDA); }
    std::memset(&c->stats, 0, sizeof c->stats);
    c->stats.bvh_build_ms = fresh ? c->scene->bvh_ms : 0.0;  // 0: the device copy kept with the scene was reused
    c->stats.n_bvh_nodes = c->ds.n_bvh_nodes;
    c->stats.bvh_node_bytes = !c->use_bvh ? 0 : (c->ds.bvh_qnodes ? sizeof(BvhQNode) : sizeof(BvhNode));
    c->stats.local_rows = fp.local_rows;
    c->stats.bytewidth = c->bytewidth;
    c->stats.accel_used = c->use_bvh ? RT_ACCEL_BVH : RT_ACCEL_LINEAR;
    c->stats.fp_mode_used = o->fp_mode == RT_FP_FAST ? RT_FP_FAST : RT_FP_STRICT;
    *out = c;
    return RT_OK;
}

void rt_context_destroy(rt_context* c) { delete c; }

namespace {
// resolve + the frame-wide counters + the frame's closing event, queued behind whatever traced the frame
int frame_finish(rt_context* c, bool closing_event = true) {
    c->prof_begin(4);
    c->k->resolve(c->stream, c->ds, c->fp, c->accum, c->final_out ? c->final_out : c->framebuffer, c->bytewidth, c->final_out != nullptr);
    c->prof_end();
    c->stats.kernel_launches += 1;
    CUDA_TRY(cudaMemcpyAsync(&c->counters_host[c->final_slot], c->counters, sizeof(Counters), cudaMemcpyDeviceToHost, c->stream));
    if (closing_event) CUDA_TRY(cudaEventRecord(c->ev1, c->stream));
    return RT_OK;
}

// Captures one device-driven frame (frame_begin .. frame_finish) on the context's own streams into an executable graph.
// The timing events stay outside: they are recorded around the graph launch on whatever stream the caller gave.
void frame_graph_capture(rt_context* c) {
    if (c->graph_exec) { cudaGraphExecDestroy(c->graph_exec); c->graph_exec = nullptr; }
    const cudaStream_t user = c->stream;
    c->stream = c->own_stream;
    cudaGraph_t g = nullptr;
    int rc = RT_OK;
    if (cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) rc = RT_ERR_CUDA;
    if (rc == RT_OK) {
        c->stats.kernel_launches = 0;
        rc = frame_begin(c);
        if (rc == RT_OK) rc = frame_device_driven(c);
        if (rc == RT_OK) rc = frame_finish(c, false);
        if (cudaStreamEndCapture(c->stream, &g) != cudaSuccess || !g) rc = RT_ERR_CUDA;
    }
    c->stream = user;
    if (rc == RT_OK && cudaGraphInstantiate(&c->graph_exec, g, 0) != cudaSuccess) { c->graph_exec = nullptr; rc = RT_ERR_CUDA; }
    if (g) cudaGraphDestroy(g);
    if (rc != RT_OK) {
        cudaGetLastError();
        c->graph_failed = true;
        return;
    }
    c->graph_kernel_launches = c->stats.kernel_launches;
}
}  // namespace

int rt_context_render_begin(rt_context* c) {
    if (!c) { set_error("null context"); return RT_ERR_INVALID; }
    if (c->in_flight) { set_error("rt_context_render_begin: the previous frame has not been collected (rt_context_render_end)"); return RT_ERR_INVALID; }
    CUDA_TRY(cudaSetDevice(c->device));
    c->stats.kernel_launches = 0;
    c->prof_class.clear();
    int rc = RT_OK;
    if (!c->host_levels) {
        // every launch of the frame, the resolve included, is queued without a single read-back: the call returns while the
        // GPU works, and the counts (and whether a queue overflowed) are looked at in rt_context_render_end
        const bool graphable = c->use_bvh && c->stream2 && !c->profiling && !c->graph_failed && (c->opts.tuning & RT_TUNE_FRAME_GRAPH);
        uint8_t* out_now = c->final_out ? c->final_out : c->framebuffer;
        const bool same = c->graph_key_valid && out_now == c->graph_out && std::memcmp(&c->ds, &c->graph_ds, sizeof(DeviceScene)) == 0;
        if (graphable && same && !c->graph_exec) frame_graph_capture(c);   // second frame from the same arguments: nothing runs yet
        if (!same && c->graph_exec) { cudaGraphExecDestroy(c->graph_exec); c->graph_exec = nullptr; }
        c->graph_ds = c->ds; c->graph_out = out_now; c->graph_key_valid = true;
        if (graphable && same && c->graph_exec) {
            c->stats.rays_total = c->stats.rays_primary = c->stats.rays_secondary = c->stats.rays_shadow = 0;
            c->stats.paths = 0;
            c->stats.kernel_launches = c->graph_kernel_launches;
            CUDA_TRY(cudaEventRecord(c->ev0, c->stream));
            CUDA_TRY(cudaGraphLaunch(c->graph_exec, c->stream));   // frame_begin .. frame_finish, the closing event apart
            CUDA_TRY(cudaEventRecord(c->ev1, c->stream));
            c->in_flight_dd = true;
            c->in_flight = true;
            return RT_OK;
        }
        c->stats.kernel_launches = 0;
        CUDA_TRY(cudaEventRecord(c->ev0, c->stream));
        if ((rc = frame_begin(c)) != RT_OK) return rc;
        if ((rc = frame_device_driven(c)) != RT_OK) return rc;
        c->in_flight_dd = true;
    } else {
        CUDA_TRY(cudaEventRecord(c->ev0, c->stream));
        if ((rc = frame_begin(c)) != RT_OK) return rc;
        if ((rc = frame_host_levels(c)) != RT_OK) return rc;  // reads counts back level by level: done when it returns
        c->in_flight_dd = false;
    }
    if ((rc = frame_finish(c)) != RT_OK) return rc;
    c->in_flight = true;
    return RT_OK;
}

int rt_context_render_end(rt_context* c, rt_stats* stats) {
    if (!c) { set_error("null context"); return RT_ERR_INVALID; }
    if (!c->in_flight) { set_error("rt_context_render_end: no frame in flight (rt_context_render_begin)"); return RT_ERR_INVALID; }
    CUDA_TRY(cudaSetDevice(c->device));
    c->in_flight = false;
    CUDA_TRY(cudaEventSynchronize(c->ev1));
    if (c->in_flight_dd && dd_collect(c)) {
        // a queue overflowed: this context (and later ones of this scene) go level by level from now on, this frame included
        c->host_levels = true;
        c->scene->dd_overflowed.store(true);
        c->stats.device_driven_overflows += 1;
        int rc = frame_begin(c);
        if (rc == RT_OK) rc = frame_host_levels(c);
        if (rc == RT_OK) rc = frame_finish(c);
        if (rc != RT_OK) return rc;
        CUDA_TRY(cudaEventSynchronize(c->ev1));
    }
    const Counters* fin = &c->counters_host[c->final_slot];
    if (fin->overflow) { set_error("internal error: ray queue or traversal stack overflow"); return RT_ERR_CUDA; }
    for (int k = 0; k < 2; k++) {
        c->stats.bvh_nodes_visited[k] = fin->bvh_nodes[k];
        c->stats.bvh_prims_tested[k] = fin->bvh_prims[k];
        c->stats.heavy_rays[k] = fin->heavy_rays[k];
        c->stats.scan_sphere_tests[k] = fin->scan_tests[k];
    }
    for (int k = 0; k < 5; k++) { c->stats.class_ms[k] = 0.0; c->stats.class_launches[k] = 0; }
    for (size_t i = 0; i < c->prof_class.size(); i++) {
        float pm = 0;
        if (cudaEventElapsedTime(&pm, c->prof_events[2 * i], c->prof_events[2 * i + 1]) == cudaSuccess) {
            c->stats.class_ms[c->prof_class[i]] += pm;
            c->stats.class_launches[c->prof_class[i]] += 1;
        }
    }
    CUDA_TRY(cudaGetLastError());
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->stats.render_ms = ms;
    c->stats.rays_total = c->stats.rays_primary + c->stats.rays_secondary + c->stats.rays_shadow;
    c->stats.host_levels = c->host_levels ? 1u : 0u;
    if (stats) *stats = c->stats;
    return RT_OK;
}

int rt_context_render(rt_context* c, rt_stats* stats) {
    const int rc = rt_context_render_begin(c);
    return rc != RT_OK ? rc : rt_context_render_end(c, stats);
}

int rt_context_stats(const rt_context* c, rt_stats* stats) {
    if (!c || !stats) { set_error("null argument"); return RT_ERR_INVALID; }
    *stats = c->stats;
    return RT_OK;
}

void* rt_context_framebuffer(rt_context* c) { return c ? c->framebuffer : nullptr; }
size_t rt_context_framebuffer_bytes(const rt_context* c) { return c ? c->fb_bytes : 0; }

int rt_context_download(rt_context* c, uint8_t* bgr_rows, int32_t* primary_ids, double* accum) {
    if (!c) { set_error("null context"); return RT_ERR_INVALID; }
    CUDA_TRY(cudaSetDevice(c->device));
    const size_t npix = (size_t)c->fp.local_rows * c->fp.width;
    if (bgr_rows) CUDA_TRY(cudaMemcpyAsync(bgr_rows, c->framebuffer, c->fb_bytes, cudaMemcpyDeviceToHost, c->stream));
    if (primary_ids) CUDA_TRY(cudaMemcpyAsync(primary_ids, c->ids, npix * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    if (accum) CUDA_TRY(cudaMemcpyAsync(accum, c->accum, npix * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_context_untile(rt_context* c, const void* gathered_dev, void* out_dev) {
    if (!c || !gathered_dev || !out_dev) { set_error("null argument"); return RT_ERR_INVALID; }
    CUDA_TRY(cudaSetDevice(c->device));
    launch_untile(c->stream, (const uint32_t*)gathered_dev, (uint32_t*)out_dev, c->fp.height, c->bytewidth / 4, c->fp.band_rows, c->fp.world,
                  c->fp.local_rows);
    CUDA_TRY(cudaGetLastError());
    return RT_OK;
}

int rt_context_set_stream(rt_context* c, void* cuda_stream) {
    if (!c) { set_error("null context"); return RT_ERR_INVALID; }
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
    return RT_OK;
}

int rt_context_set_profiling(rt_context* c, int on) {
    if (!c) { set_error("null context"); return RT_ERR_INVALID; }
    c->profiling = on != 0;
    return RT_OK;
}

int rt_context_frame_create(rt_context* c, void** dev_ptr) {
    if (!c) { set_error("null context"); return RT_ERR_INVALID; }
    CUDA_TRY(cudaSetDevice(c->device));
    if (!c->frame_own) {
        CUDA_TRY(cudaMalloc((void**)&c->frame_own, (size_t)c->bytewidth * c->fp.height + 16));
        CUDA_TRY(cudaMemset(c->frame_own, 0, (size_t)c->bytewidth * c->fp.height + 16));
    }
    c->final_out = c->frame_own;
    if (dev_ptr) *dev_ptr = c->frame_own;
    return RT_OK;
}

int rt_context_frame_export(rt_context* c, uint8_t* handle64) {
    if (!c || !handle64) { set_error("null argument"); return RT_ERR_INVALID; }
    if (!c->frame_own) { set_error("rt_context_frame_export: no frame (call rt_context_frame_create first)"); return RT_ERR_INVALID; }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
    CUDA_TRY(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, c->frame_own));
    std::memcpy(handle64, &h, 64);
    return RT_OK;
}

int rt_context_frame_import(rt_context* c, const uint8_t* handle64) {
    if (!c || !handle64) { set_error("null argument"); return RT_ERR_INVALID; }
    CUDA_TRY(cudaSetDevice(c->device));
    if (c->frame_ipc) { CUDA_TRY(cudaIpcCloseMemHandle(c->frame_ipc)); c->frame_ipc = nullptr; }
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, 64);
    CUDA_TRY(cudaIpcOpenMemHandle(&c->frame_ipc, h, cudaIpcMemLazyEnablePeerAccess));
    c->final_out = (uint8_t*)c->frame_ipc;
    return RT_OK;
}

int rt_context_frame_set(rt_context* c, void* dev_ptr) {
    if (!c) { set_error("null context"); return RT_ERR_INVALID; }
    c->final_out = (uint8_t*)dev_ptr;
    return RT_OK;
}

int rt_context_synchronize(rt_context* c) {
    if (!c) { set_error("null context"); return RT_ERR_INVALID; }
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    return RT_OK;
}

}  // extern "C"

namespace {
int ctx_rebind_scene(rt_context* c, const rt_scene* s, double* built_ms = nullptr);  // below, with the multi-GPU code

// rt_render is one-shot by contract -- a context per call, as upstream's main would have it -- but a caller rendering frame
// after frame would pay for the context's streams, events and ~30 queue blocks every time (0.29 ms of a 16.5 ms frame: create
// 0.16, destroy 0.13).  The last two contexts are kept, keyed by (scene handle, options), and re-bound to the scene's current
// revision on the next call (two: two callers with a scene each keep two frames in flight).  rt_release_cached_memory and
// rt_scene_free drop them.
struct ContextCache {
    struct Entry { const rt_scene* scene; rt_render_opts opts; rt_context* c; };
    std::mutex mu;
    std::vector<Entry> entries;
    rt_context* take(const rt_scene* s, const rt_render_opts& o) {
        std::lock_guard<std::mutex> g(mu);
        for (size_t i = 0; i < entries.size(); i++)
            if (entries[i].scene == s && std::memcmp(&entries[i].opts, &o, sizeof o) == 0) {
                rt_context* c = entries[i].c;
                entries.erase(entries.begin() + i);
                return c;
            }
        return nullptr;
    }
    void put(const rt_scene* s, const rt_render_opts& o, rt_context* c) {
        rt_context* evict = nullptr;
        {
            std::lock_guard<std::mutex> g(mu);
            entries.push_back({s, o, c});
            if (entries.size() > 2) { evict = entries.front().c; entries.erase(entries.begin()); }
        }
        delete evict;
    }
    void drop(const rt_scene* s) {  // s == nullptr: everything
        std::vector<rt_context*> dead;
        {
            std::lock_guard<std::mutex> g(mu);
            for (size_t i = 0; i < entries.size();)
                if (!s || entries[i].scene == s) { dead.push_back(entries[i].c); entries.erase(entries.begin() + i); }
                else i++;
        }
        for (rt_context* c : dead) delete c;
    }
};
ContextCache g_context_cache;
}  // namespace

extern "C" {

int rt_render(const rt_scene* s, const rt_render_opts* o, uint8_t* bgr_rows, int32_t* primary_ids, double* accum, rt_stats* stats) {
    if (!s || !o) { set_error("rt_render: null argument"); return RT_ERR_INVALID; }
    if (o->world != 1) { set_error("rt_render renders whole frames on one GPU: world must be 1 (rt_render_multi shards a frame over GPUs)"); return RT_ERR_INVALID; }
    rt_render_opts key;
    std::memset(&key, 0, sizeof key);  // padding bytes too: the key is compared with memcmp
    key = *o;
    rt_context* c = g_context_cache.take(s, key);
    double built_ms = 0.0;  // a build done for the kept context that then could not serve the scene still happened in this call
    if (c) {
        if (ctx_rebind_scene(c, s, &built_ms) != RT_OK) { delete c; c = nullptr; }  // cannot serve the scene as it is now: make a new one
        else c->stats.device_driven_overflows = 0;                                  // a count per call, as with a fresh context
    }
    int rc = RT_OK;
    if (!c) {
        rc = rt_context_create(s, o, &c);
        if (rc != RT_OK) return rc;
        if (built_ms > 0.0 && c->stats.bvh_build_ms == 0.0) c->stats.bvh_build_ms = built_ms;
    }
    rc = rt_context_render(c, stats);
    if (rc == RT_OK) {
        // world == 1: local rows are the image rows (plus band padding at the end, which is not copied)
        const size_t npix = (size_t)c->fp.height * c->fp.width;
        cudaError_t e = cudaSuccess;
        if (bgr_rows) e = cudaMemcpyAsync(bgr_rows, c->framebuffer, (size_t)c->fp.height * c->bytewidth, cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess && primary_ids) e = cudaMemcpyAsync(primary_ids, c->ids, npix * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess && accum) e = cudaMemcpyAsync(accum, c->accum, npix * 3 * sizeof(double), cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) { set_error(std::string("copy back: ") + cudaGetErrorString(e)); rc = RT_ERR_CUDA; }
    }
    if (rc == RT_OK) g_context_cache.put(s, key, c);
    else {
        const std::string keep = get_error();
        rt_context_destroy(c);
        set_error(keep);
    }
    return rc;
}

}  // extern "C"

// =========================================================================================================
// Multi-GPU frames inside the library: the frame loop of src/main.rs:39-59 with the image sharded by
// interleaved row bands over the GPUs of one node.  One host thread and one stream per device; the scene is replicated
// (each device keeps its copy with the rt_scene); every pixel is independent under the counter RNG, so tracing needs no
// exchange.  What differs between the gather modes is only how the quantised BGR bands become one frame:
//   DIRECT  every device DMAs its own bands straight to their rows of the caller's host buffer (one strided copy per
//           device, all PCIe links in parallel) -- the frame is never assembled on a device;
//   PEER    every device's k_resolve stores its rows into device[0]'s whole-frame buffer over NVLink peer access -- the
//           gather is the resolve kernel's own 128-bit stores, no collective and no de-interleave pass;
//   NCCL    ncclAllGather of the shards (single process, ncclCommInitAll) + k_untile on every device -- every GPU ends up
//           with the whole frame.  NCCL is loaded at run time (dlopen libnccl.so.2): the library has no link-time
//           dependency on it and fails loudly, for this mode only, where it is absent.
// =========================================================================================================
namespace {

typedef struct ncclComm* nccl_comm_t;
struct NcclApi {
    void* so = nullptr;
    int (*CommInitAll)(nccl_comm_t*, int, const int*) = nullptr;
    int (*CommDestroy)(nccl_comm_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*AllGather)(const void*, void*, size_t, int /* ncclDataType_t */, nccl_comm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
    std::string error;
    bool load() {
        if (so) return true;
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            so = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (so) break;
        }
        if (!so) { error = "NCCL gather requested but libnccl.so.2 cannot be loaded"; return false; }
        auto sym = [&](const char* n) { return dlsym(so, n); };
        CommInitAll = (decltype(CommInitAll))sym("ncclCommInitAll");
        CommDestroy = (decltype(CommDestroy))sym("ncclCommDestroy");
        GroupStart = (decltype(GroupStart))sym("ncclGroupStart");
        GroupEnd = (decltype(GroupEnd))sym("ncclGroupEnd");
        AllGather = (decltype(AllGather))sym("ncclAllGather");
        GetErrorString = (decltype(GetErrorString))sym("ncclGetErrorString");
        GetVersion = (decltype(GetVersion))sym("ncclGetVersion");
        if (!CommInitAll || !CommDestroy || !GroupStart || !GroupEnd || !AllGather || !GetErrorString) {
            error = "libnccl.so.2 lacks an expected symbol";
            so = nullptr;
            return false;
        }
        return true;
    }
};
constexpr int NCCL_UINT8 = 1;  // ncclUint8 (nccl.h ncclDataType_t)

// communicators are expensive (ncclCommInitAll: ~1 s) and rt_render_multi is one-shot: keep them per device list
struct NcclCache {
    std::mutex m;
    NcclApi api;
    std::map<std::vector<int>, std::vector<nccl_comm_t>> comms;
    int get(const std::vector<int>& devices, std::vector<nccl_comm_t>* out) {
        std::lock_guard<std::mutex> g(m);
        if (!api.load()) { set_error(api.error); return RT_ERR_UNSUPPORTED; }
        auto it = comms.find(devices);
        if (it == comms.end()) {
            std::vector<nccl_comm_t> v(devices.size(), nullptr);
            const int r = api.CommInitAll(v.data(), (int)devices.size(), devices.data());
            if (r != 0) { set_error(std::string("ncclCommInitAll: ") + api.GetErrorString(r)); return RT_ERR_CUDA; }
            it = comms.emplace(devices, v).first;
        }
        *out = it->second;
        return RT_OK;
    }
};
NcclCache g_nccl;

}  // namespace

struct rt_multi {
    int n = 0;
    std::vector<int> devices;
    int gather = RT_GATHER_DIRECT;
    rt_render_opts opts{};
    const rt_scene* scene = nullptr;
    std::vector<rt_context*> ctx;
    uint32_t width = 0, height = 0, bytewidth = 0, band_rows = 4;
    size_t frame_bytes = 0;
    std::vector<uint8_t*> frames;    // PEER: [0] only; NCCL: one whole frame per device
    std::vector<uint8_t*> gathered;  // NCCL: rank-major shards per device
    std::vector<nccl_comm_t> comms;
    rt_stats last{};
    // workers: one per device, running whatever `job` is posted
    std::vector<std::thread> threads;
    std::mutex m;
    std::condition_variable cv_go, cv_done;
    uint64_t generation = 0;
    int pending = 0;
    bool quit = false;
    std::function<int(int)> job;
    std::vector<int> rc;
    std::vector<std::string> err;
    std::vector<rt_stats> st;

    void worker(int i) {
        uint64_t seen = 0;
        for (;;) {
            std::function<int(int)> j;
            {
                std::unique_lock<std::mutex> lk(m);
                cv_go.wait(lk, [&] { return quit || generation != seen; });
                if (quit) return;
                seen = generation;
                j = job;
            }
            const int r = j(i);
            {
                std::lock_guard<std::mutex> lk(m);
                rc[i] = r;
                if (r != RT_OK) err[i] = get_error();
                if (--pending == 0) cv_done.notify_all();
            }
        }
    }
    // runs f(i) on every device's thread; returns the first failure (its message becomes the caller's rt_last_error)
    int run(std::function<int(int)> f) {
        {
            std::lock_guard<std::mutex> lk(m);
            job = std::move(f);
            pending = n;
            generation++;
        }
        cv_go.notify_all();
        std::unique_lock<std::mutex> lk(m);
        cv_done.wait(lk, [&] { return pending == 0; });
        for (int i = 0; i < n; i++)
            if (rc[i] != RT_OK) { set_error("device " + std::to_string(devices[i]) + ": " + err[i]); return rc[i]; }
        return RT_OK;
    }
    ~rt_multi() {
        if (!threads.empty()) {
            run([&](int i) {
                if (ctx[i]) { rt_context_destroy(ctx[i]); ctx[i] = nullptr; }
                cudaSetDevice(devices[i]);
                if (i < (int)frames.size()) g_cache.release(frames[i]);
                if (i < (int)gathered.size()) g_cache.release(gathered[i]);
                return (int)RT_OK;
            });
            {
                std::lock_guard<std::mutex> lk(m);
                quit = true;
            }
            cv_go.notify_all();
            for (std::thread& t : threads) t.join();
        }
    }
};

namespace {

// Copies the bands a shard holds (local row order, `row_bytes` per row) to their rows of a whole-frame host array: one strided
// DMA for the whole bands plus one for a cut-off last band.
cudaError_t copy_bands_to_host(const FrameParams& fp, const void* local, void* host, size_t row_bytes, cudaStream_t st) {
    const uint32_t band = fp.band_rows;
    const uint32_t n_bands = (fp.height + band - 1) / band;          // bands of the whole frame
    const uint32_t mine = fp.rank < n_bands ? (n_bands - fp.rank + fp.world - 1) / fp.world : 0u;  // ... of this shard
    if (!mine) return cudaSuccess;
    const uint32_t last_rows = fp.height - (n_bands - 1) * band;     // rows of the frame's last band
    const bool have_last = (n_bands - 1) % fp.world == fp.rank && last_rows != band;
    const uint32_t whole = have_last ? mine - 1 : mine;
    const size_t band_bytes = (size_t)band * row_bytes;
    char* dst = (char*)host + (size_t)fp.rank * band_bytes;
    cudaError_t e = cudaSuccess;
    if (whole) e = cudaMemcpy2DAsync(dst, (size_t)fp.world * band_bytes, local, band_bytes, band_bytes, whole, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && have_last)
        e = cudaMemcpyAsync(dst + (size_t)whole * fp.world * band_bytes, (const char*)local + (size_t)whole * band_bytes, (size_t)last_rows * row_bytes,
                            cudaMemcpyDeviceToHost, st);
    return e;
}

}  // namespace

extern "C" {

int rt_multi_create(const rt_scene* s, const rt_render_opts* o, int n_devices, const int* devices, int gather, rt_multi** out) {
    if (!s || !o || !out || n_devices < 1) { set_error("rt_multi_create: bad argument"); return RT_ERR_INVALID; }
    *out = nullptr;
    const int have = rt_device_count();
    if (have <= 0) { set_error("no CUDA device: this library has no CPU fallback"); return RT_ERR_CUDA; }
    if (gather < RT_GATHER_AUTO || gather > RT_GATHER_NCCL) { set_error("rt_multi_create: unknown gather mode"); return RT_ERR_INVALID; }
    std::unique_ptr<rt_multi> m(new rt_multi());
    m->n = n_devices;
    for (int i = 0; i < n_devices; i++) {
        const int d = devices ? devices[i] : i;
        if (d < 0 || d >= have) { set_error("rt_multi_create: device ordinal out of range"); return RT_ERR_INVALID; }
        // DIRECT / PEER: an ordinal may appear more than once (its shards then share that GPU -- how the sharding and the
        // gather are exercised on a one-GPU machine); NCCL wants one rank per GPU
        if (gather == RT_GATHER_NCCL)
            for (int k : m->devices)
                if (k == d) { set_error("rt_multi_create: RT_GATHER_NCCL needs distinct devices"); return RT_ERR_INVALID; }
        m->devices.push_back(d);
    }
    m->gather = gather == RT_GATHER_AUTO ? RT_GATHER_DIRECT : gather;
    m->opts = *o;
    m->scene = s;
    m->ctx.assign(n_devices, nullptr);
    m->rc.assign(n_devices, RT_OK);
    m->err.assign(n_devices, std::string());
    m->st.assign(n_devices, rt_stats{});
    FrameParams fp{};
    rt_render_opts probe = *o;
    probe.rank = 0; probe.world = n_devices;
    int rc = make_frame_params(s, &probe, &fp);
    if (rc != RT_OK) return rc;
    m->width = fp.width; m->height = fp.height; m->band_rows = fp.band_rows;
    m->bytewidth = rt_bmp_bytewidth(fp.width);
    m->frame_bytes = (size_t)m->bytewidth * fp.height;
    if (m->gather == RT_GATHER_NCCL && n_devices > 1) {
        if ((rc = g_nccl.get(m->devices, &m->comms)) != RT_OK) return rc;
    }
    if (m->gather == RT_GATHER_PEER) m->frames.assign(1, nullptr);
    if (m->gather == RT_GATHER_NCCL) { m->frames.assign(n_devices, nullptr); m->gathered.assign(n_devices, nullptr); }
    for (int i = 0; i < n_devices; i++) m->threads.emplace_back([p = m.get(), i] { p->worker(i); });
    rt_multi* M = m.get();
    // device 0 first: in PEER mode the others store into its frame buffer
    if (M->gather == RT_GATHER_PEER) {
        if (cudaSetDevice(M->devices[0]) != cudaSuccess || g_cache.alloc((void**)&M->frames[0], M->frame_bytes + 16) != cudaSuccess) {
            set_error("rt_multi_create: cannot allocate the frame on the first device");
            return RT_ERR_CUDA;
        }
    }
    rc = M->run([M](int i) {
        if (cudaSetDevice(M->devices[i]) != cudaSuccess) { set_error("cudaSetDevice failed"); return (int)RT_ERR_CUDA; }
        rt_render_opts oi = M->opts;
        oi.device = M->devices[i]; oi.rank = i; oi.world = M->n;
        int r = rt_context_create(M->scene, &oi, &M->ctx[i]);
        if (r != RT_OK) return r;
        if (M->gather == RT_GATHER_PEER) {
            if (M->devices[i] != M->devices[0]) {
                int can = 0;
                cudaDeviceCanAccessPeer(&can, M->devices[i], M->devices[0]);
                if (!can) { set_error("no peer access to device " + std::to_string(M->devices[0]) + " (RT_GATHER_PEER needs NVLink/PCIe P2P)"); return (int)RT_ERR_UNSUPPORTED; }
                const cudaError_t e = cudaDeviceEnablePeerAccess(M->devices[0], 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { set_error(std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e)); return (int)RT_ERR_CUDA; }
                cudaGetLastError();
            }
            M->ctx[i]->final_out = M->frames[0];
        } else if (M->gather == RT_GATHER_NCCL) {
            if (g_cache.alloc((void**)&M->frames[i], M->frame_bytes + 16) != cudaSuccess ||
                g_cache.alloc((void**)&M->gathered[i], (size_t)M->n * M->ctx[i]->fb_bytes + 16) != cudaSuccess) {
                set_error("rt_multi_create: cannot allocate the gather buffers");
                return (int)RT_ERR_CUDA;
            }
        }
        return (int)RT_OK;
    });
    if (rc != RT_OK) return rc;  // ~rt_multi destroys what was created
    *out = m.release();
    return RT_OK;
}

void rt_multi_destroy(rt_multi* m) { delete m; }

int rt_multi_gather_used(const rt_multi* m) { return m ? m->gather : RT_ERR_INVALID; }

int rt_multi_render(rt_multi* m, uint8_t* bgr_rows, int32_t* primary_ids, rt_stats* stats) {
    if (!m) { set_error("null rt_multi"); return RT_ERR_INVALID; }
    const auto t0 = std::chrono::steady_clock::now();
    int rc = m->run([m, bgr_rows, primary_ids](int i) {
        rt_context* c = m->ctx[i];
        int r = rt_context_render(c, &m->st[i]);  // returns with the device idle: peer stores of k_resolve have landed
        if (r != RT_OK) return r;
        cudaError_t e = cudaSuccess;
        if (bgr_rows && m->gather == RT_GATHER_DIRECT) e = copy_bands_to_host(c->fp, c->framebuffer, bgr_rows, m->bytewidth, c->stream);
        if (e == cudaSuccess && primary_ids) e = copy_bands_to_host(c->fp, c->ids, primary_ids, (size_t)c->fp.width * sizeof(int32_t), c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) { set_error(std::string("copy back: ") + cudaGetErrorString(e)); return (int)RT_ERR_CUDA; }
        return (int)RT_OK;
    });
    if (rc != RT_OK) return rc;
    if (m->gather == RT_GATHER_NCCL) {
        if (m->n > 1) {
            NcclApi& N = g_nccl.api;
            int r = N.GroupStart();
            for (int i = 0; i < m->n && r == 0; i++)
                r = N.AllGather(m->ctx[i]->framebuffer, m->gathered[i], m->ctx[i]->fb_bytes, NCCL_UINT8, m->comms[i], m->ctx[i]->stream);
            const int r2 = N.GroupEnd();
            if (r == 0) r = r2;
            if (r != 0) { set_error(std::string("ncclAllGather: ") + N.GetErrorString(r)); return RT_ERR_CUDA; }
        }
        for (int i = 0; i < m->n; i++) {
            rt_context* c = m->ctx[i];
            CUDA_TRY(cudaSetDevice(m->devices[i]));
            if (m->n > 1) {
                launch_untile(c->stream, (const uint32_t*)m->gathered[i], (uint32_t*)m->frames[i], c->fp.height, m->bytewidth / 4, c->fp.band_rows, c->fp.world,
                              c->fp.local_rows);
            } else {
                CUDA_TRY(cudaMemcpyAsync(m->frames[i], c->framebuffer, m->frame_bytes, cudaMemcpyDeviceToDevice, c->stream));
            }
        }
        for (int i = 0; i < m->n; i++) {
            CUDA_TRY(cudaSetDevice(m->devices[i]));
            CUDA_TRY(cudaStreamSynchronize(m->ctx[i]->stream));
        }
    }
    if (bgr_rows && m->gather != RT_GATHER_DIRECT) {
        CUDA_TRY(cudaSetDevice(m->devices[0]));
        CUDA_TRY(cudaMemcpyAsync(bgr_rows, m->frames[0], m->frame_bytes, cudaMemcpyDeviceToHost, m->ctx[0]->stream));
        CUDA_TRY(cudaStreamSynchronize(m->ctx[0]->stream));
    }
    rt_stats sum = m->st[0];
    for (int i = 1; i < m->n; i++) {
        const rt_stats& a = m->st[i];
        sum.rays_total += a.rays_total; sum.rays_primary += a.rays_primary; sum.rays_secondary += a.rays_secondary; sum.rays_shadow += a.rays_shadow;
        sum.paths += a.paths; sum.kernel_launches += a.kernel_launches;
        sum.render_ms = std::max(sum.render_ms, a.render_ms);
        sum.bvh_build_ms = std::max(sum.bvh_build_ms, a.bvh_build_ms);
        sum.local_rows += a.local_rows;
        for (int k = 0; k < 2; k++) {
            sum.bvh_nodes_visited[k] += a.bvh_nodes_visited[k]; sum.bvh_prims_tested[k] += a.bvh_prims_tested[k]; sum.heavy_rays[k] += a.heavy_rays[k];
            sum.scan_sphere_tests[k] += a.scan_sphere_tests[k];
        }
        for (int k = 0; k < 5; k++) { sum.class_ms[k] = std::max(sum.class_ms[k], a.class_ms[k]); sum.class_launches[k] += a.class_launches[k]; }
        sum.host_levels |= a.host_levels; sum.device_driven_overflows += a.device_driven_overflows;
    }
    if (m->gather == RT_GATHER_NCCL && m->n > 1) sum.kernel_launches += (uint64_t)m->n;  // k_untile per device
    (void)t0;
    m->last = sum;
    if (stats) *stats = sum;
    return RT_OK;
}

int rt_multi_device_stats(const rt_multi* m, int index, rt_stats* stats) {
    if (!m || !stats || index < 0 || index >= m->n) { set_error("rt_multi_device_stats: bad argument"); return RT_ERR_INVALID; }
    *stats = m->st[index];
    return RT_OK;
}

const void* rt_multi_frame_device(const rt_multi* m, int index) {
    if (!m || index < 0 || index >= (int)m->frames.size()) return nullptr;
    return m->frames[index];
}

}  // extern "C"

namespace {

// A context made for an earlier frame of the same scene handle is brought up to date: the device copy of the scene's
// current revision (uploaded and built now if the scene changed since), camera and background.  Returns RT_ERR_UNSUPPORTED
// when the context cannot serve the scene as it is now (frame size, branching factor or light count differ from what its
// queues were sized for): the caller makes a new one.
// built_ms: receives the BVH build time when this call had to upload and build the scene -- also when it then finds that the
// context cannot serve the scene (the device copy stays with the scene: the context made in its place will find it ready)
int ctx_rebind_scene(rt_context* c, const rt_scene* s, double* built_ms) {
    CUDA_TRY(cudaSetDevice(c->device));
    FrameParams fp{};
    int rc = make_frame_params(s, &c->opts, &fp);
    if (rc != RT_OK) return rc;
    if (std::memcmp(&fp, &c->fp, sizeof fp) != 0) { set_error("frame parameters changed"); return RT_ERR_UNSUPPORTED; }
    std::shared_ptr<DeviceSceneHolder> h;
    bool fresh = false;
    rc = acquire_device_scene(s, c->opts.accel, c->opts.tuning, c->device, c->stream, &h, &fresh);
    if (rc != RT_OK) return rc;
    if (built_ms && fresh) *built_ms = h->bvh_ms;
    if (h->use_bvh != c->use_bvh || h->ds.max_children != c->ds.max_children || h->ds.n_lights != c->ds.n_lights) {
        set_error("scene structure changed");
        return RT_ERR_UNSUPPORTED;
    }
    c->scene = h;
    uint8_t* keep_out = c->final_out;
    c->ds = h->ds;
    c->ds.camera = s->camera;
    c->ds.bg_kind = s->bg_kind;
    std::memcpy(c->ds.bg_color, s->bg_color, sizeof c->ds.bg_color);
    c->final_out = keep_out;
    c->host_levels = (c->opts.tuning & RT_TUNE_HOST_LEVELS) != 0 || h->dd_overflowed.load();
    c->stats.bvh_build_ms = fresh ? h->bvh_ms : 0.0;
    c->stats.n_bvh_nodes = c->ds.n_bvh_nodes;
    c->stats.bvh_node_bytes = !c->use_bvh ? 0 : (c->ds.bvh_qnodes ? sizeof(BvhQNode) : sizeof(BvhNode));
    return RT_OK;
}

// rt_render_multi is one-shot by contract, but a caller rendering frame after frame (the reference's main in a loop) would
// pay for eight threads, eight contexts and their streams / events per frame (0.9 ms of a 5 ms frame at four GPUs).  The
// last few rt_multi objects are kept, keyed by (scene handle, options, devices, gather), and re-bound to the scene's current
// revision on the next call.  rt_release_cached_memory and rt_scene_free drop them.
struct MultiCache {
    struct Entry {
        const rt_scene* scene;
        rt_render_opts opts;
        std::vector<int> devices;
        int gather;
        rt_multi* m;
    };
    std::mutex mu;
    std::vector<Entry> entries;
    static bool same_opts(const rt_render_opts& a, const rt_render_opts& b) {
        rt_render_opts x = a, y = b;
        x.device = y.device = 0; x.rank = y.rank = 0; x.world = y.world = 1;  // set per device by rt_multi
        return std::memcmp(&x, &y, sizeof x) == 0;
    }
    rt_multi* take(const rt_scene* s, const rt_render_opts& o, const std::vector<int>& devs, int gather) {
        std::lock_guard<std::mutex> g(mu);
        for (size_t i = 0; i < entries.size(); i++)
            if (entries[i].scene == s && entries[i].gather == gather && entries[i].devices == devs && same_opts(entries[i].opts, o)) {
                rt_multi* m = entries[i].m;
                entries.erase(entries.begin() + i);
                return m;
            }
        return nullptr;
    }
    void put(const rt_scene* s, const rt_render_opts& o, const std::vector<int>& devs, int gather, rt_multi* m) {
        rt_multi* evict = nullptr;
        {
            std::lock_guard<std::mutex> g(mu);
            entries.push_back({s, o, devs, gather, m});
            if (entries.size() > 2) { evict = entries.front().m; entries.erase(entries.begin()); }
        }
        delete evict;
    }
    void drop(const rt_scene* s) {  // s == nullptr: everything
        std::vector<rt_multi*> dead;
        {
            std::lock_guard<std::mutex> g(mu);
            for (size_t i = 0; i < entries.size();)
                if (!s || entries[i].scene == s) { dead.push_back(entries[i].m); entries.erase(entries.begin() + i); }
                else i++;
        }
        for (rt_multi* m : dead) delete m;
    }
};
MultiCache g_multi_cache;
struct MultiCacheHook {
    MultiCacheHook() {
        rt::g_scene_free_hook = [](const rt_scene* s) { g_multi_cache.drop(s); g_context_cache.drop(s); };
        g_drop_cached_multi = [] { g_multi_cache.drop(nullptr); g_context_cache.drop(nullptr); };
    }
} g_multi_cache_hook;

}  // namespace

extern "C" {

int rt_render_multi(const rt_scene* s, const rt_render_opts* o, int n_devices, const int* devices, int gather, uint8_t* bgr_rows, int32_t* primary_ids,
                    rt_stats* stats) {
    if (!s || !o || n_devices < 1) { set_error("rt_render_multi: bad argument"); return RT_ERR_INVALID; }
    std::vector<int> devs;
    for (int i = 0; i < n_devices; i++) devs.push_back(devices ? devices[i] : i);
    const int mode = gather == RT_GATHER_AUTO ? RT_GATHER_DIRECT : gather;
    rt_multi* m = g_multi_cache.take(s, *o, devs, mode);
    if (m) {
        m->scene = s;
        const int rc = m->run([m](int i) { return ctx_rebind_scene(m->ctx[i], m->scene); });
        if (rc != RT_OK) { delete m; m = nullptr; }  // cannot serve the scene as it is now: make a new one
    }
    if (!m) {
        const int rc = rt_multi_create(s, o, n_devices, devices, gather, &m);
        if (rc != RT_OK) return rc;
    }
    const int rc = rt_multi_render(m, bgr_rows, primary_ids, stats);
    if (rc != RT_OK) {
        const std::string keep = get_error();
        delete m;
        set_error(keep);
        return rc;
    }
    g_multi_cache.put(s, *o, devs, mode, m);
    return RT_OK;
}

// ---------------------------------------------------------------------------------------------------------
// test hooks: batch Scene::intersect, Camera::project, Color::write_bgr
// ---------------------------------------------------------------------------------------------------------
int rt_trace_rays(const rt_scene* s, const rt_render_opts* o, uint64_t n, const double* origin_dir, int32_t* ids, double* t, double* normal) {
    if (!s || !o || (n && (!origin_dir || !ids))) { set_error("rt_trace_rays: null argument"); return RT_ERR_INVALID; }
    if (rt_device_count() <= 0) { set_error("no CUDA device: this library has no CPU fallback"); return RT_ERR_CUDA; }
    if (n == 0) return RT_OK;
    if (n > 0x7FFFFFFFull) { set_error("too many rays"); return RT_ERR_INVALID; }
    CUDA_TRY(cudaSetDevice(o->device));
    cudaStream_t st;
    CUDA_TRY(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    int sm_count = 148;
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, o->device);
    int rc = RT_OK;
    {
        DeviceSceneHolder holder;
        holder.device = o->device;
        uint32_t* heavy = nullptr;
        const rt_scene* use = s;  // scenes used only for intersection tests may lack camera/background
        rt_scene patched;
        if (!s->camera_set || !s->background_set) {
            patched = *s;
            patched.camera_set = true; patched.background_set = true; patched.bg_kind = BG_SOLID;
            use = &patched;
        }
        rc = holder.init(use, o->accel, o->tuning & SCENE_TUNING, st);
        if (rc == RT_OK && holder.use_bvh) rc = dev_alloc(&heavy, HEAVY_WORDS);
        std::vector<void*> allocs;
        RayQueue q{};
        double *d_t = nullptr, *d_n = nullptr;
        Counters* cursor = nullptr;
        if (rc == RT_OK) rc = alloc_ray_queue(&q, (uint32_t)n, &allocs, false);
        if (rc == RT_OK) rc = dev_alloc(&d_t, n);
        if (rc == RT_OK) rc = dev_alloc(&d_n, 3 * n);
        if (rc == RT_OK) rc = dev_alloc(&cursor, 1);
        if (rc == RT_OK) {
            std::vector<double> soa(6 * n);
            for (uint64_t i = 0; i < n; i++)
                for (int k = 0; k < 6; k++) soa[(size_t)k * n + i] = origin_dir[6 * i + k];
            cudaError_t e = cudaMemcpyAsync(q.ox, soa.data(), sizeof(double) * 6 * n, cudaMemcpyHostToDevice, st);
            if (e == cudaSuccess) e = cudaMemsetAsync(cursor, 0, sizeof(Counters), st);
            if (e == cudaSuccess && heavy) e = cudaMemsetAsync(heavy + HEAVY_CAP, 0, sizeof(HeavyScratch), st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
            const KernelTable* k = o->fp_mode == RT_FP_FAST ? kernels_fast() : kernels_strict();
            if (e == cudaSuccess) {
                k->closest(st, holder.ds, holder.use_bvh, q, 0, (uint32_t)n, cursor, sm_count, heavy, nullptr, &cursor->lc[LC_SHARED], false);
                k->hit_normals(st, holder.ds, q, (uint32_t)n, d_t, d_n);
                e = cudaMemcpyAsync(ids, q.hit_id, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, st);
            }
            if (e == cudaSuccess && t) e = cudaMemcpyAsync(t, d_t, sizeof(double) * n, cudaMemcpyDeviceToHost, st);
            if (e == cudaSuccess && normal) e = cudaMemcpyAsync(normal, d_n, sizeof(double) * 3 * n, cudaMemcpyDeviceToHost, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
            if (e == cudaSuccess) e = cudaGetLastError();
            if (e != cudaSuccess) { set_error(std::string("rt_trace_rays: ") + cudaGetErrorString(e)); rc = RT_ERR_CUDA; }
        }
        for (void* p : allocs) g_cache.release(p);
        g_cache.release(d_t); g_cache.release(d_n); g_cache.release(cursor); g_cache.release(heavy);
    }
    cudaStreamDestroy(st);
    return rc;
}

int rt_project(const rt_scene* s, const rt_render_opts* o, uint64_t n, const double* pos_xy, double* origin_dir) {
    if (!s || !o || (n && (!pos_xy || !origin_dir))) { set_error("rt_project: null argument"); return RT_ERR_INVALID; }
    if (!s->camera_set) { set_error("rt_project: scene has no camera"); return RT_ERR_INVALID; }
    if (rt_device_count() <= 0) { set_error("no CUDA device: this library has no CPU fallback"); return RT_ERR_CUDA; }
    if (n == 0) return RT_OK;
    CUDA_TRY(cudaSetDevice(o->device));
    DeviceScene ds{};
    ds.camera = s->camera;
    double *d_pos = nullptr, *d_od = nullptr;
    CUDA_TRY(cudaMalloc((void**)&d_pos, sizeof(double) * 2 * n));
    cudaError_t e = cudaMalloc((void**)&d_od, sizeof(double) * 6 * n);
    if (e == cudaSuccess) e = cudaMemcpy(d_pos, pos_xy, sizeof(double) * 2 * n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        const KernelTable* k = o->fp_mode == RT_FP_FAST ? kernels_fast() : kernels_strict();
        k->project(0, ds, o->seed, (uint32_t)n, d_pos, d_od);
        e = cudaMemcpy(origin_dir, d_od, sizeof(double) * 6 * n, cudaMemcpyDeviceToHost);
    }
    cudaFree(d_pos); cudaFree(d_od);
    if (e != cudaSuccess) { set_error(std::string("rt_project: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
    return RT_OK;
}

int rt_quantize_bgr(const double* rgb, uint64_t n, uint8_t* bgr, int device) {
    if (n && (!rgb || !bgr)) { set_error("rt_quantize_bgr: null argument"); return RT_ERR_INVALID; }
    if (rt_device_count() <= 0) { set_error("no CUDA device: this library has no CPU fallback"); return RT_ERR_CUDA; }
    if (n == 0) return RT_OK;
    CUDA_TRY(cudaSetDevice(device));
    double *d_rgb = nullptr, *d_avg = nullptr;
    uint8_t* d_bgr = nullptr;
    CUDA_TRY(cudaMalloc((void**)&d_rgb, sizeof(double) * 3 * n));
    cudaError_t e = cudaMalloc((void**)&d_bgr, 3 * n);
    if (e == cudaSuccess) e = cudaMalloc((void**)&d_avg, sizeof(double) * 255);
    if (e == cudaSuccess) e = cudaMemcpy(d_rgb, rgb, sizeof(double) * 3 * n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d_avg, SRGB_AVERAGE, sizeof(double) * 255, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        kernels_strict()->quantize(0, d_avg, d_rgb, n, d_bgr);
        e = cudaMemcpy(bgr, d_bgr, 3 * n, cudaMemcpyDeviceToHost);
    }
    cudaFree(d_rgb); cudaFree(d_bgr); cudaFree(d_avg);
    if (e != cudaSuccess) { set_error(std::string("rt_quantize_bgr: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
    return RT_OK;
}

}  // extern "C"
