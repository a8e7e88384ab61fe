// kernels.cu -- the wavefront ray-queue kernels (sm_100a).  Compiled twice, see rt_device.cuh / build.py:
//   -DRT_NS=strict -fmad=false   and   -DRT_NS=fast -fmad=true
//
// Kernel            replaces (reference file:line)
//   k_primary       src/main.rs:50-53 + src/camera.rs:77-79,110-121      one thread per path, coalesced SoA ray writes
//   k_closest_scan  src/scene.rs:247-249 + src/shapes.rs:50-112          sphere records staged through shared memory
//                                                                        with 1-D TMA bulk copies (cp.async.bulk + mbarrier)
//   k_closest_bvh   same, through the device-built LBVH                  persistent warps pulling rays off a cursor
//   k_shade         src/raytrace.rs:30-256, src/scene.rs:117-155,        material switch on a tag; queue slots reserved once per
//                   src/texture.rs:39-58, src/color.rs ops               CTA, shared out by __ballot_sync/__popc ranks
//   k_shadow_*      src/raytrace.rs:43-50 (and siblings)                 any-hit with range, adds the pending contribution
//   k_resolve       src/main.rs:56 + src/color.rs:593-600,628-632        threshold-table search, BGR pack, 128-bit stores
#include <cuda_runtime.h>

#include <type_traits>

#include "kernels.h"
#include "rt_device.cuh"

namespace rt {
namespace RT_NS {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_TILE = 1024;  // spheres per shared-memory tile (32 KB), double buffered
#define RT_PREFETCH_L2(p) asm volatile("prefetch.global.L2 [%0];" ::"l"(p))
#ifndef RT_RAY_PREFETCH
#define RT_RAY_PREFETCH 8192  // k_closest_bvh refills prefetch the queue records this many rays ahead of the cursor (0: off): 5.01 -> 4.97 ms (10k),
                              // 5.24 -> 5.19 ms (1M), the same at 2048 and 32768; the shadow kernel (reserved blocks) did not gain: profiles/r02_v5_ray_prefetch_ab.txt
#endif
#ifndef RT_BVH_THREADS
#define RT_BVH_THREADS 128  // with RT_BVH_MIN_BLOCKS: 64 x 16 is 0.7-1 % faster on a 1/8 shard (CTAs leave sooner at the end of a launch), 256 x 4
#endif                      // 1-2 % slower; whole frames within noise (profiles/r02_v8_bvh_threads_shard_ab.txt)
constexpr int BVH_THREADS = RT_BVH_THREADS;
#ifndef RT_SHADE_THREADS
#define RT_SHADE_THREADS 128
#endif
constexpr int SHADE_THREADS = RT_SHADE_THREADS;

// ---------------------------------------------------------------------------------------------------------
// small PTX wrappers: mbarrier + 1-D TMA bulk copy (global -> shared)
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
                 "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!ok);
}

// ---------------------------------------------------------------------------------------------------------
// queue append with one atomicAdd per CTA (returning atomics on one counter queue up in its L2 slice: see k_shade).
// Returns the slot for threads with want == true (0xFFFFFFFF when the queue is full -> overflow flag).  Must be
// called by all THREADS threads of the block, once.
// ---------------------------------------------------------------------------------------------------------
template <int THREADS>
__device__ __forceinline__ uint32_t block_append(bool want, uint32_t* counter, uint32_t cap, uint32_t* overflow) {
    __shared__ uint32_t w_count[THREADS / 32];
    __shared__ uint32_t base_s;
    const unsigned m = __ballot_sync(0xFFFFFFFFu, want);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) w_count[warp] = (uint32_t)__popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t total = 0;
#pragma unroll
        for (int k = 0; k < THREADS / 32; k++) total += w_count[k];
        base_s = total ? atomicAdd(counter, total) : 0u;
    }
    __syncthreads();
    uint32_t slot = base_s + (uint32_t)__popc(m & ((1u << lane) - 1u));
#pragma unroll
    for (int k = 0; k < THREADS / 32; k++)
        if (k < warp) slot += w_count[k];
    if (want && slot >= cap) {
        *overflow = 1u;
        return 0xFFFFFFFFu;
    }
    return want ? slot : 0xFFFFFFFFu;
}

__device__ __forceinline__ void accum_add(double* accum, uint32_t pixel, C3 c) {
    if (c.r == 0.0 && c.g == 0.0 && c.b == 0.0) return;  // x + 0 == x; NaN is != 0 and still added
    double* p = accum + (size_t)pixel * 3;
    atomicAdd(p, c.r);
    atomicAdd(p + 1, c.g);
    atomicAdd(p + 2, c.b);
}

// ---------------------------------------------------------------------------------------------------------
// k_primary
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_primary(const __grid_constant__ DeviceScene sc, const __grid_constant__ FrameParams fp, RayQueue q,
                                                 uint32_t* q_count, uint32_t* overflow, uint64_t path_base, uint32_t n_paths) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = i < n_paths;
    uint32_t x = 0, y = 0, lp = 0, aa = 0, cs = 0;
    if (valid) {
        const uint64_t g = path_base + i;
        const uint64_t plane = g / fp.npix_padded;
        const uint64_t slot = g - plane * fp.npix_padded;
        aa = (uint32_t)(plane / fp.cam_samples);
        cs = (uint32_t)(plane - (uint64_t)aa * fp.cam_samples);
        valid = slot_to_pixel(fp, slot, x, y, lp);
    }
    V3 o = v3(0, 0, 0), d = v3(0, 0, 0);
    uint64_t ck = 0;
    if (valid) {
        const uint64_t pixel = (uint64_t)y * fp.width + x;
        const uint64_t sk = key_sample(fp.seed, pixel, aa);
        double px, py;
        pixel_position(fp, x, y, sk, px, py);
        ck = child_key(sk, CODE_CAMERA0 + cs);
        camera_project(sc.camera, px, py, ck, o, d);
    }
    const uint32_t s = block_append<256>(valid, q_count, q.cap, overflow);
    if (s != 0xFFFFFFFFu) {
        q.ox[s] = o.x; q.oy[s] = o.y; q.oz[s] = o.z;
        q.dx[s] = d.x; q.dy[s] = d.y; q.dz[s] = d.z;
        q.tr[s] = 1.0; q.tg[s] = 1.0; q.tb[s] = 1.0;
        q.sig[s] = 1.0;
        q.key[s] = ck;
        q.pixel[s] = lp;
        q.meta[s] = (aa == 0 && cs == 0) ? META_RECORD_ID : 0u;
    }
}

// ---------------------------------------------------------------------------------------------------------
// linear scan with shared-memory staging.  All lanes test the same sphere (shared-memory broadcast); the
// sphere buffer streams through two 32 KB tiles filled by 1-D TMA bulk copies that overlap the arithmetic.
// ---------------------------------------------------------------------------------------------------------
struct ScanSmem {
    SphereRec tile[2][SCAN_TILE];
    uint64_t bar[2];
};

// Returns the number of exact sphere tests this thread made (a shadow query stops at its first occluder): the
// DFMA-equivalent work the scan's roofline fraction is computed from (bench.py).
template <bool SHADOW>
__device__ __forceinline__ uint32_t scan_spheres(const DeviceScene& sc, ScanSmem* sm, const RayQ& r, bool active, Closest& best, double range2, Shadow& sh) {
    const uint32_t n = sc.n_spheres;
    uint32_t tested = 0;
    if (n == 0) return 0;
    const uint32_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (threadIdx.x == 0) {
        mbar_init(&sm->bar[0], 1);
        mbar_init(&sm->bar[1], 1);
        mbar_fence_init();
        const uint32_t cnt = n < (uint32_t)SCAN_TILE ? n : (uint32_t)SCAN_TILE;
        mbar_expect_tx(&sm->bar[0], cnt * (uint32_t)sizeof(SphereRec));
        tma_load_1d(sm->tile[0], sc.sph, cnt * (uint32_t)sizeof(SphereRec), &sm->bar[0]);
    }
    __syncthreads();
    bool done = !active;
    for (uint32_t k = 0; k < ntiles; k++) {
        const uint32_t buf = k & 1u;
        if (threadIdx.x == 0 && k + 1 < ntiles) {
            const uint32_t first = (k + 1) * SCAN_TILE;
            const uint32_t cnt = (n - first) < (uint32_t)SCAN_TILE ? (n - first) : (uint32_t)SCAN_TILE;
            mbar_expect_tx(&sm->bar[buf ^ 1u], cnt * (uint32_t)sizeof(SphereRec));
            tma_load_1d(sm->tile[buf ^ 1u], sc.sph + first, cnt * (uint32_t)sizeof(SphereRec), &sm->bar[buf ^ 1u]);
        }
        mbar_wait(&sm->bar[buf], (k >> 1) & 1u);
        const uint32_t first = k * SCAN_TILE;
        const uint32_t cnt = (n - first) < (uint32_t)SCAN_TILE ? (n - first) : (uint32_t)SCAN_TILE;
        if (!done) {
            const SphereRec* t = sm->tile[buf];
            tested += cnt;
#pragma unroll 4
            for (uint32_t j = 0; j < cnt; j++) {
                const double2 c01 = *reinterpret_cast<const double2*>(&t[j].cx);
                const double2 c23 = *reinterpret_cast<const double2*>(&t[j].cz);
                double th;
                if (sphere_hit(r, c01.x, c01.y, c23.x, c23.y, th)) {
                    if (SHADOW) {
                        if (shadow_hit_counts(th, range2)) { sh.any = true; done = true; tested -= cnt - (j + 1u); break; }
                    } else {
                        closest_update(best, th, __ldg(sc.sph_obj + first + j));
                    }
                }
            }
        }
        // all reads of tile[buf] are complete before it is refilled two iterations later
        const int all_done = __syncthreads_and(done ? 1 : 0);
        if (SHADOW && all_done) {
            if (k + 1 < ntiles) mbar_wait(&sm->bar[buf ^ 1u], ((k + 1) >> 1) & 1u);  // drain the in-flight copy before exit
            break;
        }
    }
    return tested;
}
__device__ __forceinline__ void flush_scan_tests(uint32_t tested, Counters* counters, int cls) {
    unsigned long long t = tested;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xFFFFFFFFu, t, o);
    if ((threadIdx.x & 31) == 0 && t) atomicAdd(&counters->scan_tests[cls], t);
}

// n_dev (both scan kernels): the ray count is read on the device and the grid was sized for an upper bound -- CTAs past
// the end leave before they touch a barrier
__global__ void __launch_bounds__(SCAN_THREADS) k_closest_scan(const __grid_constant__ DeviceScene sc, RayQueue q, uint32_t first, uint32_t n,
                                                               const uint32_t* n_dev, Counters* counters) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    ScanSmem* sm = reinterpret_cast<ScanSmem*>(smem_raw);
    if (n_dev) { n = *n_dev < q.cap ? *n_dev : q.cap; if (blockIdx.x * SCAN_THREADS >= n) return; }
    const uint32_t li = blockIdx.x * SCAN_THREADS + threadIdx.x;
    const bool active = li < n;
    const uint32_t i = first + (active ? li : 0);
    RayQ r = make_rayq(v3(q.ox[i], q.oy[i], q.oz[i]), v3(q.dx[i], q.dy[i], q.dz[i]));
    Closest best = closest_none();
    planes_closest(sc, r, best);
    Shadow dummy; dummy.nan = false; dummy.any = false;
    const uint32_t tested = scan_spheres<false>(sc, sm, r, active && !best.nan, best, -1.0, dummy);
    if (active) {
        q.hit_id[i] = closest_id(best);
        q.hit_t[i] = best.t;
    }
    flush_scan_tests(tested, counters, 0);
}

__global__ void __launch_bounds__(SCAN_THREADS) k_shadow_scan(const __grid_constant__ DeviceScene sc, ShadowQueue sq, uint32_t n, double* accum,
                                                              const uint32_t* n_dev, Counters* counters) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    ScanSmem* sm = reinterpret_cast<ScanSmem*>(smem_raw);
    if (n_dev) { n = *n_dev < sq.cap ? *n_dev : sq.cap; if (blockIdx.x * SCAN_THREADS >= n) return; }
    const uint32_t li = blockIdx.x * SCAN_THREADS + threadIdx.x;
    const bool active = li < n;
    const uint32_t i = active ? li : 0;
    RayQ r = make_rayq(v3(sq.ox[i], sq.oy[i], sq.oz[i]), v3(sq.dx[i], sq.dy[i], sq.dz[i]));
    const double range2 = sq.r2[i];
    Shadow s; s.nan = false; s.any = false;
    planes_shadow(sc, r, range2, s);
    Closest dummy = closest_none();
    const uint32_t tested = scan_spheres<true>(sc, sm, r, active && !s.nan && !s.any, dummy, range2, s);
    if (active && !shadow_result(s, range2)) accum_add(accum, sq.pixel[i], c3(sq.cr[i], sq.cg[i], sq.cb[i]));
    flush_scan_tests(tested, counters, 1);
}

// ---------------------------------------------------------------------------------------------------------
// BVH traversal: persistent warps, each pulling 32 rays at a time off a global cursor
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void flush_traverse_count(const TraverseCount& tc, Counters* counters, int cls) {
    unsigned long long nodes = tc.nodes, prims = tc.prims;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        nodes += __shfl_xor_sync(0xFFFFFFFFu, nodes, o);
        prims += __shfl_xor_sync(0xFFFFFFFFu, prims, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&counters->bvh_nodes[cls], nodes);
        atomicAdd(&counters->bvh_prims[cls], prims);
    }
}

// Warp-aggregated refill: every idle lane takes the next ray index.  The warp reserves a block of `chunk` consecutive
// rays with one atomicAdd and hands them to its idle lanes over the next few refills; only the refill that finds the
// block short waits for the round trip to the cursor (one per ~24 rays was 5.5 % of k_shadow_bvh's stall samples).
// Blocks are aligned powers of two, so `next` alone describes what is left (up to the next multiple of chunk;
// FETCH_NONE = nothing).  Big launches take bigger blocks; small ones stay at 32 so that no warp sits on rays the others
// are waiting to finish.  Must be called by all 32 lanes.  Returns the ray index for lanes that were idle (may be >= n).
// k_closest_bvh keeps the plain form (one atomicAdd per refill): its launches are short, a warp sitting on a reserved block
// at the end of one costs more than the waits (measured: 5.1 -> 5.3 ms per frame with blocks), and it has no register to spare.
__device__ __forceinline__ uint32_t warp_fetch_each(unsigned idle_mask, uint32_t* cursor) {
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(idle_mask) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(cursor, (uint32_t)__popc(idle_mask));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    return base + (uint32_t)__popc(idle_mask & ((1u << lane) - 1u));
}
constexpr uint32_t FETCH_NONE = 0xFFFFFFFFu;
__device__ __forceinline__ uint32_t fetch_chunk(uint32_t n) { return n >= (1u << 24) ? 128u : (n >= (1u << 22) ? 64u : 32u); }
__device__ __forceinline__ uint32_t warp_fetch(unsigned idle_mask, uint32_t* cursor, uint32_t& next, uint32_t chunk) {
    const int lane = threadIdx.x & 31;
    const uint32_t need = (uint32_t)__popc(idle_mask);
    const uint32_t rank = (uint32_t)__popc(idle_mask & ((1u << lane) - 1u));
    const uint32_t avail = next == FETCH_NONE ? 0u : chunk - (next & (chunk - 1u));
    uint32_t idx = next + rank;
    if (avail > need) {
        next += need;
    } else if (avail == need) {
        next = FETCH_NONE;
    } else {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(cursor, chunk);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (rank >= avail) idx = base + (rank - avail);
        next = need - avail == chunk ? FETCH_NONE : base + (need - avail);
    }
    return idx;
}

#ifndef RT_BVH_MIN_BLOCKS
#define RT_BVH_MIN_BLOCKS 8  // 64 registers; 9 and 10 CTAs per SM (56 / 48 registers) spill in the walk: measured slower
#endif
#ifndef RT_FETCH_MIN_IDLE
#define RT_FETCH_MIN_IDLE 16
#endif
#ifndef RT_DESCEND_STEPS
#define RT_DESCEND_STEPS 16  // walk at most this many nodes before looking at the warp again (sweep: profiles/r01_v5_sweep.md)
#endif
constexpr int FETCH_MIN_IDLE = RT_FETCH_MIN_IDLE;  // refill when at least this many lanes of the warp have finished their ray
// Per kernel (profiles/r02_v9_per_kernel_sweep_ab.txt, r02_v9_closest_refill_*.txt): the closest-hit kernel refills a little later
// (its refill is the costlier interruption: plane tests in f64, the ordered walk's state) -- 8 / 12 / 16 / 20 / 24 / 28 idle lanes:
// 5.34 / 5.11 / 4.96 / 4.90 / 4.90 / 4.97 ms per 10k frame, a 1/8 shard 2.43 against 2.47 ms; walk lengths 12 / 24 for it and
// 6 / 12 wide visits for the shadow kernel are worse on the 10k frame and within noise at 1M.
#ifndef RT_FETCH_MIN_IDLE_CLOSEST
#define RT_FETCH_MIN_IDLE_CLOSEST 20
#endif
#ifndef RT_DESCEND_STEPS_CLOSEST
#define RT_DESCEND_STEPS_CLOSEST RT_DESCEND_STEPS
#endif
#ifndef RT_SHADOW_WIDE_STEPS
#define RT_SHADOW_WIDE_STEPS (RT_DESCEND_STEPS / 2)
#endif

// Persistent "while-while" traversal.  Each lane owns one ray at a time; the warp alternates between
//   phase 1  every lane walks internal nodes (f32 box filter) until its next job is a leaf test or it is done,
//   phase 2  every lane with a pending leaf runs the exact f64 sphere test,
// and lanes whose ray has finished are refilled from the global cursor instead of idling until the slowest lane of
// a fixed 32-ray batch is done.
//
// Registers: the walk needs (cur, sp, bound, the twelve plane constants); the f64 ray, the running best hit / shadow
// range and t0 are only needed at a leaf or at the end, so they are parked in shared memory (one column per thread, SoA:
// conflict-free) -- 64 registers per thread, 8 CTAs of 128 threads per SM.
struct TravSmem {
    double ray[6][BVH_THREADS];  // origin, direction
    double aux[BVH_THREADS];     // closest: t of the best hit so far; shadow: squared range
};                               // 7 KB + 1 KB the runtime reserves per CTA: 8 CTAs fit the 64 KB carve-out, 192 KB stay L1
__device__ __forceinline__ void park_ray(TravSmem& sm, V3 o, V3 d, double aux) {
    const int t = threadIdx.x;
    sm.ray[0][t] = o.x; sm.ray[1][t] = o.y; sm.ray[2][t] = o.z;
    sm.ray[3][t] = d.x; sm.ray[4][t] = d.y; sm.ray[5][t] = d.z;
    sm.aux[t] = aux;
}
__device__ __forceinline__ RayQ parked_ray(const TravSmem& sm) {
    const int t = threadIdx.x;
    return make_rayq(v3(sm.ray[0][t], sm.ray[1][t], sm.ray[2][t]), v3(sm.ray[3][t], sm.ray[4][t], sm.ray[5][t]));
}

template <bool QNODES>
__global__ void __launch_bounds__(BVH_THREADS, RT_BVH_MIN_BLOCKS) k_closest_bvh(const __grid_constant__ DeviceScene sc, RayQueue q, uint32_t first, uint32_t n, Counters* counters,
                                                                uint32_t* heavy, uint32_t heavy_walks, const uint32_t* n_dev, uint32_t lc_index) {
    __shared__ TravSmem sm;
    // launched ahead of the host's count read-back (context.cu).  The producer's counter runs past the capacity when its
    // queue overflows (the appends are dropped and the frame fails): never index beyond what the queue holds.
    // The count lives in shared memory: it is only looked at when lanes are refilled, and the walk has no register to spare.
    __shared__ uint32_t n_s;
    if (threadIdx.x == 0) n_s = n_dev ? (*n_dev < q.cap ? *n_dev : q.cap) : n;
    __syncthreads();
    uint32_t walks = 0;  // bounded walks spent on the current ray
    TraverseCount tc = traverse_count_zero();
    TravStack st; int sp = 0;
    BoxRay32 br = boxray32_none();
    int32_t cur = TRAV_EMPTY;
    int32_t best_obj = 0x7FFFFFFF;  // the best hit's t is parked in sm.aux
    bool best_nan = false;
    float bound = 0.0f;
    float t0f = 0.0f;  // t0 of the advanced origin rounded down: a later bound (t - t0f) can only be looser than (t - t0)
    uint32_t ray = 0;
    bool active = false, exhausted = false;
    while (true) {
        const unsigned idle = __ballot_sync(0xFFFFFFFFu, !active);
        if (idle != 0u && !exhausted && (__popc(idle) >= RT_FETCH_MIN_IDLE_CLOSEST || idle == 0xFFFFFFFFu)) {
            const uint32_t li = warp_fetch_each(idle, &counters->lc[lc_index].cursor);
            const uint32_t n_now = n_s;
            if (!active) {
#if RT_RAY_PREFETCH
                if (li + RT_RAY_PREFETCH < n_now) {  // the rays some warp will take RT_RAY_PREFETCH refills from now: HBM -> L2
                    const uint32_t pf = first + li + RT_RAY_PREFETCH;
                    RT_PREFETCH_L2(q.ox + pf); RT_PREFETCH_L2(q.oy + pf); RT_PREFETCH_L2(q.oz + pf);
                    RT_PREFETCH_L2(q.dx + pf); RT_PREFETCH_L2(q.dy + pf); RT_PREFETCH_L2(q.dz + pf);
                }
#endif
                if (li < n_now) {
                    ray = first + li;
                    RT_CHECK(ray < q.cap);
                    const RayQ r = make_rayq(v3(q.ox[ray], q.oy[ray], q.oz[ray]), v3(q.dx[ray], q.dy[ray], q.dz[ray]));
                    Closest best = closest_none();
                    planes_closest(sc, r, best);
                    walks = 0;
                    sp = 0;
                    cur = TRAV_EMPTY;
                    t0f = 0.0f;
                    if (!best.nan && sc.n_spheres > 0) {
                        br = make_boxray32_t<QNODES>(sc, r.o, r.d);
                        t0f = f64_to_f32_down(br.t0);
                        if (br.enter) { cur = 0; bound = closest_bound32(best.t, br.t0); }
                    }
                    best_obj = best.obj; best_nan = best.nan;
                    park_ray(sm, r.o, r.d, best.t);
                    active = true;
                }
            }
            // the cursor only grows: once any lane is handed an index past the end, the queue is drained
            exhausted = __any_sync(0xFFFFFFFFu, li >= n_now && (idle >> (threadIdx.x & 31) & 1u));
        }
        if (__all_sync(0xFFFFFFFFu, !active)) {
            if (exhausted) break;
            continue;
        }
        if (active) {
            cur = trav_descend<QNODES>(sc, br, cur, bound, st, sp, tc, RT_DESCEND_STEPS_CLOSEST);
        }
        __syncwarp();
        if (active) {
            if (cur == TRAV_EMPTY) {
                q.hit_id[ray] = best_obj == 0x7FFFFFFF ? -1 : best_obj;
                q.hit_t[ray] = sm.aux[threadIdx.x];
                active = false;
            } else if (cur < 0) {
                const RayQ r = parked_ray(sm);
                Closest best; best.t = sm.aux[threadIdx.x]; best.obj = best_obj; best.nan = best_nan;
                leaf_closest(sc, r, ~cur, best, tc);
                sm.aux[threadIdx.x] = best.t; best_obj = best.obj; best_nan = best.nan;
                bound = closest_bound32(best.t, (double)t0f);
                cur = trav_pop(st, sp, bound);
            }
            if (active && ++walks > heavy_walks) {
                // runaway traversal: hand the ray to the block-wide scan (k_heavy_scan), which redoes it from scratch
                const uint32_t slot = atomicAdd(&counters->lc[lc_index].heavy, 1u);
                if (slot < HEAVY_CAP) { heavy[slot] = ray; active = false; }
                else walks = 0;  // list full: keep traversing
            }
        }
    }
    flush_traverse_count(tc, counters, 0);
}

// WIDE: the walk goes over sc.bvh_wnodes, four boxes per visit (rt_device.cuh trav_descend_wide); everything else is the
// same kernel.
template <bool QNODES, bool WIDE = false>
__global__ void __launch_bounds__(BVH_THREADS, RT_BVH_MIN_BLOCKS) k_shadow_bvh(const __grid_constant__ DeviceScene sc, ShadowQueue sq, uint32_t n, double* accum, Counters* counters,
                                                               uint32_t* heavy, uint32_t heavy_walks, const uint32_t* n_dev, uint32_t lc_index,
                                                               const uint32_t* first_dev) {
    __shared__ TravSmem sm;
    __shared__ uint32_t n_s, first_s;  // see k_closest_bvh; first_s: where this launch's segment of the queue starts
    if (threadIdx.x == 0) {
        const uint32_t end = n_dev ? (*n_dev < sq.cap ? *n_dev : sq.cap) : n;
        const uint32_t first = first_dev ? (*first_dev < end ? *first_dev : end) : 0u;
        first_s = first;
        n_s = end - first;
    }
    __syncthreads();
    uint32_t range_next = FETCH_NONE;
    uint32_t walks = 0;
    TraverseCount tc = traverse_count_zero();
#ifdef RT_SHADOW_STACK64
    TravStack st; int sp = 0;
#else
    typename std::conditional<WIDE, TravStackWide, TravStackFixed>::type st; int sp = 0;
#endif
    BoxRay32 br = boxray32_none();
    int32_t cur = TRAV_EMPTY;
    float bound = 0.0f;
    uint32_t ray = 0;
    bool active = false, exhausted = false, occluded = false;
    while (true) {
        const unsigned idle = __ballot_sync(0xFFFFFFFFu, !active);
        if (idle != 0u && !exhausted && (__popc(idle) >= FETCH_MIN_IDLE || idle == 0xFFFFFFFFu)) {
            const uint32_t n_now = n_s;
            const uint32_t li = warp_fetch(idle, &counters->lc[lc_index].cursor, range_next, fetch_chunk(n_now));
            if (!active) {
                if (li < n_now) {
                    ray = first_s + li;
                    RT_CHECK(ray < sq.cap);
                    const RayQ r = make_rayq(v3(sq.ox[ray], sq.oy[ray], sq.oz[ray]), v3(sq.dx[ray], sq.dy[ray], sq.dz[ray]));
                    const double range2 = sq.r2[ray];
                    Shadow s; s.nan = false; s.any = false;
                    planes_shadow(sc, r, range2, s);
                    walks = 0;
                    sp = 0;
                    cur = TRAV_EMPTY;
                    occluded = shadow_result(s, range2);
                    if (!s.nan && !s.any && sc.n_spheres > 0) {
                        br = make_boxray32_t<QNODES>(sc, r.o, r.d);
                        if (br.enter) { cur = 0; bound = shadow_bound32(range2, br.t0); }
                    }
                    park_ray(sm, r.o, r.d, range2);
                    active = true;
                }
            }
            exhausted = __any_sync(0xFFFFFFFFu, li >= n_now && (idle >> (threadIdx.x & 31) & 1u));
        }
        if (__all_sync(0xFFFFFFFFu, !active)) {
            if (exhausted) break;
            continue;
        }
        if (active) {
            if constexpr (WIDE) cur = trav_descend_wide(sc, br, cur, bound, st, sp, tc, RT_SHADOW_WIDE_STEPS);
            else cur = trav_descend<QNODES>(sc, br, cur, bound, st, sp, tc, RT_DESCEND_STEPS);
        }
        __syncwarp();
        if (active) {
            if (cur != TRAV_EMPTY && cur < 0) {
                if (leaf_shadow(sc, parked_ray(sm), ~cur, sm.aux[threadIdx.x], tc)) { occluded = true; cur = TRAV_EMPTY; }
                else cur = trav_pop(st, sp, bound);
            }
            if (cur == TRAV_EMPTY) {
                if (!occluded) accum_add(accum, sq.pixel[ray], c3(sq.cr[ray], sq.cg[ray], sq.cb[ray]));
                active = false;
            } else if (++walks > heavy_walks) {
                const uint32_t slot = atomicAdd(&counters->lc[lc_index].heavy, 1u);
                if (slot < HEAVY_CAP) { heavy[slot] = ray; active = false; }
                else walks = 0;
            }
        }
    }
    flush_traverse_count(tc, counters, 1);
}

// ---------------------------------------------------------------------------------------------------------
// k_heavy_scan: the rays a traversal kernel gave up on, one CTA per ray.  A ray that starts 10^7+ units from the scene
// (a ground-plane hit near the horizon) makes the exact test's own rounding error large, so the conservative filter must
// pad every box by many units and the walk touches a large part of the tree: milliseconds in ONE lane.  Here the 256
// threads of a CTA stride over the sphere buffer with the scan's exact test and merge -- the same answer as the scan by
// construction, in microseconds.
// ---------------------------------------------------------------------------------------------------------
constexpr int HEAVY_THREADS = 256;

__device__ __forceinline__ bool closest_before(const Closest& a, const Closest& b) {  // ordering of closest_update
    if (a.nan != b.nan) return a.nan;
    if (a.nan) return a.obj < b.obj;
    return a.t < b.t || (a.t == b.t && a.obj < b.obj);
}

// exact tests of spheres [k0, k1) by the 256 threads of the CTA; SHADOW: returns "an occluder was found" (CTA-wide),
// else the CTA's best hit in `best` of thread 0
template <bool SHADOW>
__device__ __forceinline__ bool heavy_scan_range(const DeviceScene& sc, const RayQ& r, double range2, uint32_t k0, uint32_t k1, Closest& best) {
    __shared__ double s_t[HEAVY_THREADS / 32];
    __shared__ int32_t s_obj[HEAVY_THREADS / 32];
    __shared__ int s_nan[HEAVY_THREADS / 32];
    if (SHADOW) {
        bool any = false;
        for (uint32_t k = k0 + threadIdx.x; k < k1 && !any; k += HEAVY_THREADS) {
            const double2 c01 = *reinterpret_cast<const double2*>(&sc.sph[k].cx), c23 = *reinterpret_cast<const double2*>(&sc.sph[k].cz);
            double th;
            if (sphere_hit(r, c01.x, c01.y, c23.x, c23.y, th) && shadow_hit_counts(th, range2)) any = true;
        }
        return __syncthreads_or(any ? 1 : 0) != 0;
    }
    for (uint32_t k = k0 + threadIdx.x; k < k1; k += HEAVY_THREADS) {
        const double2 c01 = *reinterpret_cast<const double2*>(&sc.sph[k].cx), c23 = *reinterpret_cast<const double2*>(&sc.sph[k].cz);
        double th;
        if (sphere_hit(r, c01.x, c01.y, c23.x, c23.y, th)) closest_update(best, th, __ldg(sc.sph_obj + k));
    }
#pragma unroll
    for (int x = 16; x > 0; x >>= 1) {
        Closest other;
        other.t = __shfl_xor_sync(0xFFFFFFFFu, best.t, x);
        other.obj = __shfl_xor_sync(0xFFFFFFFFu, best.obj, x);
        other.nan = __shfl_xor_sync(0xFFFFFFFFu, best.nan ? 1 : 0, x) != 0;
        if (closest_before(other, best)) best = other;
    }
    if ((threadIdx.x & 31) == 0) { s_t[threadIdx.x >> 5] = best.t; s_obj[threadIdx.x >> 5] = best.obj; s_nan[threadIdx.x >> 5] = best.nan ? 1 : 0; }
    __syncthreads();
    if (threadIdx.x == 0)
        for (int w = 1; w < HEAVY_THREADS / 32; w++) {
            Closest other; other.t = s_t[w]; other.obj = s_obj[w]; other.nan = s_nan[w] != 0;
            if (closest_before(other, best)) best = other;
        }
    __syncthreads();
    return false;
}

// Two shapes.  Many heavy rays: one CTA per ray, CTAs loop over the list.  A few (<= HEAVY_SPLIT_RAYS, the usual case: a
// handful of horizon rays per frame): the sphere buffer of each ray is cut into G = grid / count slices, one CTA per
// slice, and the last CTA to finish a ray (arrival counter behind a __threadfence) merges the slices -- a 1M-sphere scan is
// then microseconds instead of 0.5 ms on one SM.  The merge is a minimum under closest_update's total order (or an OR), so
// it does not depend on which CTA comes last.
template <bool SHADOW>
__global__ void __launch_bounds__(HEAVY_THREADS) k_heavy_scan(const __grid_constant__ DeviceScene sc, RayQueue q, ShadowQueue sq, uint32_t first,
                                                              double* accum, Counters* counters, uint32_t* heavy, uint32_t lc_index) {
    const uint32_t count = counters->lc[lc_index].heavy < HEAVY_CAP ? counters->lc[lc_index].heavy : HEAVY_CAP;
    if (count == 0) return;
    HeavyScratch* hs = reinterpret_cast<HeavyScratch*>(heavy + HEAVY_CAP);
    uint32_t G = 1;
    if (count <= HEAVY_SPLIT_RAYS && gridDim.x <= HEAVY_SPLIT_ITEMS) G = gridDim.x / count;
    if (G < 1) G = 1;
    const uint32_t items = count * G;
    for (uint32_t item = blockIdx.x; item < items; item += gridDim.x) {  // G > 1: items <= gridDim.x, one pass
        const uint32_t e = item / G, j = item - e * G;
        const uint32_t chunk = (sc.n_spheres + G - 1) / G;
        const uint32_t k0 = j * chunk < sc.n_spheres ? j * chunk : sc.n_spheres;
        const uint32_t k1 = k0 + chunk < sc.n_spheres ? k0 + chunk : sc.n_spheres;
        const uint32_t i = heavy[e];  // closest: already the absolute queue index (first + local)
        RT_CHECK(SHADOW ? i < sq.cap : i < q.cap);
        RT_CHECK(item < HEAVY_SPLIT_ITEMS || G == 1);
        const V3 o = SHADOW ? v3(sq.ox[i], sq.oy[i], sq.oz[i]) : v3(q.ox[i], q.oy[i], q.oz[i]);
        const V3 d = SHADOW ? v3(sq.dx[i], sq.dy[i], sq.dz[i]) : v3(q.dx[i], q.dy[i], q.dz[i]);
        const RayQ r = make_rayq(o, d);
        const double range2 = SHADOW ? sq.r2[i] : -1.0;
        Closest best = closest_none();
        bool any = heavy_scan_range<SHADOW>(sc, r, range2, k0, k1, best);
        if (threadIdx.x != 0) continue;
        if (G > 1) {
            if (SHADOW) {
                if (any) atomicOr(&hs->flag[e], 1u);
            } else {
                hs->t[item] = best.t; hs->obj[item] = best.obj; hs->nan[item] = best.nan ? 1 : 0;
            }
            __threadfence();
            if (atomicAdd(&hs->done[e], 1u) != G - 1) continue;  // not the last slice of this ray
            __threadfence();
            if (SHADOW) {
                any = atomicExch(&hs->flag[e], 0u) != 0u;
            } else {
                for (uint32_t w = e * G; w < e * G + G; w++) {
                    Closest other;
                    other.t = __ldcg(&hs->t[w]); other.obj = __ldcg(&hs->obj[w]); other.nan = __ldcg(&hs->nan[w]) != 0;
                    if (closest_before(other, best)) best = other;
                }
            }
            hs->done[e] = 0u;
        }
        if (SHADOW) {
            Shadow s; s.nan = false; s.any = false;
            planes_shadow(sc, r, range2, s);   // the traversal kernel handed the ray over only if no plane decided it; same result
            s.any = s.any || any;
            if (!shadow_result(s, range2)) accum_add(accum, sq.pixel[i], c3(sq.cr[i], sq.cg[i], sq.cb[i]));
            atomicAdd(&counters->heavy_rays[1], 1ull);
        } else {
            // a NaN plane hit wins whatever the spheres say (src/scene.rs:227-233); spheres never produce NaN
            Closest pl = closest_none();
            planes_closest(sc, r, pl);
            if (closest_before(pl, best)) best = pl;
            q.hit_id[i] = closest_id(best);
            q.hit_t[i] = best.t;
            atomicAdd(&counters->heavy_rays[0], 1ull);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// k_shade
// ---------------------------------------------------------------------------------------------------------
// 4 CTAs per SM = 128 registers (52 bytes spilled; the kernel wants 176): 16 warps instead of 12 to cover the HBM latency of
// the ray record and the dependent gathers.  Measured once the queue reservations had stopped queueing in L2 (before that
// more warps only meant more contention: 7.75 ms vs 6.15): 3 / 4 / 5 CTAs = 5.20 / 4.23 / 4.53 ms per 10k frame, 4.00 / 3.28 /
// 3.52 ms at 1M.
#ifndef RT_SHADE_MIN_BLOCKS
#define RT_SHADE_MIN_BLOCKS 4
#endif
#ifndef RT_SHADE_PREFETCH
#define RT_SHADE_PREFETCH 1  // LOOP: prefetch the next tile's ray record into L2 while this tile emits its rays.  Measured
#endif                       // (profiles/r02_v5_shade_prefetch_ab.txt): shade class 4.57 -> 4.23 ms (10k), 3.55 -> 3.36 ms (1M); into L1
                             // instead, or with level 0 looping too: the same times (r02_v5_shade_prefetch2_ab.txt)
#ifndef RT_SHADE_WAVES
#define RT_SHADE_WAVES 1  // LOOP grid = SMs x resident CTAs x this (tiles come off a cursor: the resident grid is enough)
#endif
// n_dev: the count is read on the device (clamped to the queue's capacity), so the launch can be queued before the host
// knows it.  Two shapes of the same body:
//   LOOP = false  one tile of SHADE_THREADS hits per CTA; with n_dev the grid was sized for an upper bound and CTAs past
//                 the end leave at once -- right when the bound is tight (level 0: the bound is the path count);
//   LOOP = true   a resident grid pulls tiles off a cursor -- for the deeper levels of a device-driven frame, whose bound
//                 (capacity of the queue) can be hundreds of times the actual count.  Tiles are handed out in order, one
//                 atomic each, so the rays a level appends keep the order of the level that spawned them, as with one CTA
//                 per tile (blocks start in index order); a strided loop scrambles it, and the next level's traversal and
//                 gathers pay for the lost coherence.
template <bool LOOP>
__global__ void __launch_bounds__(SHADE_THREADS, RT_SHADE_MIN_BLOCKS) k_shade(const __grid_constant__ DeviceScene sc, const __grid_constant__ FrameParams fp, RayQueue in,
                                                         uint32_t first, uint32_t n, RayQueue out, uint32_t* out_count, ShadowQueue sq,
                                                         uint32_t* sq_count, uint32_t* overflow, double* accum, int32_t* ids, const uint32_t* n_dev, uint32_t* tile_cursor) {
    constexpr int SHADE_WARPS = SHADE_THREADS / 32;
    // double-buffered by tile parity: a warp that is one tile ahead writes the other half (it cannot get two ahead: both
    // barriers of the tile in between need every warp)
    __shared__ uint32_t w_shadow[2][SHADE_WARPS], w_child[2][SHADE_WARPS], cta_base[2][2];
    __shared__ uint32_t n_s;  // LOOP: the count stays out of the registers (128 of the 176 this kernel wants)
    __shared__ uint32_t tile_s[2];
    if (LOOP) {
        if (threadIdx.x == 0) {
            n_s = n_dev ? (*n_dev < in.cap ? *n_dev : in.cap) : n;
            tile_s[0] = atomicAdd(tile_cursor, 1u);
        }
        __syncthreads();
    }
    const unsigned FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned below = (1u << lane) - 1u;
    uint32_t par = 0;
    while (true) {
        const uint32_t tile = LOOP ? tile_s[par] : blockIdx.x;
        const uint32_t nn = LOOP ? n_s : (n_dev ? (*n_dev < in.cap ? *n_dev : in.cap) : n);
        if ((uint64_t)tile * SHADE_THREADS >= nn) break;
        if (LOOP && threadIdx.x == 0) tile_s[par ^ 1u] = atomicAdd(tile_cursor, 1u);  // next tile; published by this tile's barriers
        const uint32_t li = tile * SHADE_THREADS + threadIdx.x;
        const bool active = li < nn;
        ShadeCtx s;
        s.alive = false; s.lit = false;
        uint32_t pixel = 0, depth = 0;
        if (active) {
            const uint32_t i = first + li;
            RT_CHECK(i < in.cap);
            const V3 o = v3(in.ox[i], in.oy[i], in.oz[i]);
            const V3 d = v3(in.dx[i], in.dy[i], in.dz[i]);
            const C3 T = c3(in.tr[i], in.tg[i], in.tb[i]);
            const uint32_t meta = in.meta[i];
            const int32_t hid = in.hit_id[i];
            pixel = in.pixel[i];
            RT_CHECK(pixel < fp.local_rows * fp.width);
            depth = meta & META_DEPTH_MASK;
            if ((meta & META_RECORD_ID) && ids) ids[pixel] = hid;
            const C3 add = shade_begin(sc, fp, o, d, T, in.sig[i], in.key[i], depth, hid, in.hit_t[i], s);
            accum_add(accum, pixel, add);
        }
        // Queue slots for everything this tile will emit are reserved up front with ONE atomic per queue.  (One per warp was
        // a third of this kernel's stall samples: 2 x 260k returning atomics per 4K level on the same line of Counters queue
        // up in one L2 slice; ncu source view, profiles/r01_v17_ncu_k_shade_stalls.md.)  The warps leave their counts in
        // shared memory, thread 0 reserves for all of them, and every warp takes its share in warp order.
        // Every lit hit casts a shadow ray at every light (src/scene.rs:117-155 tests occlusion only), so the count is known.
        const bool lit = s.alive && s.lit;
        const unsigned lit_m = __ballot_sync(FULL, lit);
        const uint32_t n_lit = (uint32_t)__popc(lit_m);
        uint32_t n_child = 0;
        for (uint32_t k = 0; k < (uint32_t)sc.max_children; k++) n_child += (uint32_t)__popc(__ballot_sync(FULL, shade_child_wanted(s, k)));
        if (lane == 0) { w_shadow[par][warp] = n_lit * sc.n_lights; w_child[par][warp] = n_child; }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t ts = 0, tc = 0;
#pragma unroll
            for (int k = 0; k < SHADE_WARPS; k++) { ts += w_shadow[par][k]; tc += w_child[par][k]; }
            uint32_t bs = 0, bc = 0;
            if (ts) bs = atomicAdd(sq_count, ts);
            if (tc) bc = atomicAdd(out_count, tc);
            cta_base[par][0] = bs; cta_base[par][1] = bc;
        }
        __syncthreads();
#if RT_SHADE_PREFETCH
        if (LOOP) {  // the next tile's record is on its way from HBM while this tile's rays are being emitted
            const uint64_t nx = (uint64_t)tile_s[par ^ 1u] * SHADE_THREADS + threadIdx.x;
            if (nx < nn) {
                const uint32_t i = first + (uint32_t)nx;
                RT_PREFETCH_L2(in.ox + i); RT_PREFETCH_L2(in.oy + i); RT_PREFETCH_L2(in.oz + i);
                RT_PREFETCH_L2(in.dx + i); RT_PREFETCH_L2(in.dy + i); RT_PREFETCH_L2(in.dz + i);
                RT_PREFETCH_L2(in.tr + i); RT_PREFETCH_L2(in.tg + i); RT_PREFETCH_L2(in.tb + i);
                RT_PREFETCH_L2(in.sig + i); RT_PREFETCH_L2(in.key + i); RT_PREFETCH_L2(in.hit_t + i);
                RT_PREFETCH_L2(in.pixel + i); RT_PREFETCH_L2(in.meta + i); RT_PREFETCH_L2(in.hit_id + i);
            }
        }
#endif
        uint32_t sbase = cta_base[par][0], cbase = cta_base[par][1];
#pragma unroll
        for (int k = 0; k < SHADE_WARPS; k++)
            if (k < warp) { sbase += w_shadow[par][k]; cbase += w_child[par][k]; }
        // light loop: light-major slots, so the lanes of a warp store side by side and rays to one light stay together
        const uint32_t lit_rank = (uint32_t)__popc(lit_m & below);
        for (uint32_t li2 = 0; li2 < sc.n_lights; li2++) {
            ShadowRay sr;
            if (shade_light(sc, s, li2, sr)) {
                const uint32_t slot = sbase + li2 * n_lit + lit_rank;
                if (slot >= sq.cap) { *overflow = 1u; continue; }
                sq.ox[slot] = sr.o.x; sq.oy[slot] = sr.o.y; sq.oz[slot] = sr.o.z;
                sq.dx[slot] = sr.d.x; sq.dy[slot] = sr.d.y; sq.dz[slot] = sr.d.z;
                sq.r2[slot] = sr.range2;
                sq.cr[slot] = sr.c.r; sq.cg[slot] = sr.c.g; sq.cb[slot] = sr.c.b;
                sq.pixel[slot] = pixel;
            }
        }
        // spawned rays: slot k of every lane in lock step
        for (uint32_t k = 0; k < (uint32_t)sc.max_children; k++) {
            ChildRay cr;
            const bool want = shade_child(sc, s, k, cr);
            const unsigned m = __ballot_sync(FULL, want);
            const uint32_t slot = cbase + (uint32_t)__popc(m & below);
            cbase += (uint32_t)__popc(m);
            if (!want) continue;
            if (slot >= out.cap) { *overflow = 1u; continue; }
            out.ox[slot] = cr.o.x; out.oy[slot] = cr.o.y; out.oz[slot] = cr.o.z;
            out.dx[slot] = cr.d.x; out.dy[slot] = cr.d.y; out.dz[slot] = cr.d.z;
            out.tr[slot] = cr.T.r; out.tg[slot] = cr.T.g; out.tb[slot] = cr.T.b;
            out.sig[slot] = cr.sig;
            out.key[slot] = cr.key;
            out.pixel[slot] = pixel;
            out.meta[slot] = depth + 1u;
        }
        if (!LOOP) break;
        par ^= 1u;
    }
}

// ---------------------------------------------------------------------------------------------------------
// k_resolve: each thread produces 16 consecutive output bytes (one 128-bit store)
// ---------------------------------------------------------------------------------------------------------
// One thread per 16-byte piece of an output row (rows are padded to a multiple of 4 bytes, so a row's last piece may be
// short).  final_rows: local row yl goes to its row of the whole frame (BMP order), `out` then being a whole-frame buffer --
// the shard's own, or device 0's over NVLink peer access (context.cu rt_multi_*), which is what makes the gather a plain
// store instead of a collective followed by a de-interleave.
__global__ void __launch_bounds__(256) k_resolve(const __grid_constant__ DeviceScene sc, const __grid_constant__ FrameParams fp, const double* accum,
                                                 uint8_t* out, uint32_t bytewidth, uint32_t pieces_per_row, uint32_t final_rows) {
    __shared__ double avg[255];
    for (int i = threadIdx.x; i < 255; i += blockDim.x) avg[i] = sc.srgb_average[i];
    __syncthreads();
    const uint64_t piece = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t row = (uint32_t)(piece / pieces_per_row);
    if (row >= fp.local_rows) return;
    const uint32_t col0 = (uint32_t)(piece - (uint64_t)row * pieces_per_row) * 16u;
    uint32_t out_row = row;
    if (final_rows) {
        const uint32_t band = row / fp.band_rows;
        out_row = (band * fp.world + fp.rank) * fp.band_rows + row % fp.band_rows;
        if (out_row >= fp.height) return;  // padding rows of the last band
    }
    const double cam_div = (double)fp.cam_samples;
    const double aa_div = (double)fp.antialias;
    uint32_t col = col0;
    uint32_t px = col / 3, ch = col - px * 3;
    uint32_t words[4] = {0, 0, 0, 0};
    const uint32_t rowbytes = fp.width * 3;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        uint32_t v = 0;
        if (col < rowbytes) {
            const double a = accum[((size_t)row * fp.width + px) * 3 + (2 - ch)];
            v = to_srgb(avg, (a / cam_div) / aa_div);  // raytrace(): res / samples (src/raytrace.rs:275); main: res / aa (src/main.rs:56)
        }
        words[k >> 2] |= v << (8 * (k & 3));
        col++; ch++;
        if (ch == 3) { ch = 0; px++; }
    }
    uint8_t* dst = out + (size_t)out_row * bytewidth + col0;
    if ((bytewidth & 15u) == 0u) {  // every piece whole and 16-byte aligned
        *reinterpret_cast<uint4*>(dst) = make_uint4(words[0], words[1], words[2], words[3]);
    } else {
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (col0 + 4 * k < bytewidth) reinterpret_cast<uint32_t*>(dst)[k] = words[k];
    }
}

// ---------------------------------------------------------------------------------------------------------
// test hooks
// ---------------------------------------------------------------------------------------------------------
__global__ void k_hit_normals(const __grid_constant__ DeviceScene sc, RayQueue q, uint32_t n, double* t_out, double* normal_out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t id = q.hit_id[i];
    V3 nrm = v3(0, 0, 0);
    double t = (double)INFINITY;
    if (id >= 0) {
        t = q.hit_t[i];
        nrm = shape_normal(sc, id, v3(q.ox[i], q.oy[i], q.oz[i]), v3(q.dx[i], q.dy[i], q.dz[i]), t);
    }
    t_out[i] = t;
    normal_out[3 * (size_t)i] = nrm.x; normal_out[3 * (size_t)i + 1] = nrm.y; normal_out[3 * (size_t)i + 2] = nrm.z;
}

__global__ void k_project(const __grid_constant__ DeviceScene sc, uint64_t seed, uint32_t n, const double* pos, double* od) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t ck = child_key(key_sample(seed, i, 0), CODE_CAMERA0);
    V3 o, d;
    camera_project(sc.camera, pos[2 * (size_t)i], pos[2 * (size_t)i + 1], ck, o, d);
    double* p = od + 6 * (size_t)i;
    p[0] = o.x; p[1] = o.y; p[2] = o.z; p[3] = d.x; p[4] = d.y; p[5] = d.z;
}

__global__ void k_quantize(const double* srgb_average, const double* rgb, uint64_t n, uint8_t* bgr) {
    __shared__ double avg[255];
    for (int i = threadIdx.x; i < 255; i += blockDim.x) avg[i] = srgb_average[i];
    __syncthreads();
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    bgr[3 * i] = to_srgb(avg, rgb[3 * i + 2]);
    bgr[3 * i + 1] = to_srgb(avg, rgb[3 * i + 1]);
    bgr[3 * i + 2] = to_srgb(avg, rgb[3 * i]);
}

// ---------------------------------------------------------------------------------------------------------
// launch wrappers
// ---------------------------------------------------------------------------------------------------------
static inline uint32_t div_up(uint64_t a, uint32_t b) { return (uint32_t)((a + b - 1) / b); }

static void l_primary(cudaStream_t st, const DeviceScene& sc, const FrameParams& fp, RayQueue q, uint32_t* q_count, uint32_t* overflow,
                      uint64_t path_base, uint32_t n_paths) {
    if (!n_paths) return;
    k_primary<<<div_up(n_paths, 256), 256, 0, st>>>(sc, fp, q, q_count, overflow, path_base, n_paths);
}

static void scan_attr_once() {
    static bool done = false;
    if (!done) {
        cudaFuncSetAttribute(k_closest_scan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScanSmem));
        cudaFuncSetAttribute(k_shadow_scan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScanSmem));
        done = true;
    }
}

// A traversal that has looked at the warp this many times (16 nodes or one leaf each) is handed to k_heavy_scan.
// A runaway ray walks ALONE at the tail of the launch, one dependent node fetch after the other (~300 cycles each), while
// the scan costs one CTA n/256 exact tests per thread (~100 cycles each) next to the other CTAs: break-even near n/12000
// looks.  Measured on the 1M-sphere frame: with the old n/256 = 3906 looks one ray held its GPU for 8 ms (and, sharded over
// 8 GPUs, the seven others in the gather); n/8192 hands over after ~2000 nodes, 80 x what an average ray needs.
#ifndef RT_HEAVY_DIV
#define RT_HEAVY_DIV 8192u
#endif
#ifndef RT_HEAVY_MIN
#define RT_HEAVY_MIN 32
#endif
static uint32_t heavy_walks(const DeviceScene& sc) {
    const uint32_t w = sc.n_spheres / RT_HEAVY_DIV;
    return w < (uint32_t)RT_HEAVY_MIN ? (uint32_t)RT_HEAVY_MIN : (w > 4096u ? 4096u : w);
}

static void l_closest(cudaStream_t st, const DeviceScene& sc, bool bvh, RayQueue q, uint32_t first, uint32_t n, Counters* counters, int sm_count,
                      uint32_t* heavy, const uint32_t* n_dev, LaunchCursor* lc, bool reset_lc) {
    if (!n) return;  // with n_dev, n is the upper bound the grid is sized for
    if (bvh) {
        if (reset_lc) cudaMemsetAsync(lc, 0, sizeof(LaunchCursor), st);
        const uint32_t lci = (uint32_t)(lc - counters->lc);  // host arithmetic on device addresses only
        uint32_t want = div_up(n, BVH_THREADS);
        uint32_t grid = (uint32_t)sm_count * (uint32_t)RT_BVH_MIN_BLOCKS;  // resident CTAs of 128 threads per SM
        if (want < grid) grid = want;
        if (sc.bvh_qnodes) k_closest_bvh<true><<<grid, BVH_THREADS, 0, st>>>(sc, q, first, n, counters, heavy, heavy_walks(sc), n_dev, lci);
        else k_closest_bvh<false><<<grid, BVH_THREADS, 0, st>>>(sc, q, first, n, counters, heavy, heavy_walks(sc), n_dev, lci);
        ShadowQueue none{};
        k_heavy_scan<false><<<(uint32_t)sm_count * 4u, HEAVY_THREADS, 0, st>>>(sc, q, none, first, nullptr, counters, heavy, lci);
    } else {
        scan_attr_once();
        k_closest_scan<<<div_up(n, SCAN_THREADS), SCAN_THREADS, sizeof(ScanSmem), st>>>(sc, q, first, n, n_dev, counters);
    }
}

static void l_shade(cudaStream_t st, const DeviceScene& sc, const FrameParams& fp, RayQueue in, uint32_t first, uint32_t n, RayQueue out,
                    uint32_t* out_count, ShadowQueue sq, uint32_t* sq_count, uint32_t* overflow, double* accum, int32_t* ids, int sm_count,
                    const uint32_t* n_dev, uint32_t* tile_cursor) {
    if (!n) return;
    const uint32_t want = div_up(n, SHADE_THREADS);
    const uint32_t grid = (uint32_t)sm_count * (uint32_t)RT_SHADE_MIN_BLOCKS * (uint32_t)RT_SHADE_WAVES;
    if (!tile_cursor || want <= grid) k_shade<false><<<want, SHADE_THREADS, 0, st>>>(sc, fp, in, first, n, out, out_count, sq, sq_count, overflow, accum, ids, n_dev, nullptr);
    else k_shade<true><<<grid, SHADE_THREADS, 0, st>>>(sc, fp, in, first, n, out, out_count, sq, sq_count, overflow, accum, ids, n_dev, tile_cursor);
}

__global__ void k_mark(const uint32_t* src, uint32_t* dst) { *dst = *src; }
static void l_mark(cudaStream_t st, const uint32_t* src, uint32_t* dst) { k_mark<<<1, 1, 0, st>>>(src, dst); }

static void l_shadow(cudaStream_t st, const DeviceScene& sc, bool bvh, ShadowQueue sq, uint32_t n, double* accum, Counters* counters, int sm_count,
                     uint32_t* heavy, const uint32_t* n_dev, LaunchCursor* lc, bool reset_lc, const uint32_t* first_dev) {
    if (!n) return;
    if (bvh) {
        if (reset_lc) cudaMemsetAsync(lc, 0, sizeof(LaunchCursor), st);
        const uint32_t lci = (uint32_t)(lc - counters->lc);
        uint32_t want = div_up(n, BVH_THREADS);
        uint32_t grid = (uint32_t)sm_count * (uint32_t)RT_BVH_MIN_BLOCKS;
        if (want < grid) grid = want;
        if (sc.bvh_qnodes && sc.bvh_wnodes) k_shadow_bvh<true, true><<<grid, BVH_THREADS, 0, st>>>(sc, sq, n, accum, counters, heavy, heavy_walks(sc), n_dev, lci, first_dev);
        else if (sc.bvh_qnodes) k_shadow_bvh<true><<<grid, BVH_THREADS, 0, st>>>(sc, sq, n, accum, counters, heavy, heavy_walks(sc), n_dev, lci, first_dev);
        else k_shadow_bvh<false><<<grid, BVH_THREADS, 0, st>>>(sc, sq, n, accum, counters, heavy, heavy_walks(sc), n_dev, lci, first_dev);
        RayQueue none{};
        k_heavy_scan<true><<<(uint32_t)sm_count * 4u, HEAVY_THREADS, 0, st>>>(sc, none, sq, 0u, accum, counters, heavy, lci);
    } else {
        scan_attr_once();
        k_shadow_scan<<<div_up(n, SCAN_THREADS), SCAN_THREADS, sizeof(ScanSmem), st>>>(sc, sq, n, accum, n_dev, counters);
    }
}

static void l_resolve(cudaStream_t st, const DeviceScene& sc, const FrameParams& fp, const double* accum, uint8_t* out, uint32_t bytewidth, bool final_rows) {
    const uint32_t pieces_per_row = (bytewidth + 15u) / 16u;
    const uint64_t total = (uint64_t)pieces_per_row * fp.local_rows;
    if (!total) return;
    k_resolve<<<div_up(total, 256), 256, 0, st>>>(sc, fp, accum, out, bytewidth, pieces_per_row, final_rows ? 1u : 0u);
}

static void l_hit_normals(cudaStream_t st, const DeviceScene& sc, RayQueue q, uint32_t n, double* t_out, double* normal_out) {
    if (!n) return;
    k_hit_normals<<<div_up(n, 256), 256, 0, st>>>(sc, q, n, t_out, normal_out);
}
static void l_project(cudaStream_t st, const DeviceScene& sc, uint64_t seed, uint32_t n, const double* pos, double* od) {
    if (!n) return;
    k_project<<<div_up(n, 256), 256, 0, st>>>(sc, seed, n, pos, od);
}
static void l_quantize(cudaStream_t st, const double* avg, const double* rgb, uint64_t n, uint8_t* bgr) {
    if (!n) return;
    k_quantize<<<div_up(n, 256), 256, 0, st>>>(avg, rgb, n, bgr);
}

static const KernelTable g_table = {l_primary, l_closest, l_shade, l_shadow, l_mark, l_resolve, l_hit_normals, l_project, l_quantize};

}  // namespace RT_NS

#define RT_CAT_(a, b) a##b
#define RT_CAT(a, b) RT_CAT_(a, b)
const KernelTable* RT_CAT(kernels_, RT_NS)() { return &RT_NS::g_table; }

}  // namespace rt
