// noalloc.cu -- rt_intersect_spheres_noalloc: Scene::intersect over a sphere list for callers that may not allocate
// device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.  It is a variant of the linear-scan
// closest-hit kernel (kernels.cu k_closest_scan: src/scene.rs:247-249 over src/shapes.rs:50-89) whose scratch -- the
// packed SphereRec buffer the scan streams through shared memory and the work cursor of its resident grid -- lives in
// ONE file-scope `static __device__` array instead of memory taken from the allocation cache.  Consequences:
//   * every GPU that loads the library holds RT_NOALLOC_SCRATCH_BYTES (2 MiB + 256 B) until the process exits;
//   * at most RT_NOALLOC_MAX_SPHERES spheres per call (more: RT_ERR_INVALID, nothing launched);
//   * every call on a device shares the array, which keeps the last call's contents.  Each call rewrites the sphere
//     records it uses and zeroes the cursor / counter words itself, in stream order before its scan kernel; two calls
//     that OVERLAP on the GPU (different streams, or two graph launches in flight) race on it.  Serialise them.
//   * the entry point calls no cudaMalloc / cudaFree / cudaStreamCreate / cudaEventCreate and does not synchronise: it
//     queues two kernels on the caller's stream and returns.
// In host code `g_noalloc_scratch` is only a host-side stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, asked once per device with that device current (scratch_for below), and is what the kernels get.
//
// Built with -fmad=false (RT_NS=strict): the same unfused f64 expression as rt_trace_rays in strict mode, so ids and
// hit distances are bit-identical to it.
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "../../include/raytrace_b200.h"
#include "host_scene.h"
#include "rt_device.cuh"

namespace {

using namespace rt;
using namespace rt::RT_NS;

constexpr uint32_t NA_MAX_SPHERES = RT_NOALLOC_MAX_SPHERES;
constexpr int NA_THREADS = 256;   // rays per CTA pass
constexpr int NA_TILE = 1024;     // spheres per shared-memory tile (32 KB)
constexpr int NA_MAX_DEVICES = 64;

// layout of the static scratch
struct NaScratch {
    SphereRec sph[NA_MAX_SPHERES];  // centre + radius*radius of the call's spheres, object order
    uint32_t cursor;                // next unclaimed 256-ray tile (zeroed by k_na_pack)
    uint32_t _pad;
    unsigned long long tests;       // exact sphere tests of the call (zeroed by k_na_pack)
    unsigned char _tail[256 - 16];
};
static_assert(sizeof(NaScratch) == RT_NOALLOC_SCRATCH_BYTES, "RT_NOALLOC_SCRATCH_BYTES in raytrace_b200.h is stale");

static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NaScratch)];

// pack (x, y, z, radius) -> SphereRec and clear the words that must start at zero
__global__ void __launch_bounds__(256) k_na_pack(NaScratch* s, const double* __restrict__ xyzr, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { s->cursor = 0; s->tests = 0ull; }
    if (i >= n) return;
    const double2 a = *reinterpret_cast<const double2*>(xyzr + 4 * (size_t)i);
    const double2 b = *reinterpret_cast<const double2*>(xyzr + 4 * (size_t)i + 2);
    SphereRec r;
    r.cx = a.x; r.cy = a.y; r.cz = b.x;
    r.r2 = b.y * b.y;  // src/shapes.rs:63 `self.radius * self.radius`, hoisted as host_scene.cpp does
    *reinterpret_cast<double4*>(&s->sph[i]) = *reinterpret_cast<const double4*>(&r);
}

// resident grid: each CTA pulls 256-ray tiles off the cursor; every lane tests the same sphere (shared-memory broadcast)
__global__ void __launch_bounds__(NA_THREADS) k_na_closest(NaScratch* s, uint32_t n_spheres, const double* __restrict__ origin_dir, uint32_t n_rays,
                                                           int32_t* __restrict__ ids, double* __restrict__ t_out) {
    __shared__ __align__(32) SphereRec tile[NA_TILE];
    __shared__ uint32_t tile_s;
    unsigned long long tested = 0;
    for (;;) {
        __syncthreads();  // tile_s and tile[] of the previous pass are no longer read
        if (threadIdx.x == 0) tile_s = atomicAdd(&s->cursor, 1u);
        __syncthreads();
        const uint64_t base = (uint64_t)tile_s * NA_THREADS;
        if (base >= n_rays) break;
        const uint32_t i = (uint32_t)base + threadIdx.x;
        const bool active = i < n_rays;
        const double* p = origin_dir + 6 * (size_t)(active ? i : (uint32_t)base);
        const RayQ r = make_rayq(v3(p[0], p[1], p[2]), v3(p[3], p[4], p[5]));
        Closest best = closest_none();
        for (uint32_t first = 0; first < n_spheres; first += NA_TILE) {
            const uint32_t cnt = (n_spheres - first) < (uint32_t)NA_TILE ? (n_spheres - first) : (uint32_t)NA_TILE;
            if (first) __syncthreads();
            for (uint32_t j = threadIdx.x; j < cnt; j += NA_THREADS)
                *reinterpret_cast<double4*>(&tile[j]) = *reinterpret_cast<const double4*>(&s->sph[first + j]);
            __syncthreads();
            if (active) {
                tested += cnt;
#pragma unroll 4
                for (uint32_t j = 0; j < cnt; j++) {
                    const double2 c01 = *reinterpret_cast<const double2*>(&tile[j].cx);
                    const double2 c23 = *reinterpret_cast<const double2*>(&tile[j].cz);
                    double th;
                    if (sphere_hit(r, c01.x, c01.y, c23.x, c23.y, th)) closest_update(best, th, (int32_t)(first + j));
                }
            }
        }
        if (active) {
            ids[i] = closest_id(best);
            if (t_out) t_out[i] = best.t;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tested += __shfl_xor_sync(0xFFFFFFFFu, tested, o);
    if ((threadIdx.x & 31) == 0 && tested) atomicAdd(&s->tests, tested);
}

// device address of the scratch on `device` (which must be current): asked once per device, then remembered
struct NaDevice {
    NaScratch* ptr = nullptr;
    int sm_count = 0;
};
std::mutex g_na_mutex;
NaDevice g_na_dev[NA_MAX_DEVICES];

cudaError_t scratch_for(int device, NaDevice* out) {
    std::lock_guard<std::mutex> lock(g_na_mutex);
    NaDevice& d = g_na_dev[device];
    if (!d.ptr) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess) return e;
        d.sm_count = sms > 0 ? sms : 148;
        d.ptr = static_cast<NaScratch*>(p);
    }
    *out = d;
    return cudaSuccess;
}

}  // namespace

extern "C" int rt_intersect_spheres_noalloc(int device, void* cuda_stream, const double* dev_center_xyz_radius, uint64_t n_spheres,
                                            const double* dev_origin_dir, uint64_t n_rays, int32_t* dev_ids, double* dev_t) {
    using rt::set_error;
    if ((n_spheres && !dev_center_xyz_radius) || (n_rays && (!dev_origin_dir || !dev_ids))) {
        set_error("rt_intersect_spheres_noalloc: null argument");
        return RT_ERR_INVALID;
    }
    if (n_spheres > NA_MAX_SPHERES) {
        set_error("rt_intersect_spheres_noalloc: " + std::to_string(n_spheres) + " spheres exceed the static scratch (RT_NOALLOC_MAX_SPHERES = " +
                  std::to_string(NA_MAX_SPHERES) + "); use rt_trace_rays");
        return RT_ERR_INVALID;
    }
    if (n_rays > 0x7FFFFFFFull) { set_error("rt_intersect_spheres_noalloc: too many rays"); return RT_ERR_INVALID; }
    if (((uintptr_t)dev_center_xyz_radius & 15u) || ((uintptr_t)dev_origin_dir & 7u) || ((uintptr_t)dev_t & 7u) || ((uintptr_t)dev_ids & 3u)) {
        set_error("rt_intersect_spheres_noalloc: misaligned pointer (spheres 16 bytes, rays / t 8, ids 4)");
        return RT_ERR_INVALID;
    }
    if (rt_device_count() <= 0) { set_error("no CUDA device: this library has no CPU fallback"); return RT_ERR_CUDA; }
    if (device < 0 || device >= NA_MAX_DEVICES || device >= rt_device_count()) { set_error("rt_intersect_spheres_noalloc: no such device"); return RT_ERR_INVALID; }
    if (n_rays == 0) return RT_OK;
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e == cudaSuccess && prev != device) e = cudaSetDevice(device);
    NaDevice d;
    if (e == cudaSuccess) e = scratch_for(device, &d);
    if (e == cudaSuccess) {
        cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
        const uint32_t ns = (uint32_t)n_spheres, nr = (uint32_t)n_rays;
        k_na_pack<<<ns ? (ns + 255u) / 256u : 1u, 256, 0, st>>>(d.ptr, dev_center_xyz_radius, ns);
        const uint32_t tiles = (nr + NA_THREADS - 1) / NA_THREADS;
        const uint32_t resident = (uint32_t)d.sm_count * 4u;  // 4 CTAs of 256 threads and 32 KB per SM
        k_na_closest<<<tiles < resident ? tiles : resident, NA_THREADS, 0, st>>>(d.ptr, ns, dev_origin_dir, nr, dev_ids, dev_t);
        e = cudaPeekAtLastError();
    }
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
    if (e != cudaSuccess) { set_error(std::string("rt_intersect_spheres_noalloc: ") + cudaGetErrorString(e)); return RT_ERR_CUDA; }
    return RT_OK;
}
