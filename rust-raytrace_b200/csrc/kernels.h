// kernels.h -- launch interface of the wavefront kernels.  kernels.cu is compiled twice (see rt_device.cuh):
// rt::kernels_strict() (-fmad=false) and rt::kernels_fast() (-fmad=true) return the two instantiations.
#pragma once
#include <cuda_runtime.h>

#include "rt_types.h"

namespace rt {

// device-side counters of one context (one allocation, zeroed/reset by the host driver)
struct Counters {
    uint32_t overflow;      // a queue append was dropped (host sizing bug) -> render fails loudly
    uint32_t shadow_count;  // shadow queue length
    uint32_t cursor;        // persistent-thread work cursor (reset before each persistent launch)
    uint32_t heavy_count;   // rays the traversal kernel handed to the block-wide scan (reset with the cursor)
    unsigned long long bvh_nodes[2];  // 64-byte BVH nodes fetched: [0] closest-hit kernel, [1] shadow kernel (whole render)
    unsigned long long bvh_prims[2];  // exact sphere tests done at BVH leaves
    uint32_t level_count[64];  // ray-queue length per wavefront level
    unsigned long long heavy_rays[2];  // rays finished by the block-wide scan (whole render): [0] closest, [1] shadow
};
constexpr uint32_t HEAVY_CAP = 1u << 16;  // entries of the heavy-ray list (rays beyond it simply keep traversing)
// Behind the list: scratch for splitting a few heavy rays over many CTAs (k_heavy_scan).  Zeroed once when allocated;
// the last CTA of a ray leaves its words zero again.
constexpr uint32_t HEAVY_SPLIT_RAYS = 64;     // split mode when at most this many rays are on the list
constexpr uint32_t HEAVY_SPLIT_ITEMS = 2048;  // >= grid size of k_heavy_scan
struct HeavyScratch {
    uint32_t done[HEAVY_SPLIT_RAYS];          // CTAs that have finished their slice of ray e
    uint32_t flag[HEAVY_SPLIT_RAYS];          // shadow: some slice found an occluder
    double t[HEAVY_SPLIT_ITEMS];              // closest: best hit of each slice
    int32_t obj[HEAVY_SPLIT_ITEMS];
    int32_t nan[HEAVY_SPLIT_ITEMS];
};
constexpr size_t HEAVY_WORDS = HEAVY_CAP + (sizeof(HeavyScratch) + 3) / 4;  // uint32 words to allocate for `heavy`

struct KernelTable {
    // camera.rs primary-ray generation: paths [path_base, path_base + n_paths) -> level-0 queue
    void (*primary)(cudaStream_t, const DeviceScene&, const FrameParams&, RayQueue q, uint32_t* q_count, uint32_t* overflow, uint64_t path_base,
                    uint32_t n_paths);
    // Scene::intersect for rays [first, first+n) of q; fills hit_id / hit_t
    // `heavy`: device scratch of HEAVY_CAP uint32 (BVH only): rays whose traversal runs away (huge filter padding for origins
    // 10^7+ units out) are finished by a block-wide scan instead of holding one lane for milliseconds
    // n_dev != nullptr (BVH only): the ray count is read on the device from *n_dev (first must be 0, n is ignored) -- the
    // launch can be queued before the host knows how many rays the level holds
    void (*closest)(cudaStream_t, const DeviceScene&, bool bvh, RayQueue q, uint32_t first, uint32_t n, Counters* counters, int sm_count,
                    uint32_t* heavy, const uint32_t* n_dev);
    // material evaluation: pixel accumulation, shadow-ray and child-ray emission with warp-aggregated appends
    void (*shade)(cudaStream_t, const DeviceScene&, const FrameParams&, RayQueue in, uint32_t first, uint32_t n, RayQueue out, uint32_t* out_count,
                  ShadowQueue sq, uint32_t* sq_count, uint32_t* overflow, double* accum, int32_t* ids);
    // shadow queries; unoccluded rays add their pending contribution to the pixel
    void (*shadow)(cudaStream_t, const DeviceScene&, bool bvh, ShadowQueue sq, uint32_t n, double* accum, Counters* counters, int sm_count,
                   uint32_t* heavy);
    // average, quantise (threshold table) and pack BGR rows with padding; 128-bit stores
    void (*resolve)(cudaStream_t, const DeviceScene&, const FrameParams&, const double* accum, uint8_t* out, uint32_t bytewidth);
    // test hooks
    void (*hit_normals)(cudaStream_t, const DeviceScene&, RayQueue q, uint32_t n, double* t_out, double* normal_out);
    void (*project)(cudaStream_t, const DeviceScene&, uint64_t seed, uint32_t n, const double* pos_xy, double* origin_dir);
    void (*quantize)(cudaStream_t, const double* srgb_average, const double* rgb, uint64_t n, uint8_t* bgr);
};

const KernelTable* kernels_strict();
const KernelTable* kernels_fast();

// mode-independent kernels (bvh_build.cu)
// where the build takes its device memory from (context.cu passes its allocation cache; release(nullptr) is a no-op)
struct DeviceAllocator {
    cudaError_t (*alloc)(void** p, size_t bytes);
    void (*release)(void* p);
};
struct BvhBuildResult {
    BvhNode* nodes;       // [n_nodes]; null when qnodes is set
    BvhQNode* qnodes;     // [n_nodes] 32-byte grid form, when asked for and every box fits the grid
    BvhWNode* wnodes;     // [n_nodes] wide form of qnodes (rt_types.h BvhWNode), when asked for and qnodes exists
    SphereRec* sph;       // Morton order
    int32_t* obj;
    uint32_t n_nodes;
    uint32_t launches;
};
// Builds the LBVH over `n` spheres (scan-order records + object ids, device pointers).  Device-side throughout:
// Morton codes -> bitonic sort -> Karras hierarchy -> bottom-up refit -> packed f32 child boxes.
// q_org/q_cell non-null: also pack the nodes on that grid (rt_types.h BvhQNode) and drop the 64-byte form.
// wide: also derive the wide nodes from the grid form (experimental, shadow queries).
cudaError_t bvh_build(cudaStream_t, const SphereRec* sph, const int32_t* sph_obj, uint32_t n, const double* centroid_lo, const double* centroid_hi,
                      const double* q_org, const double* q_cell, const DeviceAllocator& A, BvhBuildResult* out, bool wide = false);
void bvh_free(BvhBuildResult* r, const DeviceAllocator& A);

// de-interleave row bands after the framebuffer all-gather (rank-major input)
void launch_untile(cudaStream_t, const uint32_t* gathered, uint32_t* out, uint32_t height, uint32_t words_per_row, uint32_t band_rows, uint32_t world,
                   uint32_t local_rows);

}  // namespace rt
