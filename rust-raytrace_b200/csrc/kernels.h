// kernels.h -- launch interface of the wavefront kernels.  kernels.cu is compiled twice (see rt_device.cuh):
// rt::kernels_strict() (-fmad=false) and rt::kernels_fast() (-fmad=true) return the two instantiations.
#pragma once
#include <cuda_runtime.h>

#include "rt_types.h"

namespace rt {

// device-side counters of one context (one allocation, zeroed/reset by the host driver)
// work cursor of ONE persistent launch + the length of its heavy-ray list.  Device-driven frames give every launch its own
// (zeroed with the whole record at the start of the frame), so nothing is reset between launches.
struct LaunchCursor {
    uint32_t cursor;  // next unclaimed ray
    uint32_t heavy;   // rays the traversal handed to the block-wide scan
};
constexpr int LC_SHADOW = 64;  // Counters.lc index of the frame's shadow launch (0..63: closest-hit pass of that level)
constexpr int LC_SHARED = 65;  // host-driven launches: one cursor, reset before each launch
constexpr int LC_SHADOW_SEG = 66;  // + d: the shadow launch of wavefront level d's segment (device-driven frames, RT_TUNE_SERIAL_SHADOW off)
struct Counters {
    // whole frame
    uint32_t overflow;      // a queue append was dropped -> the frame is redone host-driven (which slices and cannot overflow)
    uint32_t _pad0;
    unsigned long long bvh_nodes[2];   // BVH nodes fetched: [0] closest-hit kernel, [1] shadow kernel
    unsigned long long bvh_prims[2];   // exact sphere tests done at BVH leaves
    unsigned long long heavy_rays[2];  // rays finished by the block-wide scan: [0] closest, [1] shadow
    unsigned long long scan_tests[2];  // exact sphere tests made by the linear-scan kernels: [0] closest, [1] shadow
    // per primary batch (a device-driven frame zeroes from here to the end of the record between batches)
    uint32_t shadow_count;     // shadow queue length
    uint32_t _pad1;
    uint32_t level_count[64];  // ray-queue length per wavefront level
    LaunchCursor lc[66 + 64];
    uint32_t shade_cursor[64];  // tile cursor of the shade pass of level d (device-driven frames)
    uint32_t shadow_mark[66];   // shadow queue length before level d's shade pass ([0] stays 0): segment d = [mark[d], mark[d + 1])
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
    // n_dev != nullptr: the ray count is read on the device from *n_dev (clamped to the queue's capacity; first must be 0
    // and n is then only an upper bound used to size the grid) -- the launch can be queued before the host knows how
    // many rays the level holds.  lc: this launch's cursor (BVH); reset_lc: zero it first (shared cursor of host-driven frames)
    void (*closest)(cudaStream_t, const DeviceScene&, bool bvh, RayQueue q, uint32_t first, uint32_t n, Counters* counters, int sm_count,
                    uint32_t* heavy, const uint32_t* n_dev, LaunchCursor* lc, bool reset_lc);
    // material evaluation: pixel accumulation, shadow-ray and child-ray emission with warp-aggregated appends
    // n_dev as for closest; tile_cursor != nullptr: n is a loose bound -- a resident grid pulls tiles off this (zeroed)
    // cursor instead of one CTA per tile
    void (*shade)(cudaStream_t, const DeviceScene&, const FrameParams&, RayQueue in, uint32_t first, uint32_t n, RayQueue out, uint32_t* out_count,
                  ShadowQueue sq, uint32_t* sq_count, uint32_t* overflow, double* accum, int32_t* ids, int sm_count, const uint32_t* n_dev,
                  uint32_t* tile_cursor);
    // shadow queries; unoccluded rays add their pending contribution to the pixel
    // first_dev != nullptr (BVH only, with n_dev): the launch traces queue entries [*first_dev, *n_dev) -- one level's segment
    void (*shadow)(cudaStream_t, const DeviceScene&, bool bvh, ShadowQueue sq, uint32_t n, double* accum, Counters* counters, int sm_count,
                   uint32_t* heavy, const uint32_t* n_dev, LaunchCursor* lc, bool reset_lc, const uint32_t* first_dev);
    // *dst = *src on the stream (one thread): pins a queue length between two launches that both append to the queue
    void (*mark)(cudaStream_t, const uint32_t* src, uint32_t* dst);
    // average, quantise (threshold table) and pack BGR rows with padding; 128-bit stores.  final_rows == false: `out` is
    // the shard's own framebuffer (local row order); true: `out` is a whole-frame buffer in BMP row order (possibly another
    // device's memory over NVLink peer access) and every local row goes to its place in it
    void (*resolve)(cudaStream_t, const DeviceScene&, const FrameParams&, const double* accum, uint8_t* out, uint32_t bytewidth, bool final_rows);
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
    uint32_t max_depth;   // internal nodes above the deepest leaf, measured on the device; <= BVH_STACK or the build fails
};
// Builds the LBVH over `n` spheres (scan-order records + object ids, device pointers).  Device-side throughout:
// Morton codes -> bitonic sort -> Karras hierarchy -> bottom-up refit -> packed f32 child boxes.
// q_org/q_cell non-null: also pack the nodes on that grid (rt_types.h BvhQNode) and drop the 64-byte form.
// wide: also derive the four-wide nodes from the grid form (shadow queries).  morton_per_axis: RT_TUNE_MORTON_AXIS.
cudaError_t bvh_build(cudaStream_t, const SphereRec* sph, const int32_t* sph_obj, uint32_t n, const double* centroid_lo, const double* centroid_hi,
                      const double* q_org, const double* q_cell, const DeviceAllocator& A, BvhBuildResult* out, bool wide, bool morton_per_axis);
void bvh_free(BvhBuildResult* r, const DeviceAllocator& A);

// de-interleave row bands after the framebuffer all-gather (rank-major input)
void launch_untile(cudaStream_t, const uint32_t* gathered, uint32_t* out, uint32_t height, uint32_t words_per_row, uint32_t band_rows, uint32_t world,
                   uint32_t local_rows);

}  // namespace rt
