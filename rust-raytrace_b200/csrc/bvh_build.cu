// bvh_build.cu -- device-side LBVH construction over the sphere buffer, plus the framebuffer de-interleave kernel.
//
// No reference counterpart: upstream only has the O(N) scan of src/scene.rs:247-249.  The tree is an
// acceleration structure whose traversal (rt_device.cuh: bvh_traverse) must return exactly the scan's result.
//
//   k_morton        63-bit Morton code of every sphere centre (21 bits per axis over the centroid bounds)
//   k_bitonic_*     sort (code, index) pairs; the pair is a total order, so the tree is deterministic
//   k_gather        sphere records / object ids into Morton order
//   k_hierarchy     Karras (2012) radix-tree node per thread
//   k_refit_pack    bottom-up: the second thread to reach a node writes the packed 64-byte node (both child boxes,
//                   f32 rounded outward) and the node's own bounds
#include <cuda_runtime.h>

#include <cstdio>

#include "kernels.h"
#define RT_NS build
#include "rt_device.cuh"

namespace rt {

using namespace rt::build;

namespace {

__global__ void k_morton(const SphereRec* sph, uint32_t n, uint32_t n_pow2, double lox, double loy, double loz, double ix, double iy, double iz,
                         uint64_t* keys, uint32_t* idx) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pow2) return;
    if (i < n) {
        const double lo[3] = {lox, loy, loz};
        const double inv[3] = {ix, iy, iz};
        keys[i] = morton63(sph[i].cx, sph[i].cy, sph[i].cz, lo, inv);
        idx[i] = i;
    } else {
        keys[i] = ~0ULL;
        idx[i] = ~0u;
    }
}

__device__ __forceinline__ bool pair_greater(uint64_t ka, uint32_t ia, uint64_t kb, uint32_t ib) { return ka > kb || (ka == kb && ia > ib); }

// one compare-exchange step of the bitonic network in global memory
__global__ void k_bitonic_global(uint64_t* keys, uint32_t* idx, uint32_t n_pow2, uint32_t j, uint32_t k) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_pow2 / 2) return;
    const uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));  // insert a 0 bit at position log2(j)
    const uint32_t p = i | j;
    const bool up = (i & k) == 0;
    const uint64_t ka = keys[i], kb = keys[p];
    const uint32_t ia = idx[i], ib = idx[p];
    if (pair_greater(ka, ia, kb, ib) == up) {
        keys[i] = kb; keys[p] = ka;
        idx[i] = ib; idx[p] = ia;
    }
}

// all steps with j < BLOCK_ELEMS for a given k, inside shared memory (BLOCK_ELEMS = 2048 pairs per CTA)
constexpr uint32_t BITONIC_ELEMS = 2048;
__global__ void __launch_bounds__(BITONIC_ELEMS / 2) k_bitonic_shared(uint64_t* keys, uint32_t* idx, uint32_t k_first, uint32_t k_last) {
    __shared__ uint64_t sk[BITONIC_ELEMS];
    __shared__ uint32_t si[BITONIC_ELEMS];
    const uint32_t base = blockIdx.x * BITONIC_ELEMS;
    const uint32_t t = threadIdx.x;
    sk[t] = keys[base + t]; si[t] = idx[base + t];
    sk[t + BITONIC_ELEMS / 2] = keys[base + t + BITONIC_ELEMS / 2]; si[t + BITONIC_ELEMS / 2] = idx[base + t + BITONIC_ELEMS / 2];
    __syncthreads();
    for (uint32_t k = k_first; k <= k_last; k <<= 1) {
        uint32_t j = (k >> 1) < (BITONIC_ELEMS / 2) ? (k >> 1) : (BITONIC_ELEMS / 2);
        for (; j >= 1; j >>= 1) {
            const uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
            const uint32_t p = i | j;
            const bool up = ((base + i) & k) == 0;
            const uint64_t ka = sk[i], kb = sk[p];
            const uint32_t ia = si[i], ib = si[p];
            if (pair_greater(ka, ia, kb, ib) == up) {
                sk[i] = kb; sk[p] = ka;
                si[i] = ib; si[p] = ia;
            }
            __syncthreads();
        }
    }
    keys[base + t] = sk[t]; idx[base + t] = si[t];
    keys[base + t + BITONIC_ELEMS / 2] = sk[t + BITONIC_ELEMS / 2]; idx[base + t + BITONIC_ELEMS / 2] = si[t + BITONIC_ELEMS / 2];
}

__global__ void k_gather(const SphereRec* sph, const int32_t* sph_obj, const uint32_t* idx, uint32_t n, SphereRec* out_sph, int32_t* out_obj) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t s = idx[i];
    out_sph[i] = sph[s];
    out_obj[i] = sph_obj[s];
}

__global__ void k_hierarchy(const uint64_t* keys, uint32_t n, int32_t* child0, int32_t* child1, int32_t* parent_internal, int32_t* parent_leaf) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 >= n) return;
    int32_t l, r;
    lbvh_node(keys, (int)n, (int)i, l, r);
    child0[i] = l; child1[i] = r;
    if (l < 0) parent_leaf[~l] = (int32_t)i; else parent_internal[l] = (int32_t)i;
    if (r < 0) parent_leaf[~r] = (int32_t)i; else parent_internal[r] = (int32_t)i;
    if (i == 0) parent_internal[0] = -1;
}

// depth of the tree = most internal nodes above any leaf.  The traversal stacks are sized for it and the walks do not
// check them (rt_device.cuh BVH_STACK): bvh_build refuses a deeper tree.
__global__ void k_tree_depth(const int32_t* parent_internal, const int32_t* parent_leaf, uint32_t n, uint32_t* max_depth) {
    const uint32_t leaf = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t d = 0;
    if (leaf < n)
        for (int32_t node = parent_leaf[leaf]; node >= 0; node = parent_internal[node]) d++;
    for (int o = 16; o > 0; o >>= 1) d = max(d, __shfl_xor_sync(0xFFFFFFFFu, d, o));
    if ((threadIdx.x & 31) == 0 && d) atomicMax(max_depth, d);
}

struct Box {
    float lo[3], hi[3];
};
__device__ __forceinline__ Box leaf_box(const SphereRec& s) {
    const double r = sqrt(s.r2) * (1.0 + 0x1.0p-50);  // sqrt is correctly rounded; the factor covers its half-ulp
    Box b;
    b.lo[0] = __double2float_rd(s.cx - r); b.lo[1] = __double2float_rd(s.cy - r); b.lo[2] = __double2float_rd(s.cz - r);
    b.hi[0] = __double2float_ru(s.cx + r); b.hi[1] = __double2float_ru(s.cy + r); b.hi[2] = __double2float_ru(s.cz + r);
    return b;
}

// child bounds were written by another thread: read them through L2 (skip the non-coherent L1)
__device__ __forceinline__ Box load_box_cg(const Box* p) {
    Box b;
    for (int a = 0; a < 3; a++) { b.lo[a] = __ldcg(&p->lo[a]); b.hi[a] = __ldcg(&p->hi[a]); }
    return b;
}

__global__ void k_refit_pack(const SphereRec* sph, uint32_t n, const int32_t* child0, const int32_t* child1, const int32_t* parent_internal,
                             const int32_t* parent_leaf, uint32_t* flags, Box* node_box, BvhNode* nodes) {
    const uint32_t leaf = blockIdx.x * blockDim.x + threadIdx.x;
    if (leaf >= n) return;
    int32_t node = parent_leaf[leaf];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&flags[node], 1u) == 0u) return;  // first arrival: the sibling subtree is not finished yet
        __threadfence();
        const int32_t c0 = child0[node], c1 = child1[node];
        const Box b0 = c0 < 0 ? leaf_box(sph[~c0]) : load_box_cg(&node_box[c0]);
        const Box b1 = c1 < 0 ? leaf_box(sph[~c1]) : load_box_cg(&node_box[c1]);
        BvhNode nd;
        Box u;
        for (int a = 0; a < 3; a++) {
            nd.lo0[a] = b0.lo[a]; nd.hi0[a] = b0.hi[a];
            nd.lo1[a] = b1.lo[a]; nd.hi1[a] = b1.hi[a];
            u.lo[a] = fminf(b0.lo[a], b1.lo[a]);
            u.hi[a] = fmaxf(b0.hi[a], b1.hi[a]);
        }
        nd.c0 = c0; nd.c1 = c1; nd._pad[0] = 0; nd._pad[1] = 0;
        nodes[node] = nd;
        node_box[node] = u;
        node = parent_internal[node];
    }
}

__global__ void k_single_leaf(const SphereRec* sph, BvhNode* nodes) {
    const Box b = leaf_box(sph[0]);
    BvhNode nd;
    for (int a = 0; a < 3; a++) { nd.lo0[a] = b.lo[a]; nd.hi0[a] = b.hi[a]; nd.lo1[a] = b.lo[a]; nd.hi1[a] = b.hi[a]; }
    nd.c0 = ~0; nd.c1 = ~0; nd._pad[0] = 0; nd._pad[1] = 0;  // the one leaf twice: testing it twice is idempotent
    nodes[0] = nd;
}

struct QuantGrid { double org[3], cell[3]; };
__global__ void k_quantise(const BvhNode* nodes, uint32_t n_nodes, QuantGrid g, BvhQNode* out, uint32_t* off_grid) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    bool ok = true;
    out[i] = quantise_node(nodes[i], g.org, g.cell, ok);
    if (!ok) atomicAdd(off_grid, 1u);
}

__global__ void k_widen(const BvhQNode* qnodes, uint32_t n_nodes, BvhWNode* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    out[i] = widen_node(qnodes, i);
}

__global__ void k_untile(const uint32_t* gathered, uint32_t* out, uint32_t height, uint32_t words_per_row, uint32_t band_rows, uint32_t world,
                         uint32_t local_rows) {
    const uint64_t w = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t total = (uint64_t)height * words_per_row;
    if (w >= total) return;
    out[w] = gathered[untile_source_word(w, words_per_row, band_rows, world, local_rows)];
}

}  // namespace

#define BVH_CHECK(x)                         \
    do {                                     \
        cudaError_t e_ = (x);                \
        if (e_ != cudaSuccess) { bvh_free(out, A); A.release(keys); A.release(idx); A.release(child0); A.release(child1); A.release(pint); A.release(pleaf); A.release(flags); A.release(boxes); A.release(depth_dev); return e_; } \
    } while (0)

cudaError_t bvh_build(cudaStream_t st, const SphereRec* sph, const int32_t* sph_obj, uint32_t n, const double* clo, const double* chi,
                      const double* q_org, const double* q_cell, const DeviceAllocator& A, BvhBuildResult* out, bool wide, bool morton_per_axis) {
    out->nodes = nullptr; out->qnodes = nullptr; out->wnodes = nullptr; out->sph = nullptr; out->obj = nullptr; out->n_nodes = 0; out->launches = 0;
    out->max_depth = n > 1 ? 0 : 1;
    uint32_t* depth_dev = nullptr;
    uint64_t* keys = nullptr; uint32_t* idx = nullptr;
    int32_t *child0 = nullptr, *child1 = nullptr, *pint = nullptr, *pleaf = nullptr;
    uint32_t* flags = nullptr; Box* boxes = nullptr;
    if (n == 0) return cudaSuccess;
    uint32_t n_pow2 = BITONIC_ELEMS;
    while (n_pow2 < n) n_pow2 <<= 1;
    const uint32_t n_nodes = n > 1 ? n - 1 : 1;
    BVH_CHECK(A.alloc((void**)&keys, sizeof(uint64_t) * n_pow2));
    BVH_CHECK(A.alloc((void**)&idx, sizeof(uint32_t) * n_pow2));
    BVH_CHECK(A.alloc((void**)&out->nodes, sizeof(BvhNode) * n_nodes));
    BVH_CHECK(A.alloc((void**)&out->sph, sizeof(SphereRec) * n));
    BVH_CHECK(A.alloc((void**)&out->obj, sizeof(int32_t) * n));
    out->n_nodes = n_nodes;
    double inv[3];
    morton_scale(clo, chi, inv, morton_per_axis);
    const uint32_t T = 256;
    k_morton<<<(n_pow2 + T - 1) / T, T, 0, st>>>(sph, n, n_pow2, clo[0], clo[1], clo[2], inv[0], inv[1], inv[2], keys, idx);
    out->launches++;
    // bitonic sort: stages with k <= BITONIC_ELEMS run entirely in shared memory; for larger k the steps with
    // j >= BITONIC_ELEMS go through global memory and the tail (j < BITONIC_ELEMS) through shared memory again
    k_bitonic_shared<<<n_pow2 / BITONIC_ELEMS, BITONIC_ELEMS / 2, 0, st>>>(keys, idx, 2, BITONIC_ELEMS);
    out->launches++;
    for (uint32_t k = BITONIC_ELEMS * 2; k <= n_pow2; k <<= 1) {
        for (uint32_t j = k >> 1; j >= BITONIC_ELEMS; j >>= 1) {
            k_bitonic_global<<<(n_pow2 / 2 + T - 1) / T, T, 0, st>>>(keys, idx, n_pow2, j, k);
            out->launches++;
        }
        k_bitonic_shared<<<n_pow2 / BITONIC_ELEMS, BITONIC_ELEMS / 2, 0, st>>>(keys, idx, k, k);
        out->launches++;
    }
    k_gather<<<(n + T - 1) / T, T, 0, st>>>(sph, sph_obj, idx, n, out->sph, out->obj);
    out->launches++;
    if (n == 1) {
        k_single_leaf<<<1, 1, 0, st>>>(out->sph, out->nodes);
        out->launches++;
    } else {
        BVH_CHECK(A.alloc((void**)&child0, sizeof(int32_t) * n_nodes));
        BVH_CHECK(A.alloc((void**)&child1, sizeof(int32_t) * n_nodes));
        BVH_CHECK(A.alloc((void**)&pint, sizeof(int32_t) * n_nodes));
        BVH_CHECK(A.alloc((void**)&pleaf, sizeof(int32_t) * n));
        BVH_CHECK(A.alloc((void**)&flags, sizeof(uint32_t) * n_nodes));
        BVH_CHECK(A.alloc((void**)&boxes, sizeof(Box) * n_nodes));
        BVH_CHECK(cudaMemsetAsync(flags, 0, sizeof(uint32_t) * n_nodes, st));
        k_hierarchy<<<(n_nodes + T - 1) / T, T, 0, st>>>(keys, n, child0, child1, pint, pleaf);
        k_refit_pack<<<(n + T - 1) / T, T, 0, st>>>(out->sph, n, child0, child1, pint, pleaf, flags, boxes, out->nodes);
        BVH_CHECK(A.alloc((void**)&depth_dev, sizeof(uint32_t)));
        BVH_CHECK(cudaMemsetAsync(depth_dev, 0, sizeof(uint32_t), st));
        k_tree_depth<<<(n + T - 1) / T, T, 0, st>>>(pint, pleaf, n, depth_dev);
        BVH_CHECK(cudaMemcpyAsync(&out->max_depth, depth_dev, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        out->launches += 3;
    }
    if (q_org && q_cell) {
        if (!flags) BVH_CHECK(A.alloc((void**)&flags, sizeof(uint32_t)));
        uint32_t* off_grid = flags;  // the refit is done with its arrival flags (stream order): word 0 becomes the counter
        BVH_CHECK(A.alloc((void**)&out->qnodes, sizeof(BvhQNode) * n_nodes));
        BVH_CHECK(cudaMemsetAsync(off_grid, 0, sizeof(uint32_t), st));
        QuantGrid g;
        for (int a = 0; a < 3; a++) { g.org[a] = q_org[a]; g.cell[a] = q_cell[a]; }
        k_quantise<<<(n_nodes + T - 1) / T, T, 0, st>>>(out->nodes, n_nodes, g, out->qnodes, off_grid);
        out->launches++;
        uint32_t bad = 0;
        BVH_CHECK(cudaMemcpyAsync(&bad, off_grid, sizeof bad, cudaMemcpyDeviceToHost, st));
        BVH_CHECK(cudaStreamSynchronize(st));
        if (bad) { A.release(out->qnodes); out->qnodes = nullptr; }   // a box off the grid: keep the 64-byte nodes
        else { A.release(out->nodes); out->nodes = nullptr; }
        if (wide && out->qnodes) {
            BVH_CHECK(A.alloc((void**)&out->wnodes, sizeof(BvhWNode) * n_nodes));
            k_widen<<<(n_nodes + T - 1) / T, T, 0, st>>>(out->qnodes, n_nodes, out->wnodes);
            out->launches++;
        }
    }
    BVH_CHECK(cudaStreamSynchronize(st));
    BVH_CHECK(cudaGetLastError());
    if (out->max_depth > (uint32_t)BVH_STACK || 3 * ((out->max_depth + 1) / 2) + 1 > (uint32_t)BVH_STACK_WIDE) {
        // cannot happen for 95-bit keys; if it ever does the scene is refused instead of walked with a short stack
        BVH_CHECK(cudaErrorInvalidValue);
    }
    A.release(depth_dev);
    A.release(keys); A.release(idx); A.release(child0); A.release(child1); A.release(pint); A.release(pleaf); A.release(flags); A.release(boxes);
    return cudaSuccess;
}

void bvh_free(BvhBuildResult* r, const DeviceAllocator& A) {
    if (!r) return;
    A.release(r->nodes); A.release(r->qnodes); A.release(r->wnodes); A.release(r->sph); A.release(r->obj);
    r->nodes = nullptr; r->qnodes = nullptr; r->wnodes = nullptr; r->sph = nullptr; r->obj = nullptr; r->n_nodes = 0;
}

void launch_untile(cudaStream_t st, const uint32_t* gathered, uint32_t* out, uint32_t height, uint32_t words_per_row, uint32_t band_rows,
                   uint32_t world, uint32_t local_rows) {
    const uint64_t total = (uint64_t)height * words_per_row;
    if (!total) return;
    k_untile<<<(uint32_t)((total + 255) / 256), 256, 0, st>>>(gathered, out, height, words_per_row, band_rows, world, local_rows);
}

}  // namespace rt
