// rt_types.h -- plain-old-data shared by the host runtime and the CUDA kernels.
//
// Layout in HBM (DESIGN.md "Data layout"): the reference's `Vec<Object>` of boxed trait objects
// (src/scene.rs:93-98, :201-212) is flattened into tagged structure-of-arrays buffers; no vtables, no
// pointer chasing per object.
#pragma once
#include <stdint.h>

namespace rt {

enum : int32_t { MAT_PHONG = 0, MAT_INDIRECT = 1, MAT_FRESNEL = 2, MAT_TRANSPARENT = 3 };
enum : int32_t { LIGHT_POINT = 0, LIGHT_DIRECTIONAL = 1, LIGHT_AREA = 2 };
enum : int32_t { CAM_SPC = 0, CAM_DOF = 1 };
enum : int32_t { BG_SOLID = 0, BG_SKYBOX = 1 };

// one sphere in a scan / BVH-leaf buffer: centre and radius*radius (the product `self.radius * self.radius`
// of src/shapes.rs:63 is a pure function of the radius, so it is hoisted to scene-build time)
struct alignas(32) SphereRec {
    double cx, cy, cz, r2;
};

struct PlaneRec {
    double p[3];
    double n[3];
    int32_t obj;
    int32_t _pad;
};

// 96 bytes = three 32-byte sectors
struct alignas(32) MaterialRec {
    int32_t kind;
    uint32_t samples;
    double diffuse[3];
    double specular[3];
    double ambient[3];
    double exponent;
    double ior;
};

struct LightRec {
    int32_t kind;
    int32_t _pad;
    double a[3];  // location | direction | origin
    double s1[3];
    double s2[3];
    double color[3];
};

struct CameraRec {
    int32_t kind;
    uint32_t samples;
    double pos[3];
    double m[9];  // row-major Mat3 (src/camera.rs:57-61)
    double focus, aperture, im_dist;
};

struct TextureRec {
    uint32_t width, height;
    const uint8_t* data;  // RGB8, row 0 = top
};

// 64-byte BVH node: both children's boxes (outward-rounded f32) + child links.
// child >= 0: internal node index; child < 0: leaf, ~child = slot in the Morton-ordered sphere buffer.
struct alignas(64) BvhNode {
    float lo0[3], hi0[3];
    float lo1[3], hi1[3];
    int32_t c0, c1;
    int32_t _pad[2];
};

// 32-byte BVH node: the same two child boxes on the scene's 16-bit grid (coordinate = q_org + q * q_cell per axis,
// lower bounds rounded down, upper bounds rounded up) -- one 256-bit load per visit.  Used when the grid is fine
// enough for the scene (host_scene.cpp choose_quantisation), else the 64-byte node above.
struct alignas(32) BvhQNode {
    uint16_t q[12];  // lo0[3], hi0[3], lo1[3], hi1[3]
    int32_t c0, c1;
};

// 64-byte wide node (shadow queries; off with RT_TUNE_BINARY_SHADOW): the four grandchildren of binary node i
// as two BvhQNode halves -- h[0] = the two children of node i's child 0, h[1] = those of its child 1 -- so a visit tests four
// boxes behind one dependent fetch (two 256-bit loads of one 64-byte line) instead of two.  A child of node i that is a
// leaf fills slot 0 of its half with its own box and link; slot 1 is then empty: link TRAV_EMPTY, box inside out.
struct alignas(64) BvhWNode {
    BvhQNode h[2];
};

// Everything a kernel needs to know about the scene (passed by value as a __grid_constant__ parameter).
struct DeviceScene {
    uint32_t n_objects, n_spheres, n_planes, n_lights;
    // linear-scan order = object order among the spheres
    const SphereRec* sph;     // [n_spheres]
    const int32_t* sph_obj;   // [n_spheres] object index of scan slot k
    const PlaneRec* planes;   // [n_planes]
    // per object: shape lookup for shading.  bit 31 set = plane (low bits index planes[]), else index into sph[]
    const uint32_t* obj_shape;     // [n_objects]
    const uint32_t* obj_material;  // [n_objects]
    const MaterialRec* materials;
    const LightRec* lights;
    CameraRec camera;
    int32_t bg_kind;
    int32_t max_children;  // max rays one shaded hit can spawn (1 Phong/Fresnel, 2 Transparent, samples Indirect)
    double bg_color[3];
    TextureRec faces[6];
    const double* srgb_values;   // [256] sRGB byte -> linear
    const double* srgb_average;  // [255] quantiser thresholds
    // BVH over the spheres (planes stay in the side list); null when accel == linear
    const BvhNode* bvh_nodes;
    const SphereRec* bvh_sph;  // Morton order
    const int32_t* bvh_obj;    // object index of Morton slot k
    uint32_t n_bvh_nodes;
    uint32_t _pad_bvh;
    double bvh_cmax;  // max over spheres of (|c|_inf + r): scale term of the per-ray box padding
    double bvh_rmin;  // min radius
    double bvh_inv2rmin;  // 0.5 / bvh_rmin, rounded up
    double bvh_lo[3], bvh_hi[3];  // bounds of all sphere boxes (f64): rays are advanced to this box before the f32 filter
    float bvh_lo_f[3], bvh_hi_f[3];  // the same rounded outward to f32 (+-inf when unusable)
    const BvhQNode* bvh_qnodes;   // non-null: traverse these instead of bvh_nodes
    double q_org[3], q_cell[3];   // the grid of bvh_qnodes
    const BvhWNode* bvh_wnodes;   // non-null (only together with bvh_qnodes): shadow queries walk these
};

// Ray queue of one wavefront level, structure of arrays (one allocation, capacity `cap`)
struct RayQueue {
    double *ox, *oy, *oz, *dx, *dy, *dz;  // origin, direction
    double *tr, *tg, *tb;                 // path throughput (product of the weights down the recursion)
    double* sig;                          // `significance` argument of src/raytrace.rs:261
    uint64_t* key;                        // counter-RNG node key
    uint32_t* pixel;                      // shard-local pixel index (local_y * width + x)
    uint32_t* meta;                       // bits 0-7 depth, bit 8: record primary id
    int32_t* hit_id;                      // filled by the closest-hit kernel
    double* hit_t;
    uint32_t cap;
};
constexpr uint32_t META_DEPTH_MASK = 0xFFu;
constexpr uint32_t META_RECORD_ID = 0x100u;

// Shadow queue: origin, direction, squared range (< 0: no range => any hit occludes), pending contribution
struct ShadowQueue {
    double *ox, *oy, *oz, *dx, *dy, *dz;
    double* r2;
    double *cr, *cg, *cb;
    uint32_t* pixel;
    uint32_t cap;
};

// image-space shard + frame parameters
struct FrameParams {
    uint32_t width, height, antialias, cam_samples;
    uint32_t rank, world, band_rows;
    uint32_t local_rows;  // rows held by this shard, padded to whole bands
    uint32_t tiles_x;     // ceil(width / 8)
    uint32_t max_depth;
    double min_significance;
    uint64_t seed;
    uint64_t npix_padded;  // tiles_x*8 * local_rows: pixel slots per sample plane (tile order)
};

}  // namespace rt
