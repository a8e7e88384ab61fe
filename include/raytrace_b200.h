/* raytrace_b200.h -- C ABI of the B200-native tracing path (libraytrace_b200.so).
 *
 * This is the drop-in boundary for the per-pixel tracing hot path of j-dong/rust-raytrace.  The upstream crate
 * has no FFI; its boundary is the public Rust API of `lib.rs` (src/lib.rs:4-13) and the call sequence of the
 * `raytrace` binary (src/main.rs:13-60).  Every entry point below names the upstream item it replaces.
 * INTEGRATION.md shows the `extern "C"` block a maintainer of the Rust crate would add to bind these.
 *
 * Conventions
 *   - plain C types only; all vectors/colours are `const double[3]` (x,y,z / r,g,b);
 *   - every function that can fail returns 0 on success or a negative rt_status; rt_last_error() returns a
 *     thread-local message for the last failure on the calling thread;
 *   - a scene (rt_scene) is host-side, mutable until it is handed to rt_context_create / rt_render; the library
 *     copies every array it is given (caller keeps ownership of its inputs);
 *   - a context (rt_context) owns the device-resident copy of one scene on one GPU (flattened primitives,
 *     device-built BVH, ray queues, accumulators); one render at a time per context;
 *   - there is NO CPU fallback: without a CUDA device every render/trace entry point fails with RT_ERR_CUDA.
 */
#ifndef RAYTRACE_B200_H
#define RAYTRACE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_API __attribute__((visibility("default")))

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_INVALID = -1, /* bad argument / scene incomplete */
    RT_ERR_SYNTAX = -2,  /* rt_scene_parse: message is the reference's `row:col: msg` */
    RT_ERR_IO = -3,      /* file open/read/write */
    RT_ERR_CUDA = -4,    /* no device, allocation or launch failure */
    RT_ERR_UNSUPPORTED = -5
} rt_status;

typedef struct rt_scene rt_scene;
typedef struct rt_context rt_context;

/* ---------------------------------------------------------------------------------------------------------
 * library
 * ------------------------------------------------------------------------------------------------------- */
RT_API const char* rt_last_error(void);
RT_API const char* rt_version(void);
/* number of visible CUDA devices (0 when there is none; never an error) */
RT_API int rt_device_count(void);
/* Device blocks released by destroyed contexts are cached per device and reused by the next context (rt_render builds
 * one per frame); this returns the cached blocks to the driver. */
RT_API void rt_release_cached_memory(void);

/* ---------------------------------------------------------------------------------------------------------
 * scene construction  (replaces the `pub` structs of src/scene.rs:32-212, src/shapes.rs:43-98)
 * ------------------------------------------------------------------------------------------------------- */
RT_API rt_scene* rt_scene_new(void);
RT_API void rt_scene_free(rt_scene* scene);

/* serialize::deserialize (src/serialize.rs:427-441).  On RT_ERR_SYNTAX `err` receives "row:col: message" exactly
 * as upstream's `impl Display for SyntaxError` (src/serialize.rs:268-272).  Warnings that upstream println!s
 * (src/serialize.rs:452,455,464) are appended to the scene's warning log, see rt_scene_warnings. */
RT_API int rt_scene_parse(const char* text, size_t len, rt_scene** out, char* err, size_t errcap);
RT_API int rt_scene_parse_file(const char* path, rt_scene** out, char* err, size_t errcap);
RT_API const char* rt_scene_warnings(const rt_scene* scene);

/* materials (src/scene.rs:32-89); each returns the new material's index (>= 0) or a negative rt_status */
RT_API int rt_scene_add_material_phong(rt_scene*, const double diffuse[3], const double specular[3], double exponent,
                                       const double ambient[3]);
RT_API int rt_scene_add_material_indirect(rt_scene*, const double diffuse[3], const double specular[3], double exponent,
                                          const double ambient[3], uint32_t samples);
RT_API int rt_scene_add_material_fresnel(rt_scene*, const double diffuse[3], const double specular[3], double exponent,
                                         const double ambient[3], double ior);
RT_API int rt_scene_add_material_transparent(rt_scene*, const double specular[3], double exponent, double ior);

/* objects (src/scene.rs:93-98, src/shapes.rs:43-48, :93-98).  Object index = order of insertion, as in
 * `Scene.objects`; it is the id reported for primary hits and the tie-break of the closest-hit search. */
RT_API int rt_scene_add_spheres(rt_scene*, uint64_t n, const double* center_xyz_radius /* 4n */, const uint32_t* material /* n */);
RT_API int rt_scene_add_plane(rt_scene*, const double point[3], const double normal[3], uint32_t material);
/* `scene.objects[i].bounds = Box::new(Sphere { .. })` for n consecutive sphere objects starting at object index
 * first_object (upstream's fields are pub: src/scene.rs:93-98, src/shapes.rs:43-48): an animation moves its spheres
 * between frames.  Materials and object order stay; the next context re-uploads the scene and rebuilds the BVH. */
RT_API int rt_scene_update_spheres(rt_scene*, uint64_t first_object, uint64_t n, const double* center_xyz_radius /* 4n */);

/* lights (src/scene.rs:109-155) */
RT_API int rt_scene_add_light_point(rt_scene*, const double location[3], const double color[3]);
RT_API int rt_scene_add_light_directional(rt_scene*, const double direction[3], const double color[3]);
RT_API int rt_scene_add_light_area(rt_scene*, const double origin[3], const double side1[3], const double side2[3],
                                   const double color[3]);

/* cameras: SimplePerspectiveCamera::new (src/camera.rs:51-63), ::look_at (:67-73), DepthOfFieldCamera::new (:97-106).
 * rt_scene_set_camera_dof wraps the perspective camera that is currently set. */
RT_API int rt_scene_set_camera_spc_new(rt_scene*, const double position[3], const double look[3], const double up[3], double im_dist);
RT_API int rt_scene_set_camera_spc_look_at(rt_scene*, const double focus[3], const double look[3], const double up[3], double pov_rad,
                                           double h);
RT_API int rt_scene_set_camera_dof(rt_scene*, double focus, double aperture, uint32_t samples);
RT_API int rt_scene_get_camera(const rt_scene*, double position[3], double matrix_rowmajor[9]);

/* backgrounds (src/scene.rs:165-188).  Skybox faces are RGB8, row 0 = top row of the decoded image
 * (what `image::open(..).to_rgb().into_raw()` yields, src/texture.rs:34-37); face order px nx py ny pz nz. */
RT_API int rt_scene_set_background_solid(rt_scene*, const double color[3]);
RT_API int rt_scene_set_skybox_face(rt_scene*, int face, uint32_t width, uint32_t height, const uint8_t* rgb);
/* Texture::load (src/texture.rs:34-37) for the formats decodable without the `image` crate: 24-bit BMP, binary PPM */
RT_API int rt_scene_load_skybox_face(rt_scene*, int face, const char* path);

/* Options (src/scene.rs:191-198) */
RT_API int rt_scene_set_options(rt_scene*, uint32_t width, uint32_t height, uint32_t antialias);
RT_API int rt_scene_get_options(const rt_scene*, uint32_t* width, uint32_t* height, uint32_t* antialias);

/* introspection (parser tests): counts and a canonical text dump of everything the scene holds */
typedef struct rt_scene_info {
    uint64_t n_objects, n_spheres, n_planes, n_materials, n_lights;
    int32_t camera_kind;     /* 0 = SimplePerspectiveCamera, 1 = DepthOfFieldCamera */
    int32_t background_kind; /* 0 = solid, 1 = skybox */
} rt_scene_info;
RT_API int rt_scene_get_info(const rt_scene*, rt_scene_info* info);
/* writes a NUL-terminated dump (hex-float numbers, one item per line) of at most cap-1 bytes; returns the full length */
RT_API size_t rt_scene_dump(const rt_scene*, char* buf, size_t cap);

/* ---------------------------------------------------------------------------------------------------------
 * rendering  (replaces src/main.rs:39-59 and everything below it: src/raytrace.rs, Scene::intersect, ...)
 * ------------------------------------------------------------------------------------------------------- */
enum { RT_ACCEL_AUTO = 0, RT_ACCEL_LINEAR = 1, RT_ACCEL_BVH = 2 };
enum { RT_FP_STRICT = 0, RT_FP_FAST = 1 };

typedef struct rt_render_opts {
    uint32_t width, height;    /* 0 = take from the scene's Options */
    uint32_t antialias;        /* 0xFFFFFFFF = take from the scene's Options (0 is a legal upstream value: NaN -> 255) */
    uint32_t max_depth;        /* upstream const MAX_DEPTH = 4 (src/raytrace.rs:18) */
    double min_significance;   /* upstream const MIN_SIGNIFICANCE = 1/512 (src/raytrace.rs:17) */
    uint64_t seed;             /* counter RNG seed (DESIGN.md "RNG") */
    int32_t accel;             /* RT_ACCEL_*: AUTO = BVH when the scene has >= 64 spheres */
    int32_t fp_mode;           /* RT_FP_STRICT: no FMA contraction (bit-parity with the reference's op order); RT_FP_FAST */
    int32_t device;            /* CUDA device ordinal */
    int32_t rank, world;       /* image-space shard: this context renders the row bands b with b % world == rank */
    uint32_t band_rows;        /* rows per band (multiple of 4); 0 = default (4) */
    uint64_t max_rays_in_flight; /* ray-queue capacity per wavefront level; 0 = default */
    uint32_t tuning;           /* RT_TUNE_* bits, 0 = the measured defaults.  They choose between formulations that give
                                  the SAME answers (tested); they exist for A/B measurements and the parity tests */
    uint32_t _reserved;        /* must be 0 */
} rt_render_opts;
enum {
    RT_TUNE_NODES_F32 = 1u,       /* BVH: keep the 64-byte f32 nodes even where the 16-bit grid is fine enough */
    RT_TUNE_NODES_Q16 = 2u,       /* BVH: force the 32-byte grid nodes wherever the bounds allow a grid at all */
    RT_TUNE_MORTON_AXIS = 4u,     /* BVH build: scale the Morton grid per axis instead of by the longest extent */
    RT_TUNE_BINARY_SHADOW = 8u,   /* shadow queries walk the binary nodes instead of the four-wide ones */
    RT_TUNE_HOST_LEVELS = 16u,    /* the host reads every wavefront level's ray count back before it launches the next
                                     level (no device-driven tail); results identical */
    RT_TUNE_SERIAL_SHADOW = 32u,  /* device-driven frames: trace the shadow queue in one launch at the end of the frame instead
                                     of level by level on a second stream beside the level loop */
    RT_TUNE_FRAME_GRAPH = 64u     /* device-driven frames: replay the frame's CUDA graph (captured once a context queues its second
                                     frame from the same arguments) instead of queueing every launch: 0.01 instead of 0.30 ms of
                                     host time per frame, but slower on the device for small shards */
};
RT_API void rt_render_opts_default(rt_render_opts* opts);

typedef struct rt_stats {
    uint64_t rays_total;      /* closest-hit + shadow queries = Scene::intersect calls upstream (src/scene.rs:247) */
    uint64_t rays_primary, rays_secondary, rays_shadow;
    uint64_t paths;           /* pixel samples x camera samples traced by this context */
    uint64_t kernel_launches; /* kernels of this library launched by the last render */
    double render_ms;         /* device time of the last render (CUDA events on the context's stream) */
    double bvh_build_ms;      /* device time of the BVH build at context creation */
    uint64_t n_bvh_nodes;
    uint32_t local_rows;      /* rows this shard produces (padded to whole bands) */
    uint32_t bytewidth;       /* bytes per output row = (3*width + 3) & ~3 (src/bmp.rs:11) */
    int32_t accel_used, fp_mode_used;
    /* BVH traversal work of the last render (device counters): 64-byte nodes fetched and exact sphere tests at leaves,
     * [0] by the closest-hit kernel, [1] by the shadow kernel */
    uint64_t bvh_nodes_visited[2], bvh_prims_tested[2];
    /* per-kernel-class device time of the last render; filled only while rt_context_set_profiling(ctx, 1) is on
     * (one CUDA event pair per launch).  Order: primary, closest, shade, shadow, resolve. */
    double class_ms[5];
    uint64_t class_launches[5];
    /* rays the BVH traversal handed to the block-wide exact scan (runaway walks of rays that start 10^7+ units out):
     * [0] closest-hit, [1] shadow */
    uint64_t heavy_rays[2];
    /* bytes per BVH node in use: 32 (box planes on the scene's 16-bit grid), 64 (f32 planes; grid too coarse for the scene) or 0 */
    uint64_t bvh_node_bytes;
    /* 1: the last frame was driven level by level from the host (RT_TUNE_HOST_LEVELS, or a queue overflowed in the
     * device-driven frame, which is then redone); device_driven_overflows counts those redone frames of this context */
    uint32_t host_levels, device_driven_overflows;
    /* exact sphere tests made by the linear-scan kernels ([0] closest-hit, [1] shadow: a shadow query stops at its first
     * occluder); with 25 flop per test (src/shapes.rs:60-65) the FP64 work their roofline fraction is computed from */
    uint64_t scan_sphere_tests[2];
} rt_stats;

/* One-call frame render: upload, (build BVH), trace, resolve, copy back.
 *   bgr_rows     host buffer of bytewidth*height bytes; row 0 is y = 0, i.e. BMP bottom-up order (src/main.rs:45-58)
 *   primary_ids  optional host buffer of width*height int32: object index hit by the first sample's primary ray, -1 = miss
 *   accum        optional host buffer of width*height*3 doubles: per-pixel SUM of the sample colours in linear RGB;
 *                the value handed to the quantiser (src/main.rs:56) is (accum / camera_samples) / antialias
 * Requires world == 1 in opts (rt_render_multi shards one frame over several GPUs).
 * The device copy of the scene (flattened buffers + BVH) stays with the rt_scene between calls, per device, until a
 * builder call changes the scene: a caller rendering frame after frame uploads and builds once.  The call's context
 * (streams, events, ray queues) is kept too, keyed by (scene, options), and re-bound to the scene's current revision by the
 * next call; rt_release_cached_memory and rt_scene_free drop it. */
RT_API int rt_render(const rt_scene*, const rt_render_opts*, uint8_t* bgr_rows, int32_t* primary_ids, double* accum, rt_stats* stats);

/* Device-resident path (bench / shards): the scene and all buffers stay in HBM between frames.
 * The first rt_context_create (or rt_render, which calls it) for a scene on a device uploads the scene and builds the BVH
 * there; that device copy, and the host-side flattened (structure-of-arrays) form it is uploaded from (page-locked in
 * place for scenes above 4 MB), stay with the rt_scene for the next context.  Any builder call that changes objects,
 * materials, lights or skybox faces drops both (stats.bvh_build_ms is 0 when the kept copy was reused).  Contexts may be
 * created from one scene concurrently, but not while another thread is changing that scene. */
RT_API int rt_context_create(const rt_scene*, const rt_render_opts*, rt_context** out);
RT_API void rt_context_destroy(rt_context*);
/* traces the shard and leaves quantised BGR rows (local_rows x bytewidth) in the context's device framebuffer */
RT_API int rt_context_render(rt_context*, rt_stats* stats);
/* The same frame in two halves, for a caller that keeps several frames in flight (one context each: upstream's frame loop,
 * src/main.rs:39-59, run for an animation): _begin queues every launch of the frame on the context's streams and returns
 * while the GPU works (a context whose scene needs host-driven levels traces the frame before it returns); _end waits for
 * the frame, reads the counts, retraces the frame level by level if a queue overflowed, and fills `stats`.  Exactly one
 * _end per _begin; nothing else may be called on the context in between.  On a shard the deep wavefront levels are bound by
 * the latency of their longest walks, not by throughput: a second frame in flight fills those gaps. */
RT_API int rt_context_render_begin(rt_context*);
RT_API int rt_context_render_end(rt_context*, rt_stats* stats);
RT_API int rt_context_stats(const rt_context*, rt_stats* stats);
/* device pointer of the local framebuffer (local_rows*bytewidth bytes), e.g. to wrap as a tensor for the NCCL gather */
RT_API void* rt_context_framebuffer(rt_context*);
RT_API size_t rt_context_framebuffer_bytes(const rt_context*);
/* copies the local framebuffer / ids / accumulators to host memory (shard-local row order) */
RT_API int rt_context_download(rt_context*, uint8_t* bgr_rows, int32_t* primary_ids, double* accum);
/* After an all-gather of every rank's local framebuffer (rank-major, world*local_rows*bytewidth bytes, device memory)
 * writes the final bytewidth*height image in BMP row order into `out_dev`.  Runs on the context's device/stream. */
RT_API int rt_context_untile(rt_context*, const void* gathered_dev, void* out_dev);
RT_API int rt_context_synchronize(rt_context*);
/* Whole-frame output.  By default a context resolves into its own shard-sized framebuffer (local row order).  With a frame
 * set, k_resolve stores every row of the shard at its place in a whole-frame buffer (bytewidth*height bytes, BMP row
 * order) instead -- the gather of a sharded frame then IS the resolve kernel's stores, over NVLink when the buffer lives
 * on another GPU.  One process per GPU (torchrun): rank 0 creates the frame and exports it, the others import the handle:
 *   rt_context_frame_create   allocate the frame on this context's device and resolve into it; *dev_ptr = its address
 *   rt_context_frame_export   64-byte CUDA IPC handle of the frame created above
 *   rt_context_frame_import   open another process's frame (peer access is enabled by the driver) and resolve into it
 *   rt_context_frame_set      resolve into any device address this GPU can store to (own or peer memory of this
 *                             process); NULL = back to the context's own framebuffer
 * A frame is complete when every shard's rt_context_render has returned (the caller synchronises the ranks). */
RT_API int rt_context_frame_create(rt_context*, void** dev_ptr);
RT_API int rt_context_frame_export(rt_context*, uint8_t handle64[64]);
RT_API int rt_context_frame_import(rt_context*, const uint8_t handle64[64]);
RT_API int rt_context_frame_set(rt_context*, void* dev_ptr);
/* run on a caller-owned CUDA stream (a cudaStream_t / CUstream handle, e.g. torch's current stream) instead of the
 * context's own; pass NULL to go back.  The caller keeps the stream alive for as long as it is set. */
RT_API int rt_context_set_stream(rt_context*, void* cuda_stream);
RT_API int rt_context_set_profiling(rt_context*, int on);

/* ---------------------------------------------------------------------------------------------------------
 * multi-GPU frames  (the same frame loop, src/main.rs:39-59, sharded over the GPUs of one node)
 *
 * The image is cut into bands of `band_rows` rows, band b rendered by device[b % n_devices] (scene replicated; with the
 * counter RNG every pixel is independent, so tracing exchanges nothing).  One host thread and stream per device inside
 * the library.  How the quantised BGR bands become one frame:
 *   RT_GATHER_DIRECT  each device DMAs its bands to their rows of the host buffer (no device ever holds the frame)
 *   RT_GATHER_PEER    each device's resolve kernel stores its rows into device[0]'s frame over NVLink peer access
 *   RT_GATHER_NCCL    ncclAllGather of the shards + de-interleave on every device (libnccl.so.2 loaded at run time);
 *                     every device ends up with the whole frame
 *   RT_GATHER_AUTO    = DIRECT (the destination of rt_multi_render is host memory)
 * opts->device / rank / world are ignored (set per device by the library).
 * ------------------------------------------------------------------------------------------------------- */
typedef struct rt_multi rt_multi;
enum { RT_GATHER_AUTO = 0, RT_GATHER_DIRECT = 1, RT_GATHER_PEER = 2, RT_GATHER_NCCL = 3 };
/* devices == NULL: ordinals 0 .. n_devices-1.  Uploads the scene and builds the BVH on every device (in parallel).
 * DIRECT / PEER accept an ordinal more than once (those shards share the GPU); NCCL needs distinct devices. */
RT_API int rt_multi_create(const rt_scene*, const rt_render_opts*, int n_devices, const int* devices, int gather, rt_multi** out);
RT_API void rt_multi_destroy(rt_multi*);
/* One frame.  bgr_rows: host buffer of bytewidth*height bytes (BMP row order) or NULL (PEER / NCCL: the frame stays on
 * the device, see rt_multi_frame_device); primary_ids: optional host buffer of width*height int32.
 * stats: sums over the devices; render_ms, bvh_build_ms and class_ms are the maximum over the devices. */
RT_API int rt_multi_render(rt_multi*, uint8_t* bgr_rows, int32_t* primary_ids, rt_stats* stats);
RT_API int rt_multi_device_stats(const rt_multi*, int index, rt_stats* stats);
/* device pointer of the assembled frame on device[index]: PEER index 0, NCCL every index; else NULL */
RT_API const void* rt_multi_frame_device(const rt_multi*, int index);
RT_API int rt_multi_gather_used(const rt_multi*);
/* rt_render for several GPUs: create, one frame, destroy.  Device memory, device scenes and NCCL communicators are
 * cached between calls (rt_release_cached_memory). */
RT_API int rt_render_multi(const rt_scene*, const rt_render_opts*, int n_devices, const int* devices, int gather, uint8_t* bgr_rows,
                           int32_t* primary_ids, rt_stats* stats);

/* Scene::intersect for a batch of rays (src/scene.rs:247-249): origin xyz + direction xyz per ray.
 * ids: object index or -1; t: hit distance (+inf on miss); normal: shape normal as upstream returns it. */
RT_API int rt_trace_rays(const rt_scene*, const rt_render_opts*, uint64_t n, const double* origin_dir, int32_t* ids, double* t,
                         double* normal);
/* Scene::intersect over a list of spheres (object i = sphere i) for callers that may not allocate device memory.
 * NO COUNTERPART IN THE REFERENCE: added for code under CUDA graph capture, tight
 * serving loops and harnesses that forbid cudaMalloc.  A variant of the linear-scan closest-hit kernel whose scratch (packed
 * sphere records, work cursor) is a static __device__ array of RT_NOALLOC_SCRATCH_BYTES that every GPU which loads the
 * library holds until exit.  All four pointers are DEVICE memory on `device`: spheres as (x, y, z, radius) (16-byte
 * aligned), rays as origin xyz + direction xyz, ids (object index or -1), dev_t (hit distance, +inf on a miss; may be
 * NULL).  Results are bit-identical to rt_trace_rays in RT_FP_STRICT on a scene holding just these spheres.
 * Calls cudaMalloc / cudaFree / stream or event creation never, and does not synchronise: two kernels are queued on
 * `cuda_stream` (a cudaStream_t; NULL = the default stream).  The first call on a device looks up the array's address
 * (cudaGetSymbolAddress): make one call before capturing a graph.  More than RT_NOALLOC_MAX_SPHERES spheres:
 * RT_ERR_INVALID, nothing is launched.  Every call on a device shares the array (each call rewrites what it reads and
 * zeroes its own counters in stream order), so calls that overlap on the GPU race: keep them on one stream. */
enum { RT_NOALLOC_MAX_SPHERES = 65536, RT_NOALLOC_SCRATCH_BYTES = 2097408 /* 65536 * 32 + 256 */ };
RT_API int rt_intersect_spheres_noalloc(int device, void* cuda_stream, const double* dev_center_xyz_radius, uint64_t n_spheres,
                                        const double* dev_origin_dir, uint64_t n_rays, int32_t* dev_ids, double* dev_t);
/* Camera::project for a batch of normalised image positions (src/camera.rs:77-79, :110-121); ray i uses the
 * counter-RNG key of (seed, pixel i, sample 0, camera sample 0). */
RT_API int rt_project(const rt_scene*, const rt_render_opts*, uint64_t n, const double* pos_xy, double* origin_dir);
/* Color::write_bgr over n colours (src/color.rs:593-600, :628-632), on the device */
RT_API int rt_quantize_bgr(const double* rgb, uint64_t n, uint8_t* bgr, int device);

/* Roofline denominators MEASURED_PEAKS.json does not carry, measured on `device` with register-resident
 * FMA chains and an L2-resident read loop (~50 ms): the FP64 pipe bounds the linear-scan kernels, L2 serves the BVH
 * kernels' node and leaf fetches.  A measurement hook: nothing on the tracing path uses it. */
typedef struct rt_peaks {
    double fp64_tflops;   /* DFMA, 2 flop each */
    double fp32_tflops;   /* FFMA */
    double l2_read_gbs;   /* ld.global.cg over a buffer of l2_buffer_bytes (a quarter of the L2) */
    uint64_t l2_buffer_bytes;
    int32_t sm_count, _pad;
} rt_peaks;
RT_API int rt_measure_peaks(int device, rt_peaks* out);

/* ---------------------------------------------------------------------------------------------------------
 * BMP sink  (src/bmp.rs:10-61 + the row writes of src/main.rs:42,58)
 * ------------------------------------------------------------------------------------------------------- */
RT_API uint32_t rt_bmp_bytewidth(uint32_t width);
RT_API void rt_bmp_header(uint32_t width, uint32_t height, uint8_t out122[122]);
RT_API int rt_bmp_write(const char* path, uint32_t width, uint32_t height, const uint8_t* rows);

#ifdef __cplusplus
}
#endif
#endif /* RAYTRACE_B200_H */
