// host_scene.h -- host-side scene model behind the opaque `rt_scene` handle.
//
// Mirrors the reference's `Scene` (src/scene.rs:201-212) as flat tagged records instead of boxed trait objects.
// The host only *describes* the scene; all tracing happens on the device.
#pragma once
#include <cstdint>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "rt_types.h"

struct rt_render_opts;

namespace rt {

struct HostTexture {
    uint32_t width = 0, height = 0;
    std::vector<uint8_t> data;  // RGB8, row 0 = top
};

struct HostObject {
    bool is_plane;
    double p[3];     // sphere centre | plane point
    double n[3];     // plane normal
    double radius;   // sphere
    uint32_t material;
};

// Flattened scene: the reference's Vec<Object> of boxed trait objects (src/scene.rs:93-98) as tagged SoA buffers.
struct FlatScene {
    std::vector<SphereRec> sph;         // scan order = object order among spheres
    std::vector<int32_t> sph_obj;
    std::vector<PlaneRec> planes;
    std::vector<uint32_t> obj_shape, obj_material;
    double centroid_lo[3], centroid_hi[3];
    double box_lo[3], box_hi[3];  // bounds of the sphere boxes (c -+ |r|); -+inf when unusable
    double cmax, rmin;
    int32_t max_children;
    // 16-bit grid for the 32-byte BVH nodes (rt_types.h BvhQNode); quant == false: grid too coarse / unusable bounds
    bool quant;
    double q_org[3], q_cell[3];
};
// What rt_context_create needs from the host scene, kept with the scene between contexts: a caller that builds a context
// per frame (the reference's main does the equivalent) flattens a 1M-object scene once, not per frame, and uploads it
// from page-locked memory (`pin`, set by context.cu, undoes the registration when the cache entry goes away).
struct FlatCache {
    FlatScene flat;
    uint64_t revision = 0;
    uint32_t node_tuning = 0;  // RT_TUNE_NODES_* bits the grid decision was taken under
    std::shared_ptr<void> pin;
};
// the scene's slot for it; copying a scene does not copy the cache
struct FlatCacheSlot {
    std::mutex m;
    std::shared_ptr<FlatCache> p;
    FlatCacheSlot() = default;
    FlatCacheSlot(const FlatCacheSlot&) {}
    FlatCacheSlot& operator=(const FlatCacheSlot&) { p.reset(); return *this; }
};

// Device copies of the scene (flattened buffers + BVH), one per (device, accel, tuning), kept with the scene so that a
// caller rendering frame after frame through rt_render -- a context per call, as the reference's main would -- uploads and
// builds once.  Entries are owned by context.cu (type-erased here: this header stays free of CUDA); scene_touch drops them.
struct DeviceSceneSlot {
    struct Entry {
        int device;
        int accel;
        uint32_t tuning;
        uint64_t revision;
        std::shared_ptr<void> holder;
    };
    std::mutex m;
    std::vector<Entry> entries;
    DeviceSceneSlot() = default;
    DeviceSceneSlot(const DeviceSceneSlot&) {}
    DeviceSceneSlot& operator=(const DeviceSceneSlot&) { entries.clear(); return *this; }
};

}  // namespace rt

// the C handle
struct rt_scene {
    std::vector<rt::HostObject> objects;
    std::vector<rt::MaterialRec> materials;
    std::vector<rt::LightRec> lights;
    rt::CameraRec camera{};
    bool camera_set = false;
    int32_t bg_kind = rt::BG_SOLID;
    double bg_color[3] = {0.0, 0.0, 0.0};
    bool background_set = false;
    rt::HostTexture faces[6];
    uint32_t width = 0, height = 0, antialias = 1;
    bool options_set = false;
    std::string warnings;
    uint64_t n_spheres = 0, n_planes = 0;
    uint64_t revision = 0;  // bumped by rt::scene_touch
    mutable rt::FlatCacheSlot flat_cache;
    mutable rt::DeviceSceneSlot device_cache;
};

namespace rt {

void set_error(const std::string& msg);
const char* get_error();

// camera constructors (src/camera.rs:51-73, :97-106)
void camera_spc_new(CameraRec* cam, const double* position, const double* look, const double* up, double im_dist);
void camera_spc_look_at(CameraRec* cam, const double* focus, const double* look, const double* up, double pov, double h);
void camera_make_dof(CameraRec* cam, double focus, double aperture, uint32_t samples);

// texture file loaders (24-bit BMP, binary PPM); returns false and sets `err` on failure
bool load_texture_file(const std::string& path, HostTexture* out, std::string* err);

// serialize::deserialize; returns RT_OK or RT_ERR_SYNTAX with `err` = "row:col: message"
int parse_scene_text(const char* text, size_t len, rt_scene* scene, std::string* err);

// Called BEFORE anything changes the objects, materials or skybox faces of a scene: drops the cache entry (and with it the
// page-lock on the vectors about to be reallocated) and bumps the revision.
void scene_touch(rt_scene* s);
// set by context.cu: called by rt_scene_free before the scene goes away (the library keeps multi-GPU render state of the
// last rt_render_multi calls, which refers to the scene it was made for)
extern void (*g_scene_free_hook)(const rt_scene*);
// called AFTER the positions / radii of sphere objects changed and nothing else: refreshes the cached flat form in place
void scene_spheres_moved(rt_scene* s);
// returns RT_OK or RT_ERR_INVALID (missing camera/background/skybox face)
int flatten_scene(const rt_scene* s, FlatScene* out, uint32_t tuning = 0);
// the same through the scene's cache: flattens on the first call after a change, validates every time
int flatten_scene_cached(const rt_scene* s, uint32_t tuning, std::shared_ptr<FlatCache>* out);
// fills the fields of DeviceScene that do not point into device memory
void fill_scene_header(const rt_scene* s, const FlatScene& f, DeviceScene* ds);
// image-space shard + frame parameters from the scene's Options and the render options
int make_frame_params(const rt_scene* s, const struct ::rt_render_opts* o, FrameParams* fp);
bool accel_uses_bvh(int accel, uint32_t n_spheres);

// sRGB tables (generated; see tools/gen_srgb_tables.py)
extern const double SRGB_VALUES[256];
extern const double SRGB_AVERAGE[255];

}  // namespace rt
