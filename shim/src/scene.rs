//! Scene description (/root/reference/src/scene.rs): materials, objects, lights, backgrounds, options, `Scene`.
//!
//! Upstream's fields are `pub` and its polymorphism is `Box<Trait>`; here the same names are plain data and closed enums
//! (the device evaluates a tagged union per material / light / background, not a vtable).  A `Scene` owns an `rt_scene`
//! handle that mirrors its vectors; `Scene::sync` rebuilds the handle after the vectors were edited in place.

use std::ffi::CString;
use std::ptr;

use crate::camera::*;
use crate::color::Color;
use crate::ffi;
use crate::shapes::*;
#[cfg(feature = "skybox")]
use crate::texture::TextureSource;
use crate::types::*;
use crate::{check, Error};

/// src/scene.rs:32-44
#[derive(Clone, Copy, Debug)]
pub struct PhongMaterial { pub diffuse: Color, pub specular: Color, pub exponent: f64, pub ambient: Color }
/// src/scene.rs:48-61
#[derive(Clone, Copy, Debug)]
pub struct FresnelMaterial { pub diffuse: Color, pub specular: Color, pub exponent: f64, pub ambient: Color, pub ior: f64 }
/// src/scene.rs:65-73
#[derive(Clone, Copy, Debug)]
pub struct TransparentMaterial { pub specular: Color, pub exponent: f64, pub ior: f64 }
/// src/scene.rs:76-89
#[derive(Clone, Copy, Debug)]
pub struct IndirectPhongMaterial { pub diffuse: Color, pub specular: Color, pub exponent: f64, pub ambient: Color, pub samples: u32 }

/// Upstream's `Box<Material>` (src/scene.rs:24-29); `Material::color` is the device's shade kernel.
#[derive(Clone, Copy, Debug)]
pub enum Material {
    Phong(PhongMaterial),
    Fresnel(FresnelMaterial),
    Transparent(TransparentMaterial),
    IndirectPhong(IndirectPhongMaterial),
}

/// src/scene.rs:93-98
#[derive(Clone, Copy, Debug)]
pub struct Object { pub bounds: Shape, pub material: Material }

/// src/scene.rs:117-120
#[derive(Clone, Copy, Debug)]
pub struct PointLight { pub location: Pnt3 }
/// src/scene.rs:130-133 (the direction is used as given, un-normalised: src/scene.rs:137)
#[derive(Clone, Copy, Debug)]
pub struct DirectionalLight { pub direction: Vec3 }
/// src/scene.rs:142-149
#[derive(Clone, Copy, Debug)]
pub struct AreaLight { pub origin: Pnt3, pub side1: Vec3, pub side2: Vec3 }

/// Upstream's `Box<LightModel>` (src/scene.rs:101-106)
#[derive(Clone, Copy, Debug)]
pub enum LightModel { Point(PointLight), Directional(DirectionalLight), Area(AreaLight) }

/// src/scene.rs:109-114
#[derive(Clone, Copy, Debug)]
pub struct Light { pub model: LightModel, pub color: Color }

/// src/scene.rs:165-168
#[derive(Clone, Copy, Debug)]
pub struct SolidColorBackground { pub color: Color }

/// src/scene.rs:175-188: six faces, positive / negative X, Y, Z
#[cfg(feature = "skybox")]
#[derive(Clone, Debug)]
pub struct SkyboxBackground {
    pub px: TextureSource, pub nx: TextureSource, pub py: TextureSource, pub ny: TextureSource, pub pz: TextureSource, pub nz: TextureSource,
}

/// Upstream's `Box<Background>` (src/scene.rs:159-162)
#[derive(Clone, Debug)]
pub enum Background {
    SolidColor(SolidColorBackground),
    #[cfg(feature = "skybox")]
    Skybox(SkyboxBackground),
}

/// src/scene.rs:191-198
#[derive(Clone, Copy, Debug)]
pub struct Options { pub width: u32, pub height: u32, pub antialias: u32 }

/// src/scene.rs:215-220
#[derive(Clone, Copy, Debug)]
pub struct SceneIntersectionResult<'a> {
    pub object: &'a Object,
    pub result: IntersectionResult,
}

/// src/scene.rs:201-212
pub struct Scene {
    pub objects: Vec<Object>,
    pub lights: Vec<Light>,
    pub camera: Camera,
    pub background: Background,
    pub options: Options,
    handle: *mut ffi::RtScene,
    warnings: String,
}

// the handle is only read by renders (include/raytrace_b200.h: "a scene handle is read-only during renders")
unsafe impl Send for Scene {}

impl Drop for Scene {
    fn drop(&mut self) {
        if !self.handle.is_null() { unsafe { ffi::rt_scene_free(self.handle) } }
    }
}

fn add_material(h: *mut ffi::RtScene, m: &Material) -> Result<u32, Error> {
    let rc = unsafe {
        match *m {
            Material::Phong(p) => ffi::rt_scene_add_material_phong(h, p.diffuse.as_array().as_ptr(), p.specular.as_array().as_ptr(), p.exponent, p.ambient.as_array().as_ptr()),
            Material::Fresnel(p) => ffi::rt_scene_add_material_fresnel(h, p.diffuse.as_array().as_ptr(), p.specular.as_array().as_ptr(), p.exponent, p.ambient.as_array().as_ptr(), p.ior),
            Material::Transparent(p) => ffi::rt_scene_add_material_transparent(h, p.specular.as_array().as_ptr(), p.exponent, p.ior),
            Material::IndirectPhong(p) => ffi::rt_scene_add_material_indirect(h, p.diffuse.as_array().as_ptr(), p.specular.as_array().as_ptr(), p.exponent, p.ambient.as_array().as_ptr(), p.samples),
        }
    };
    Ok(check(rc)? as u32)
}

impl Scene {
    /// Builds the library-side scene from the vectors.  Object order = `objects` order: it is the id reported for primary
    /// hits and the tie-break of the closest-hit search (src/scene.rs:248).
    pub fn new(objects: Vec<Object>, lights: Vec<Light>, camera: Camera, background: Background, options: Options) -> Result<Scene, Error> {
        let mut s = Scene { objects: objects, lights: lights, camera: camera, background: background, options: options, handle: ptr::null_mut(), warnings: String::new() };
        s.sync()?;
        Ok(s)
    }

    /// Wraps a handle the library built itself (`serialize::deserialize`): the Rust-side vectors stay empty.
    pub(crate) fn from_handle(handle: *mut ffi::RtScene, options: Options, warnings: String) -> Scene {
        Scene { objects: Vec::new(), lights: Vec::new(), camera: Camera::SimplePerspective(SimplePerspectiveCamera::new(&Pnt3::new(0.0, 0.0, 0.0), &Vec3::new(0.0, 0.0, -1.0), &Vec3::new(0.0, 1.0, 0.0), 1.0)),
                background: Background::SolidColor(SolidColorBackground { color: crate::color::BLACK }), options: options, handle: handle, warnings: warnings }
    }

    pub fn handle(&self) -> *const ffi::RtScene { self.handle }
    /// What upstream `println!`s while parsing (src/serialize.rs:452,455,464)
    pub fn warnings(&self) -> &str { &self.warnings }

    /// Rebuilds the library-side scene after `objects`, `lights`, `camera`, `background` or `options` were edited in place.
    pub fn sync(&mut self) -> Result<(), Error> {
        let h = unsafe { ffi::rt_scene_new() };
        // consecutive spheres go over in one call (1M spheres = one FFI call, not a million)
        let mut batch: Vec<f64> = Vec::new();
        let mut batch_mat: Vec<u32> = Vec::new();
        let flush = |batch: &mut Vec<f64>, batch_mat: &mut Vec<u32>| -> Result<(), Error> {
            if !batch_mat.is_empty() {
                check(unsafe { ffi::rt_scene_add_spheres(h, batch_mat.len() as u64, batch.as_ptr(), batch_mat.as_ptr()) })?;
                batch.clear();
                batch_mat.clear();
            }
            Ok(())
        };
        let result = (|| -> Result<(), Error> {
            for o in &self.objects {
                let m = add_material(h, &o.material)?;
                match o.bounds {
                    Shape::Sphere(s) => { batch.extend_from_slice(&[s.center.x, s.center.y, s.center.z, s.radius]); batch_mat.push(m); }
                    Shape::Plane(p) => {
                        flush(&mut batch, &mut batch_mat)?;
                        check(unsafe { ffi::rt_scene_add_plane(h, p.point.as_array().as_ptr(), p.normal.as_array().as_ptr(), m) })?;
                    }
                }
            }
            flush(&mut batch, &mut batch_mat)?;
            for l in &self.lights {
                let c = l.color.as_array();
                check(unsafe {
                    match l.model {
                        LightModel::Point(p) => ffi::rt_scene_add_light_point(h, p.location.as_array().as_ptr(), c.as_ptr()),
                        LightModel::Directional(d) => ffi::rt_scene_add_light_directional(h, d.direction.as_array().as_ptr(), c.as_ptr()),
                        LightModel::Area(a) => ffi::rt_scene_add_light_area(h, a.origin.as_array().as_ptr(), a.side1.as_array().as_ptr(), a.side2.as_array().as_ptr(), c.as_ptr()),
                    }
                })?;
            }
            let spc = |c: &SimplePerspectiveCamera| -> i32 {
                unsafe {
                    match *c {
                        SimplePerspectiveCamera::New { position, look, up, im_dist } =>
                            ffi::rt_scene_set_camera_spc_new(h, position.as_array().as_ptr(), look.as_array().as_ptr(), up.as_array().as_ptr(), im_dist),
                        SimplePerspectiveCamera::LookAt { focus, look, up, pov, h: half } =>
                            ffi::rt_scene_set_camera_spc_look_at(h, focus.as_array().as_ptr(), look.as_array().as_ptr(), up.as_array().as_ptr(), pov, half),
                    }
                }
            };
            match self.camera {
                Camera::SimplePerspective(ref c) => { check(spc(c))?; }
                Camera::DepthOfField(ref d) => {
                    check(spc(&d.camera))?;
                    check(unsafe { ffi::rt_scene_set_camera_dof(h, d.focus, d.aperture, d.samples) })?;
                }
            }
            match self.background {
                Background::SolidColor(b) => { check(unsafe { ffi::rt_scene_set_background_solid(h, b.color.as_array().as_ptr()) })?; }
                #[cfg(feature = "skybox")]
                Background::Skybox(ref sky) => {
                    let faces = [&sky.px, &sky.nx, &sky.py, &sky.ny, &sky.pz, &sky.nz];
                    for (k, f) in faces.iter().enumerate() {
                        match **f {
                            TextureSource::Pixels(ref t) => { check(unsafe { ffi::rt_scene_set_skybox_face(h, k as i32, t.width, t.height, t.data.as_ptr()) })?; }
                            TextureSource::File(ref p) => {
                                let c = CString::new(p.to_string_lossy().into_owned()).map_err(|_| Error { code: ffi::RT_ERR_INVALID, message: "path contains NUL".to_string() })?;
                                check(unsafe { ffi::rt_scene_load_skybox_face(h, k as i32, c.as_ptr()) })?;
                            }
                        }
                    }
                }
            }
            check(unsafe { ffi::rt_scene_set_options(h, self.options.width, self.options.height, self.options.antialias) })?;
            Ok(())
        })();
        match result {
            Ok(()) => {
                if !self.handle.is_null() { unsafe { ffi::rt_scene_free(self.handle) } }
                self.handle = h;
                Ok(())
            }
            Err(e) => { unsafe { ffi::rt_scene_free(h) }; Err(e) }
        }
    }

    /// `Scene::intersect` (src/scene.rs:247-249) for one ray; `raytrace::intersect_batch` is the efficient form.
    /// `None` when nothing is hit.  Only for scenes built with `Scene::new` (the hit object is borrowed from `objects`).
    pub fn intersect(&self, ray: &Ray) -> Option<SceneIntersectionResult> {
        let hits = crate::raytrace::intersect_batch(self, &[*ray], &crate::raytrace::RenderOptions::default()).ok()?;
        let (id, t, n) = hits[0];
        if id < 0 { return None; }
        self.objects.get(id as usize).map(|o| SceneIntersectionResult { object: o, result: IntersectionResult { t: t, normal: n } })
    }
}
