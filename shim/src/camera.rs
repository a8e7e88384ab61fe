//! Cameras (/root/reference/src/camera.rs).  The constructors keep upstream's names and arguments; the camera matrix is
//! built by the library in the reference's operation order (`rt_scene_set_camera_*`), and `project` runs on the device
//! (`rt_project`), so a Rust caller sees exactly the rays the renderer traces.

use crate::types::*;

/// src/camera.rs:31-36 -- described by its constructor arguments; the library derives position + matrix
#[derive(Clone, Copy, Debug)]
pub enum SimplePerspectiveCamera {
    /// `SimplePerspectiveCamera::new(position, look, up, im_dist)` (src/camera.rs:51-63)
    New { position: Pnt3, look: Vec3, up: Vec3, im_dist: f64 },
    /// `SimplePerspectiveCamera::look_at(focus, look, up, pov, h)` (src/camera.rs:67-73)
    LookAt { focus: Pnt3, look: Vec3, up: Vec3, pov: f64, h: f64 },
}

impl SimplePerspectiveCamera {
    pub fn new(position: &Pnt3, look: &Vec3, up: &Vec3, im_dist: f64) -> SimplePerspectiveCamera {
        SimplePerspectiveCamera::New { position: *position, look: *look, up: *up, im_dist: im_dist }
    }
    pub fn look_at(focus: &Pnt3, look: &Vec3, up: &Vec3, pov: f64, h: f64) -> SimplePerspectiveCamera {
        SimplePerspectiveCamera::LookAt { focus: *focus, look: *look, up: *up, pov: pov, h: h }
    }
}

/// src/camera.rs:83-106
#[derive(Clone, Copy, Debug)]
pub struct DepthOfFieldCamera {
    pub camera: SimplePerspectiveCamera,
    /// The distance to the focal plane (from the camera).
    pub focus: f64,
    /// The aperture radius. The aperture is a circle.
    pub aperture: f64,
    /// The number of samples.
    pub samples: u32,
}

impl DepthOfFieldCamera {
    pub fn new(camera: SimplePerspectiveCamera, focus: f64, aperture: f64, samples: u32) -> DepthOfFieldCamera {
        DepthOfFieldCamera { camera: camera, focus: focus, aperture: aperture, samples: samples }
    }
}

/// Upstream's `Box<Camera>` trait object (src/camera.rs:19-27) as a closed enum.
#[derive(Clone, Copy, Debug)]
pub enum Camera {
    SimplePerspective(SimplePerspectiveCamera),
    DepthOfField(DepthOfFieldCamera),
}

impl Camera {
    /// `Camera::samples` (src/camera.rs:25-26, :122)
    pub fn samples(&self) -> u32 {
        match *self { Camera::SimplePerspective(_) => 1, Camera::DepthOfField(d) => d.samples }
    }
}

impl From<SimplePerspectiveCamera> for Camera { fn from(c: SimplePerspectiveCamera) -> Camera { Camera::SimplePerspective(c) } }
impl From<DepthOfFieldCamera> for Camera { fn from(c: DepthOfFieldCamera) -> Camera { Camera::DepthOfField(c) } }
