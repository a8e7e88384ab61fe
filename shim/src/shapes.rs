//! Shapes that rays can intersect (/root/reference/src/shapes.rs).  The intersection arithmetic (src/shapes.rs:50-112)
//! runs on the device in the reference's operation order; these types only describe the geometry.

use crate::types::*;

/// src/shapes.rs:10-19
#[derive(Clone, Copy, Debug)]
pub struct Ray { pub origin: Pnt3, pub direction: Vec3 }

impl Ray {
    /// src/shapes.rs:22-24
    pub fn cast(&self, t: f64) -> Pnt3 {
        Pnt3::new(self.origin.x + self.direction.x * t, self.origin.y + self.direction.y * t, self.origin.z + self.direction.z * t)
    }
}

/// src/shapes.rs:28-33
#[derive(Clone, Copy, Debug)]
pub struct IntersectionResult { pub t: f64, pub normal: Vec3 }

/// src/shapes.rs:43-48
#[derive(Clone, Copy, Debug)]
pub struct Sphere { pub center: Pnt3, pub radius: f64 }

/// src/shapes.rs:93-98
#[derive(Clone, Copy, Debug)]
pub struct Plane { pub point: Pnt3, pub normal: Vec3 }

/// Upstream's `Box<Shape>` trait object (src/shapes.rs:36-41), as a closed enum: the device flattens objects into tagged
/// structure-of-arrays buffers, so the set of shapes is the set the kernels implement.
#[derive(Clone, Copy, Debug)]
pub enum Shape { Sphere(Sphere), Plane(Plane) }

impl From<Sphere> for Shape { fn from(s: Sphere) -> Shape { Shape::Sphere(s) } }
impl From<Plane> for Shape { fn from(p: Plane) -> Shape { Shape::Plane(p) } }
