//! Type aliases for convenience (/root/reference/src/types.rs).  Upstream aliases nalgebra 0.4 and rand 0.3 types; the
//! shim needs neither crate -- vectors only carry coordinates across the FFI, and the random numbers are drawn on the
//! device by a counter RNG keyed by `RenderOptions::seed`.

/// 2D point
#[derive(Clone, Copy, Debug, PartialEq)]
pub struct Pnt2 { pub x: f64, pub y: f64 }
/// 3D point
#[derive(Clone, Copy, Debug, PartialEq)]
pub struct Pnt3 { pub x: f64, pub y: f64, pub z: f64 }
/// 3D vector
#[derive(Clone, Copy, Debug, PartialEq)]
pub struct Vec3 { pub x: f64, pub y: f64, pub z: f64 }
/// 3x3 matrix, row-major (the argument order of nalgebra's `Mat3::new`, src/camera.rs:57-61)
#[derive(Clone, Copy, Debug, PartialEq)]
pub struct Mat3 { pub m: [f64; 9] }

impl Pnt2 { pub fn new(x: f64, y: f64) -> Pnt2 { Pnt2 { x: x, y: y } } }
impl Pnt3 {
    pub fn new(x: f64, y: f64, z: f64) -> Pnt3 { Pnt3 { x: x, y: y, z: z } }
    pub fn as_array(&self) -> [f64; 3] { [self.x, self.y, self.z] }
}
impl Vec3 {
    pub fn new(x: f64, y: f64, z: f64) -> Vec3 { Vec3 { x: x, y: y, z: z } }
    pub fn as_array(&self) -> [f64; 3] { [self.x, self.y, self.z] }
}

/// What stands in for upstream's `RngT = XorShiftRng` (src/types.rs:27): the seed of the device's counter RNG.
/// Upstream threads one OS-seeded generator through the whole frame; here every pixel sample derives its numbers
/// from (seed, pixel, sample), so frames are reproducible and pixels independent (which is what lets the image shard).
#[derive(Clone, Copy, Debug)]
pub struct RngT { pub seed: u64 }

impl RngT {
    pub fn from_seed(seed: u64) -> RngT { RngT { seed: seed } }
    /// `rand::weak_rng()` (src/main.rs:43): seeded from the clock
    pub fn weak() -> RngT {
        let t = std::time::SystemTime::now().duration_since(std::time::UNIX_EPOCH).map(|d| d.as_nanos() as u64).unwrap_or(1);
        RngT { seed: t | 1 }
    }
}
