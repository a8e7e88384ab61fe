//! Library for Rust raytracer -- the module names and public items of j-dong/rust-raytrace's `lib.rs`
//! (/root/reference/src/lib.rs:4-13), implemented over the C ABI of `libraytrace_b200.so`
//! (include/raytrace_b200.h): scene description stays in Rust, everything from the frame loop of
//! `main.rs:39-59` down (`raytrace::raytrace`, `ray_color`, `Scene::intersect`, materials, lights, backgrounds,
//! quantiser) runs as hand-written sm_100a CUDA kernels.  There is no CPU fallback.
//!
//! SOURCE ONLY: written against the generated `ffi.rs` but never compiled -- the build image has no Rust
//! toolchain.  `ffi.rs` is generated from the header and checked by the test suite.

pub mod ffi;

pub mod shapes;
pub mod types;
pub mod color;
pub mod scene;
pub mod camera;
pub mod raytrace;
pub mod bmp;
pub mod serialize;
#[cfg(feature = "skybox")]
pub mod texture;

use std::ffi::CStr;

/// The library's thread-local message for the last failure (`rt_last_error`).
pub fn last_error() -> String {
    unsafe { CStr::from_ptr(ffi::rt_last_error()) }.to_string_lossy().into_owned()
}

/// Error of a library call: status code (`ffi::RT_ERR_*`) and message.
#[derive(Debug, Clone)]
pub struct Error {
    pub code: i32,
    pub message: String,
}

impl std::fmt::Display for Error {
    fn fmt(&self, f: &mut std::fmt::Formatter) -> std::fmt::Result {
        write!(f, "{}", self.message)
    }
}

impl std::error::Error for Error {}

pub(crate) fn check(rc: i32) -> Result<i32, Error> {
    if rc < 0 { Err(Error { code: rc, message: last_error() }) } else { Ok(rc) }
}
