//! Structs and functions for working with color (/root/reference/src/color.rs).
//! The sRGB threshold tables (src/color.rs:75-591) live in the library; `to_srgb` runs there (`rt_quantize_bgr`).

use std::ops::{Add, Div, Mul, Sub};

use crate::ffi;

/// A color in a linear RGB color space (src/color.rs:15-22).
#[derive(Clone, Copy, Debug, PartialEq)]
pub struct Color { pub r: f64, pub g: f64, pub b: f64 }

/// src/color.rs:25
pub const BLACK: Color = Color { r: 0.0, g: 0.0, b: 0.0 };

impl Add for Color { type Output = Color; fn add(self, o: Color) -> Color { Color { r: self.r + o.r, g: self.g + o.g, b: self.b + o.b } } }
impl Sub for Color { type Output = Color; fn sub(self, o: Color) -> Color { Color { r: self.r - o.r, g: self.g - o.g, b: self.b - o.b } } }
impl Mul for Color { type Output = Color; fn mul(self, o: Color) -> Color { Color { r: self.r * o.r, g: self.g * o.g, b: self.b * o.b } } }
impl Mul<f64> for Color { type Output = Color; fn mul(self, s: f64) -> Color { Color { r: self.r * s, g: self.g * s, b: self.b * s } } }
impl Div<f64> for Color { type Output = Color; fn div(self, s: f64) -> Color { Color { r: self.r / s, g: self.g / s, b: self.b / s } } }

impl Color {
    /// src/color.rs:605
    pub fn from_rgb(r: f64, g: f64, b: f64) -> Color { Color { r: r, g: g, b: b } }
    pub fn as_array(&self) -> [f64; 3] { [self.r, self.g, self.b] }
    /// Quantised `[b, g, r]` (src/color.rs:593-600, :623): smallest `i` with `v < SRGB_AVERAGE[i]`, else 255.
    pub fn bgr(&self) -> [u8; 3] {
        let rgb = self.as_array();
        let mut out = [0u8; 3];
        let rc = unsafe { ffi::rt_quantize_bgr(rgb.as_ptr(), 1, out.as_mut_ptr(), 0) };
        assert!(rc == 0, "rt_quantize_bgr: {}", crate::last_error());
        out
    }
    /// src/color.rs:617
    pub fn rgb(&self) -> [u8; 3] { let c = self.bgr(); [c[2], c[1], c[0]] }
    /// src/color.rs:628-632: stores b, g, r at `3i..3i+2` (panics when the buffer is too short, as upstream)
    pub fn write_bgr(&self, buf: &mut Vec<u8>, i: usize) {
        let c = self.bgr();
        buf[i * 3] = c[0];
        buf[i * 3 + 1] = c[1];
        buf[i * 3 + 2] = c[2];
    }
    /// src/color.rs:637-639
    pub fn significance(&self) -> f64 { self.r + self.g + self.b }
}

/// A whole buffer of colours at once (what the frame's resolve kernel does on the device): `rgb` holds 3n doubles.
pub fn quantize_bgr(rgb: &[f64], device: i32) -> Result<Vec<u8>, crate::Error> {
    assert!(rgb.len() % 3 == 0);
    let mut out = vec![0u8; rgb.len()];
    crate::check(unsafe { ffi::rt_quantize_bgr(rgb.as_ptr(), (rgb.len() / 3) as u64, out.as_mut_ptr(), device) })?;
    Ok(out)
}
