//! The tracer (/root/reference/src/raytrace.rs).
//!
//! Upstream exposes the per-sample functions `raytrace(scene, pos, significance, rng)` and `ray_color(scene, ray,
//! significance, depth, rng)` and lets `main` run the serial y / x / antialias loops over them (src/main.rs:39-59).  One
//! FFI call per pixel sample would spend its time in the call, so the unit of work here is the FRAME: `render` replaces
//! the whole loop nest, on one GPU or sharded by row bands over several (`render_multi`).  `project` and
//! `intersect_batch` expose `Camera::project` and `Scene::intersect` for batches, for callers and tests that want the
//! intermediate results.

use std::ptr;

use crate::ffi;
use crate::scene::Scene;
use crate::shapes::Ray;
use crate::types::*;
use crate::{check, Error};

/// src/raytrace.rs:17
pub const MIN_SIGNIFICANCE: f64 = 1.0 / 256.0 / 2.0;
/// src/raytrace.rs:18
pub const MAX_DEPTH: u32 = 4;

/// Runtime versions of upstream's compile-time constants, plus what the device needs to know.
#[derive(Clone, Copy, Debug)]
pub struct RenderOptions {
    /// 0 = the scene's `Options`
    pub width: u32,
    pub height: u32,
    /// `None` = the scene's `Options::antialias`
    pub antialias: Option<u32>,
    pub max_depth: u32,
    pub min_significance: f64,
    pub rng: RngT,
    /// `ffi::RT_ACCEL_*`: the device-built BVH returns exactly the linear scan's closest hit
    pub accel: i32,
    /// `ffi::RT_FP_STRICT` = the reference's unfused f64 arithmetic
    pub fp_mode: i32,
    pub device: i32,
}

impl Default for RenderOptions {
    fn default() -> RenderOptions {
        RenderOptions { width: 0, height: 0, antialias: None, max_depth: MAX_DEPTH, min_significance: MIN_SIGNIFICANCE, rng: RngT::from_seed(1),
                        accel: ffi::RT_ACCEL_AUTO, fp_mode: ffi::RT_FP_STRICT, device: 0 }
    }
}

impl RenderOptions {
    pub(crate) fn to_ffi(&self) -> ffi::RtRenderOpts {
        let mut o: ffi::RtRenderOpts = unsafe { std::mem::zeroed() };
        unsafe { ffi::rt_render_opts_default(&mut o) };
        o.width = self.width;
        o.height = self.height;
        if let Some(aa) = self.antialias { o.antialias = aa; }
        o.max_depth = self.max_depth;
        o.min_significance = self.min_significance;
        o.seed = self.rng.seed;
        o.accel = self.accel;
        o.fp_mode = self.fp_mode;
        o.device = self.device;
        o
    }
}

/// A rendered frame: rows exactly as `main` writes them after the BMP header.
pub struct Frame {
    pub width: u32,
    pub height: u32,
    /// bytes per row, `(3 * width + 3) & !3` (src/bmp.rs:11)
    pub bytewidth: u32,
    /// `bytewidth * height` bytes, BGR, row 0 = bottom of the image (y = 0)
    pub rows: Vec<u8>,
    /// object index hit by the first sample's primary ray of every pixel, -1 = background
    pub primary_ids: Vec<i32>,
    pub stats: ffi::RtStats,
}

fn frame_size(scene: &Scene, o: &RenderOptions) -> (u32, u32) {
    (if o.width != 0 { o.width } else { scene.options.width }, if o.height != 0 { o.height } else { scene.options.height })
}

/// The frame loop of src/main.rs:39-59 over `raytrace` (src/raytrace.rs:270-276), on one GPU.
pub fn render(scene: &Scene, opts: &RenderOptions) -> Result<Frame, Error> {
    let (w, h) = frame_size(scene, opts);
    let bw = unsafe { ffi::rt_bmp_bytewidth(w) };
    let mut rows = vec![0u8; bw as usize * h as usize];
    let mut ids = vec![-1i32; w as usize * h as usize];
    let mut stats: ffi::RtStats = unsafe { std::mem::zeroed() };
    let o = opts.to_ffi();
    check(unsafe { ffi::rt_render(scene.handle(), &o, rows.as_mut_ptr(), ids.as_mut_ptr(), ptr::null_mut(), &mut stats) })?;
    Ok(Frame { width: w, height: h, bytewidth: bw, rows: rows, primary_ids: ids, stats: stats })
}

/// The same frame sharded by interleaved row bands over `devices` (CUDA ordinals) of one node; `gather` = `ffi::RT_GATHER_*`.
pub fn render_multi(scene: &Scene, opts: &RenderOptions, devices: &[i32], gather: i32) -> Result<Frame, Error> {
    let (w, h) = frame_size(scene, opts);
    let bw = unsafe { ffi::rt_bmp_bytewidth(w) };
    let mut rows = vec![0u8; bw as usize * h as usize];
    let mut ids = vec![-1i32; w as usize * h as usize];
    let mut stats: ffi::RtStats = unsafe { std::mem::zeroed() };
    let o = opts.to_ffi();
    check(unsafe { ffi::rt_render_multi(scene.handle(), &o, devices.len() as i32, devices.as_ptr(), gather, rows.as_mut_ptr(), ids.as_mut_ptr(), &mut stats) })?;
    Ok(Frame { width: w, height: h, bytewidth: bw, rows: rows, primary_ids: ids, stats: stats })
}

/// A frame that is being traced: `begin` queues it on the GPU and returns, `finish` waits for it and copies it back.  An
/// animation keeps two of these alive -- frame k + 1 is begun before frame k is finished -- so that the GPU is never idle in
/// the latency-bound deep levels of a frame.  One GPU; the scene must not be changed while a frame is in flight.
pub struct FrameInFlight {
    ctx: *mut ffi::RtContext,
    width: u32,
    height: u32,
}

impl FrameInFlight {
    pub fn begin(scene: &Scene, opts: &RenderOptions) -> Result<FrameInFlight, Error> {
        let (w, h) = frame_size(scene, opts);
        let o = opts.to_ffi();
        let mut ctx: *mut ffi::RtContext = ptr::null_mut();
        check(unsafe { ffi::rt_context_create(scene.handle(), &o, &mut ctx) })?;
        let f = FrameInFlight { ctx: ctx, width: w, height: h };
        check(unsafe { ffi::rt_context_render_begin(f.ctx) })?;
        Ok(f)
    }

    pub fn finish(self) -> Result<Frame, Error> {
        let (w, h) = (self.width, self.height);
        let bw = unsafe { ffi::rt_bmp_bytewidth(w) };
        let mut stats: ffi::RtStats = unsafe { std::mem::zeroed() };
        check(unsafe { ffi::rt_context_render_end(self.ctx, &mut stats) })?;
        // world == 1: the local rows are the image rows, plus band padding at the end that is not part of the frame
        let local = stats.local_rows as usize;
        let mut rows = vec![0u8; bw as usize * local];
        let mut ids = vec![-1i32; w as usize * local];
        check(unsafe { ffi::rt_context_download(self.ctx, rows.as_mut_ptr(), ids.as_mut_ptr(), ptr::null_mut()) })?;
        rows.truncate(bw as usize * h as usize);
        ids.truncate(w as usize * h as usize);
        Ok(Frame { width: w, height: h, bytewidth: bw, rows: rows, primary_ids: ids, stats: stats })
    }
}

impl Drop for FrameInFlight {
    fn drop(&mut self) {
        unsafe { ffi::rt_context_destroy(self.ctx) };   // waits for whatever is still queued
    }
}

/// `Scene::intersect` (src/scene.rs:247-249) for a batch: (object index or -1, t, normal) per ray.
pub fn intersect_batch(scene: &Scene, rays: &[Ray], opts: &RenderOptions) -> Result<Vec<(i32, f64, Vec3)>, Error> {
    let n = rays.len();
    let mut od = Vec::with_capacity(6 * n);
    for r in rays { od.extend_from_slice(&[r.origin.x, r.origin.y, r.origin.z, r.direction.x, r.direction.y, r.direction.z]); }
    let mut ids = vec![-1i32; n];
    let mut t = vec![0f64; n];
    let mut nrm = vec![0f64; 3 * n];
    let o = opts.to_ffi();
    check(unsafe { ffi::rt_trace_rays(scene.handle(), &o, n as u64, od.as_ptr(), ids.as_mut_ptr(), t.as_mut_ptr(), nrm.as_mut_ptr()) })?;
    Ok((0..n).map(|i| (ids[i], t[i], Vec3::new(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]))).collect())
}

/// `Camera::project` (src/camera.rs:77-79, :110-121) for a batch of normalised image positions.
pub fn project(scene: &Scene, positions: &[Pnt2], opts: &RenderOptions) -> Result<Vec<Ray>, Error> {
    let n = positions.len();
    let mut p = Vec::with_capacity(2 * n);
    for q in positions { p.push(q.x); p.push(q.y); }
    let mut od = vec![0f64; 6 * n];
    let o = opts.to_ffi();
    check(unsafe { ffi::rt_project(scene.handle(), &o, n as u64, p.as_ptr(), od.as_mut_ptr()) })?;
    Ok((0..n).map(|i| Ray { origin: Pnt3::new(od[6 * i], od[6 * i + 1], od[6 * i + 2]), direction: Vec3::new(od[6 * i + 3], od[6 * i + 4], od[6 * i + 5]) }).collect())
}
