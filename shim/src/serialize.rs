//! Scene text format (/root/reference/src/serialize.rs).  The grammar (src/serialize.rs:367-417, :443-814) is read by
//! the library's own parser (rust-raytrace_b200/csrc/serialize.cpp), which restates upstream's tokenizer, `f64::from_str`
//! and error messages; this module keeps the crate's entry point and error type.

use std::ffi::CStr;
use std::fmt;
use std::os::raw::c_char;
use std::ptr;

use crate::ffi;
use crate::scene::{Options, Scene};

/// Syntax error that is thrown when invalid input is passed to the `deserialize()` function (src/serialize.rs:259-272).
/// `Display` is upstream's: `"{row}:{col}: {message}"`; a lexer error overrides a parser error (src/serialize.rs:436-440).
#[derive(Debug, Clone)]
pub struct SyntaxError {
    /// `row:col: message`, exactly as upstream formats it
    text: String,
}

impl SyntaxError {
    /// 1-based (row, column) parsed back out of the message
    pub fn location(&self) -> Option<(u32, u32)> {
        let mut it = self.text.splitn(3, ':');
        let r = it.next()?.trim().parse().ok()?;
        let c = it.next()?.trim().parse().ok()?;
        Some((r, c))
    }
}

impl fmt::Display for SyntaxError {
    fn fmt(&self, fmt: &mut fmt::Formatter) -> fmt::Result { write!(fmt, "{}", self.text) }
}

impl std::error::Error for SyntaxError {
    fn description(&self) -> &str { "syntax error" }
}

/// Deserialize the given text as a `Scene` object (src/serialize.rs:427-441).
pub fn deserialize(text: &String) -> Result<Scene, SyntaxError> {
    let mut handle: *mut ffi::RtScene = ptr::null_mut();
    let mut err = [0 as c_char; 1024];
    let rc = unsafe { ffi::rt_scene_parse(text.as_ptr() as *const c_char, text.len(), &mut handle, err.as_mut_ptr(), err.len()) };
    if rc != ffi::RT_OK {
        let msg = unsafe { CStr::from_ptr(err.as_ptr()) }.to_string_lossy().into_owned();
        return Err(SyntaxError { text: if msg.is_empty() { crate::last_error() } else { msg } });
    }
    let (mut w, mut h, mut aa) = (0u32, 0u32, 0u32);
    unsafe { ffi::rt_scene_get_options(handle, &mut w, &mut h, &mut aa) };
    let warnings = unsafe { CStr::from_ptr(ffi::rt_scene_warnings(handle)) }.to_string_lossy().into_owned();
    if !warnings.is_empty() { print!("{}", warnings); }   // upstream println!s them while parsing
    Ok(Scene::from_handle(handle, Options { width: w, height: h, antialias: aa }, warnings))
}
