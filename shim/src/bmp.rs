//! BMP output (/root/reference/src/bmp.rs): a 122-byte header (14 + 108-byte V4 info header), 24-bit BGR rows padded to
//! four bytes, bottom-up.  The bytes come from the library (`rt_bmp_header`) and equal upstream's (the first 122 bytes of
//! its out.bmp; tests/test_oracle_kats.py).

use std::fs::File;
use std::io;
use std::io::prelude::*;

use crate::ffi;

/// Writes the header and returns the byte width of a row, `(3 * width + 3) & !3` (src/bmp.rs:10-61).
pub fn write_header(f: &mut File, width: u32, height: u32) -> io::Result<u32> {
    let mut h = [0u8; 122];
    unsafe { ffi::rt_bmp_header(width, height, h.as_mut_ptr()) };
    f.write_all(&h)?;
    Ok(unsafe { ffi::rt_bmp_bytewidth(width) })
}

/// Header + all rows (row 0 = bottom of the image, as `main` writes them: src/main.rs:45-58).
pub fn write(path: &str, width: u32, height: u32, rows: &[u8]) -> Result<(), crate::Error> {
    let bw = unsafe { ffi::rt_bmp_bytewidth(width) } as usize;
    assert_eq!(rows.len(), bw * height as usize);
    let c = std::ffi::CString::new(path).map_err(|_| crate::Error { code: ffi::RT_ERR_INVALID, message: "path contains NUL".to_string() })?;
    crate::check(unsafe { ffi::rt_bmp_write(c.as_ptr(), width, height, rows.as_ptr()) })?;
    Ok(())
}
