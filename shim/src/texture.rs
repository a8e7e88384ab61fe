//! Textures (/root/reference/src/texture.rs): RGB8 in the sRGB colour space, row 0 = top row of the decoded image
//! (`image::open(path).to_rgb().into_raw()`, src/texture.rs:34-37).  Sampling (`Texture::sample`, src/texture.rs:46-58:
//! clamp, truncate, bilinear in y then x, texels decoded through SRGB_VALUES) runs on the device in the hit kernel.

/// src/texture.rs:20-24
#[derive(Clone, Debug)]
pub struct Texture {
    pub width: u32,
    pub height: u32,
    /// RGB8, `width * height * 3` bytes
    pub data: Box<[u8]>,
}

/// Upstream: `image::ImageError`.  Here: the library's message for a file it cannot decode.
pub type LoadError = crate::Error;

pub fn error_description(err: LoadError) -> String { format!("error #{}", err) }

impl Texture {
    pub fn from_rgb8(width: u32, height: u32, data: Vec<u8>) -> Texture {
        assert_eq!(data.len(), (width as usize) * (height as usize) * 3);
        Texture { width: width, height: height, data: data.into_boxed_slice() }
    }
    /// `Texture::load` for the formats the library decodes itself: 24-bit BMP and binary PPM (the `image` crate's other
    /// decoders are out of scope: decode in Rust and use `from_rgb8`).  The file is read when the scene is built
    /// (`rt_scene_load_skybox_face`), so this only records the path.
    pub fn load<P: AsRef<std::path::Path>>(path: P) -> Result<TextureSource, LoadError> {
        Ok(TextureSource::File(path.as_ref().to_path_buf()))
    }
}

/// A skybox face: pixels already in memory, or a BMP / PPM file for the library to decode.
#[derive(Clone, Debug)]
pub enum TextureSource {
    Pixels(Texture),
    File(std::path::PathBuf),
}

impl From<Texture> for TextureSource { fn from(t: Texture) -> TextureSource { TextureSource::Pixels(t) } }
