// The `raytrace` binary (/root/reference/src/main.rs) over the library: same file names, same messages, same exit
// status; lines 39-59 of upstream's main (the y / x / antialias loops) are the one `raytrace::render` call.
extern crate libraytrace;

use std::fs::File;
use std::io::prelude::*;
use std::io::BufReader;

use libraytrace::bmp;
use libraytrace::raytrace;
use libraytrace::serialize;
use libraytrace::types::*;

fn main() {
    // read a file
    let scene = {
        let file = match File::open("test_scene.txt") {
            Ok(f) => f,
            Err(e) => { println!("error: {}", e); return },
        };
        let mut reader = BufReader::new(file);
        let mut contents = String::new();
        match reader.read_to_string(&mut contents) {
            Ok(_) => (),
            Err(e) => { println!("error: {}", e); return },
        }
        match serialize::deserialize(&contents) {
            Ok(s) => s,
            Err(e) => { println!("error: {}", e); return },
        }
    };
    let width = scene.options.width;
    let height = scene.options.height;
    // write BMP
    let mut file_handle = File::create("out.bmp").ok().expect("error creating out.bmp");
    let bytewidth = bmp::write_header(&mut file_handle, width, height).ok().expect("error writing BMP header");
    // render image: upstream's loops over y, x and the antialiasing samples (src/main.rs:39-57) run on the GPU
    let mut opts = raytrace::RenderOptions::default();
    opts.rng = RngT::weak(); // rand::weak_rng() (src/main.rs:43)
    let frame = match raytrace::render(&scene, &opts) {
        Ok(f) => f,
        Err(e) => panic!("render failed: {}", e),
    };
    assert_eq!(frame.bytewidth, bytewidth);
    for row in frame.rows.chunks(bytewidth as usize) {
        file_handle.write_all(row).ok().expect("error writing row");
    }
}
