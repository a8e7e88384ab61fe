// Links libraytrace_b200.so.  RAYTRACE_B200_LIB_DIR overrides the default location next to this crate.
use std::env;
use std::path::PathBuf;

fn main() {
    let dir = env::var("RAYTRACE_B200_LIB_DIR").map(PathBuf::from).unwrap_or_else(|_| {
        PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("..").join("rust-raytrace_b200").join("lib")
    });
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=raytrace_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
    println!("cargo:rerun-if-env-changed=RAYTRACE_B200_LIB_DIR");
}
