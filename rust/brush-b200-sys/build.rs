// Links libbrush_b200.so (built by `python -m brush_b200.build`, i.e. nvcc for sm_100a).
//   BRUSH_B200_LIB_DIR   directory holding libbrush_b200.so   (default: <repo>/brush_b200/lib)
use std::env;
use std::path::PathBuf;

fn main() {
    let manifest = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap());
    let default_dir = manifest.join("../../brush_b200/lib");
    let lib_dir = env::var("BRUSH_B200_LIB_DIR").map(PathBuf::from).unwrap_or(default_dir);
    println!("cargo:rustc-link-search=native={}", lib_dir.display());
    println!("cargo:rustc-link-lib=dylib=brush_b200");
    // the library is linked with `-cudart shared`: libcudart.so.12 must be on the loader path at run time
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", lib_dir.display());
    println!("cargo:rerun-if-env-changed=BRUSH_B200_LIB_DIR");
    println!("cargo:rerun-if-changed=../../include/brush_b200.h");
    println!("cargo:rerun-if-changed=src/ffi.rs");
}
