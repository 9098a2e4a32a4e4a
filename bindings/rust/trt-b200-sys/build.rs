// Links libtrt_b200.so.  TRT_B200_LIB_DIR = directory holding the library (defaults to the in-tree build
// terminal-raytracer_b200/ of the repository this crate ships in).
fn main() {
    let dir = std::env::var("TRT_B200_LIB_DIR").unwrap_or_else(|_| {
        let here = std::path::PathBuf::from(std::env::var("CARGO_MANIFEST_DIR").unwrap());
        here.join("../../../terminal-raytracer_b200").to_string_lossy().into_owned()
    });
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=trt_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{dir}");
    println!("cargo:rerun-if-env-changed=TRT_B200_LIB_DIR");
}
