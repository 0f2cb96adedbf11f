// build.rs  (UNTESTED — no Rust toolchain in the build container)
fn main() {
    let dir = std::env::var("RTGPU_LIB_DIR").expect("set RTGPU_LIB_DIR to the directory holding librtgpu.so");
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=rtgpu");
    println!("cargo:rerun-if-env-changed=RTGPU_LIB_DIR");
}
