// Links ray_tracer_b200/librt_b200.so (built by `make -C ray_tracer_b200/csrc`, nvcc sm_100a).
fn main() {
    let dir = std::env::var("RT_B200_LIB_DIR").unwrap_or_else(|_| format!("{}/../ray_tracer_b200", env!("CARGO_MANIFEST_DIR")));
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=rt_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{dir}");
    println!("cargo:rerun-if-env-changed=RT_B200_LIB_DIR");
}
