// build.rs — links libfluid_b200.so when the `cuda` feature is on.
// (Shipped as source; no Rust toolchain exists in the build container, so this file has never been compiled.)
fn main() {
    if std::env::var("CARGO_FEATURE_CUDA").is_ok() {
        let dir = std::env::var("FLUID_B200_LIB_DIR").unwrap_or_else(|_| "../gpu_fluid_simulation_b200".into());
        println!("cargo:rustc-link-search=native={dir}");
        println!("cargo:rustc-link-lib=dylib=fluid_b200");
        println!("cargo:rerun-if-env-changed=FLUID_B200_LIB_DIR");
    }
}
