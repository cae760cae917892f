// Link flags for libb200rt.so (built in-tree by `python rust_raytracer_b200/build.py`: nvcc, sm_100a).  No `cc` step:
// the CUDA sources are compiled by the repo's own build, this crate only binds the result.
fn main() {
    let root = std::env::var("B200RT_ROOT").unwrap_or_else(|_| {
        let here = std::path::PathBuf::from(std::env::var("CARGO_MANIFEST_DIR").unwrap());
        here.join("../..").canonicalize().expect("b200rt checkout").to_string_lossy().into_owned()
    });
    println!("cargo:rerun-if-env-changed=B200RT_ROOT");
    println!("cargo:rerun-if-changed={root}/include/b200rt.h");
    println!("cargo:rustc-link-search=native={root}/rust_raytracer_b200");
    println!("cargo:rustc-link-lib=dylib=b200rt");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{root}/rust_raytracer_b200");
    #[cfg(feature = "bindgen")]
    {
        let out = std::path::PathBuf::from(std::env::var("OUT_DIR").unwrap()).join("bindings.rs");
        bindgen::Builder::default()
            .header(format!("{root}/include/b200rt.h"))
            .allowlist_function("b200rt_.*")
            .allowlist_type("b200rt_.*")
            .allowlist_var("B200RT_.*")
            .generate()
            .expect("bindgen on include/b200rt.h")
            .write_to_file(out)
            .expect("write bindings.rs");
    }
}
