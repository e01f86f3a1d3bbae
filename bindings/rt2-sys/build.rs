// Link against the in-tree librt2cuda.so.  RT2_LIB_DIR overrides the default (<repo>/raytracer2_b200).
fn main() {
    let dir = std::env::var("RT2_LIB_DIR").unwrap_or_else(|_| {
        let manifest = std::env::var("CARGO_MANIFEST_DIR").unwrap();
        format!("{manifest}/../../raytracer2_b200")
    });
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=rt2cuda");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{dir}");
    println!("cargo:rerun-if-env-changed=RT2_LIB_DIR");
}
