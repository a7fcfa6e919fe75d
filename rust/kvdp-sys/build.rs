// Links against libkvdp.so. KVDP_LIB_DIR = directory holding the library (dynamo_exp_b200/ in this repository).
fn main() {
    let dir = std::env::var("KVDP_LIB_DIR").expect("set KVDP_LIB_DIR to the directory that holds libkvdp.so");
    println!("cargo:rustc-link-search=native={dir}");
    println!("cargo:rustc-link-lib=dylib=kvdp");
    println!("cargo:rustc-link-lib=dylib=stdc++");
    println!("cargo:rerun-if-env-changed=KVDP_LIB_DIR");
}
