// Links librptx.so (built in-tree by `make -C rpt_b200/csrc`).  RPTX_LIB_DIR overrides the search path.
fn main() {
    let dir = std::env::var("RPTX_LIB_DIR").unwrap_or_else(|_| "../../rpt_b200".to_string());
    println!("cargo:rustc-link-search=native={}", dir);
    println!("cargo:rustc-link-lib=dylib=rptx");
    println!("cargo:rerun-if-env-changed=RPTX_LIB_DIR");
}
