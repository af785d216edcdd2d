// Links libswb200.so.  SWB200_LIB_DIR points at the directory that holds it (default: the in-tree build,
// smithwaterman_b200/); SWB200_BUILD=1 runs the repository's own build step first (needs nvcc + python).
use std::env;
use std::path::PathBuf;
use std::process::Command;

fn main() {
    let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../..");
    if env::var("SWB200_BUILD").map(|v| v == "1").unwrap_or(false) {
        let st = Command::new("python")
            .current_dir(&root)
            .args(&["-c", "import __graft_entry__ as g; g.build()"])
            .status()
            .expect("python not found");
        assert!(st.success(), "building libswb200.so failed");
    }
    let dir = env::var("SWB200_LIB_DIR").map(PathBuf::from).unwrap_or_else(|_| root.join("smithwaterman_b200"));
    println!("cargo:rustc-link-search=native={}", dir.display());
    println!("cargo:rustc-link-lib=dylib=swb200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", dir.display());
    println!("cargo:rerun-if-env-changed=SWB200_LIB_DIR");
    println!("cargo:rerun-if-changed=../../include/swb200.h");
}
