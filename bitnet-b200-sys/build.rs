// Links libbitnet_b200.so. Mirrors how crates/bitnet-kernels/build.rs links the microsoft/BitNet C++ bridge
// (reference: crates/bitnet-kernels/src/ffi/bridge.rs:17-39 binds the same six bitnet_cpp_* symbols).
//
//   BITNET_B200_LIB_DIR   directory holding libbitnet_b200.so   (default: ../bitnet-rs_b200)
//   BITNET_B200_BUILD=1   run `make -C ../bitnet-rs_b200/csrc` first (needs nvcc, sm_100a cross-compiles without a GPU)
use std::{env, path::PathBuf, process::Command};

fn main() {
    let manifest = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap());
    let default_dir = manifest.join("..").join("bitnet-rs_b200");
    let lib_dir = env::var("BITNET_B200_LIB_DIR").map(PathBuf::from).unwrap_or(default_dir.clone());
    if env::var("BITNET_B200_BUILD").as_deref() == Ok("1") {
        let status = Command::new("make")
            .arg("-j8")
            .arg("-C")
            .arg(default_dir.join("csrc"))
            .status()
            .expect("failed to run make for libbitnet_b200.so");
        assert!(status.success(), "building libbitnet_b200.so failed");
    }
    println!("cargo:rustc-link-search=native={}", lib_dir.display());
    println!("cargo:rustc-link-lib=dylib=bitnet_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", lib_dir.display());
    println!("cargo:rerun-if-env-changed=BITNET_B200_LIB_DIR");
    println!("cargo:rerun-if-changed=../include/bitnet_b200.h");
}
