// crates/core/build.rs addition -- call build_w4a16_b200() next to the existing `nvcc --ptx` loop (build.rs:263-279,
// flags build.rs:335-339).  Builds libw4a16_b200.so from the vendored sources with nvcc for sm_100a and links it.
// The sources are expected under third_party/b200-w4a16 (a checkout of this repository).
use std::path::{Path, PathBuf};
use std::process::Command;

pub fn build_w4a16_b200() {
    let root = Path::new(env!("CARGO_MANIFEST_DIR")).join("third_party/b200-w4a16");
    let src = root.join("vllm-rust_b200/csrc");
    let out = PathBuf::from(std::env::var("OUT_DIR").expect("OUT_DIR"));
    // The Makefile compiles every .cu with `-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo` and links
    // vllm-rust_b200/lib/libw4a16_b200.so (cudart dynamically, NCCL through dlopen like nccl.rs:479-485).
    let status = Command::new("make")
        .arg("-C")
        .arg(&src)
        .arg("-j8")
        .status()
        .expect("failed to start make (nvcc for sm_100a is required)");
    assert!(status.success(), "libw4a16_b200.so did not build");
    let lib = root.join("vllm-rust_b200/lib/libw4a16_b200.so");
    std::fs::copy(&lib, out.join("libw4a16_b200.so")).expect("copy libw4a16_b200.so");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=w4a16_b200");
    println!("cargo:rerun-if-changed={}", src.display());
    println!("cargo:rerun-if-changed={}", root.join("include/w4a16_b200.h").display());
}
