// build.rs — compiles the CUDA sources for sm_100a and links them.  UNBUILT here (no cargo in the image).
// Mirrors __graft_entry__.build(): -fmad=false / -ffp-contract=off are part of the parity contract.
use std::{env, path::PathBuf, process::Command};

fn main() {
    let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../..");
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let csrc = root.join("csrc");
    let lib = out.join("librtx_cuda.so");
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "nvcc".into());
    let status = Command::new(nvcc)
        .args(["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false"])
        .args(["-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math", "-shared", "-o"])
        .arg(&lib)
        .args(["rtx_api.cu", "rtx_render.cu", "rtx_debug.cu"].iter().map(|f| csrc.join(f)))
        .status()
        .expect("nvcc not found");
    assert!(status.success(), "nvcc failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=rtx_cuda");
    println!("cargo:rerun-if-changed={}", csrc.display());
}
