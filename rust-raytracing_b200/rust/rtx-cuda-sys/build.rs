// build.rs — compiles the CUDA sources for sm_100a and links them.  UNBUILT here (no cargo in the image).
// Mirrors __graft_entry__.build(): every csrc/*.cu goes into ONE shared library, and -fmad=false /
// -ffp-contract=off are part of the parity contract.  tests/test_rust_crate.py checks (without cargo) that this
// file globs csrc/ and passes the same flags as __graft_entry__.NVCC_FLAGS.
use std::{env, fs, path::PathBuf, process::Command};

fn main() {
    let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../..");
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let csrc = root.join("csrc");
    let lib = out.join("librtx_cuda.so");
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "nvcc".into());
    // every translation unit of the library: csrc/*.cu (the same glob as __graft_entry__.build())
    let mut sources: Vec<PathBuf> = fs::read_dir(&csrc)
        .expect("csrc/ not found")
        .filter_map(|e| e.ok().map(|e| e.path()))
        .filter(|p| p.extension().map_or(false, |x| x == "cu"))
        .collect();
    sources.sort();
    assert!(!sources.is_empty(), "no .cu sources under {}", csrc.display());
    let status = Command::new(nvcc)
        .args(["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false"])
        .args(["-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math", "-shared", "-o"])
        .arg(&lib)
        .args(&sources)
        .status()
        .expect("nvcc not found");
    assert!(status.success(), "nvcc failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=rtx_cuda");
    println!("cargo:rerun-if-changed={}", csrc.display());
    println!("cargo:rerun-if-changed={}", root.join("../include/rtx_cuda.h").display());
}
