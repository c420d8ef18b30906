// build.rs — compiles the CUDA sources for sm_100a with nvcc and links the result.
// UNTESTED: no cargo/rustc in the build image. Mirrors csrc/Makefile.
use std::{env, path::PathBuf, process::Command};

fn main() {
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let csrc = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../../csrc");
    let lib = out.join("libgpu_ec.so");
    let status = Command::new(env::var("NVCC").unwrap_or_else(|_| "nvcc".into()))
        .args(["-O3", "-std=c++17", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a",
               "-Xcompiler", "-fPIC", "-cudart", "static", "-shared", "-o"])
        .arg(&lib)
        .arg(csrc.join("gpu_ec.cu"))
        .args(["-lpthread", "-ldl", "-lrt"])
        .status()
        .expect("nvcc not found");
    assert!(status.success(), "nvcc failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=gpu_ec");
    for f in ["gpu_ec.cu", "rs_kernels.cuh", "crc_kernels.cuh", "lut_common.cuh", "host_math.hpp"] {
        println!("cargo:rerun-if-changed={}", csrc.join(f).display());
    }
}
