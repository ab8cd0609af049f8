// build.rs of the `nvllm-kernels` crate a nano-vllm-rs maintainer would add (north_star: "Host code stays in Rust and
// calls a new kernels crate through a thin extern "C" FFI, built by build.rs with nvcc for sm_100a").
// COMPILE-UNVERIFIED: this image has no Rust toolchain; csrc/Makefile is the recipe that actually runs here and this
// file mirrors it one-to-one.
use std::{env, path::PathBuf, process::Command};

fn main() {
    let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("..");
    let csrc = root.join("nano-vllm-rs_b200").join("csrc");
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".into());
    let mut objs = Vec::new();
    let mut srcs: Vec<PathBuf> = Vec::new();
    for dir in [csrc.clone(), csrc.join("host")] {
        for e in std::fs::read_dir(&dir).unwrap() {
            let p = e.unwrap().path();
            if p.extension().map(|x| x == "cu").unwrap_or(false) {
                println!("cargo:rerun-if-changed={}", p.display());
                srcs.push(p);
            }
        }
    }
    for src in &srcs {
        let obj = out.join(format!("{}.o", src.file_stem().unwrap().to_string_lossy()));
        let ok = Command::new(&nvcc)
            .args(["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
                   "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-c"])
            .arg(src).arg("-o").arg(&obj)
            .status().expect("nvcc not found").success();
        assert!(ok, "nvcc failed on {}", src.display());
        objs.push(obj);
    }
    let lib = out.join("libnvllm.so");
    let ok = Command::new(&nvcc)
        .args(["-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o"]).arg(&lib).args(&objs)
        .args(["-lcudart", "-ldl"]).status().unwrap().success();
    assert!(ok, "link failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=nvllm");
    println!("cargo:rerun-if-changed={}", root.join("include").join("nvllm.h").display());
}
