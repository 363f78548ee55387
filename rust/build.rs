// build.rs for the reference crate with the B200 codec: proto generation as before, then nvcc for sm_100a.
// UNCOMPILED on the build box of this repository (no rustc); see rust/README.md.
use std::{env, path::PathBuf, process::Command};

const CUDA_SOURCES: &[&str] = &["tfs_ctx.cu", "tfs_copy.cu", "tfs_encode.cu", "tfs_decode.cu", "tfs_host.cu", "tfs_noalloc.cu", "tfs_engine.cu"];

fn main() -> Result<(), Box<dyn std::error::Error>> {
    // 1. unchanged: generate the prost structs and the tonic client stubs
    let protos = [
        "protos/tensorflow_serving/apis/prediction_service.proto",
        "protos/tensorflow_serving/apis/model_service.proto",
        "protos/tensorflow/core/lib/core/error_codes.proto",
    ];
    tonic_build::configure().compile(&protos, &["protos".to_string()])?;

    // 2. new: the CUDA library, hand-written for sm_100a
    let out = PathBuf::from(env::var("OUT_DIR")?);
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".into());
    let mut objects = Vec::new();
    for src in CUDA_SOURCES {
        let obj = out.join(format!("{}.o", src));
        let status = Command::new(&nvcc)
            .args(&["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "--expt-relaxed-constexpr"])
            .args(&["-Xcompiler", "-fPIC", "-Iinclude", "-c"])
            .arg(format!("csrc/{}", src))
            .arg("-o")
            .arg(&obj)
            .status()?;
        if !status.success() {
            return Err(format!("nvcc failed on {}", src).into());
        }
        println!("cargo:rerun-if-changed=csrc/{}", src);
        objects.push(obj);
    }
    let lib = out.join("libtfs_gpu.a");
    let status = Command::new("ar").arg("rcs").arg(&lib).args(&objects).status()?;
    if !status.success() {
        return Err("ar failed".into());
    }
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=static=tfs_gpu");
    println!("cargo:rustc-link-search=native=/usr/local/cuda/lib64");
    println!("cargo:rustc-link-lib=dylib=cudart");
    println!("cargo:rustc-link-lib=dylib=stdc++");
    println!("cargo:rustc-link-lib=dylib=pthread"); // the engine lanes, the staging helper threads
    println!("cargo:rerun-if-changed=include/tfs_gpu.h");
    Ok(())
}
