// build.rs — compiles the CUDA / C++ sources of librayca_cuda for sm_100a by invoking nvcc directly (no cmake, no cc
// crate: the flags are the parity contract and are spelled out), mirroring rayca_b200/build.py.
//   --fmad=false                 device code must not contract a*b+c (Rust never fuses; bit parity with rayca-soft)
//   -Xcompiler -ffp-contract=off the same rule for the host flattening / BVH code
//   -gencode arch=compute_100a,code=sm_100a   B200 only: no other architecture, no PTX fallback
// Environment: NVCC (default /usr/local/cuda/bin/nvcc), RAYCA_CUDA_SRC (default ../rayca_b200/csrc relative to this
// crate, i.e. the layout of this repository), RAYCA_CUDA_LIB_DIR with feature `prebuilt`.
use std::env;
use std::path::PathBuf;
use std::process::Command;

const SOURCES: [&str; 7] = ["rayca_cuda.cu", "host_scene.cpp", "bvh_build.cpp", "fast_bvh.cpp", "gpu_fast_build.cu", "png_encode.cpp", "wavefront.cu"];
const HEADERS: [&str; 8] = ["rmath.h", "gpu_types.h", "host_scene.h", "gpu_fast_build.h", "trace.cuh", "shade.cuh", "render.cuh", "wavefront.h"];

fn main() {
    let out_dir = PathBuf::from(env::var("OUT_DIR").expect("OUT_DIR"));
    if env::var("CARGO_FEATURE_PREBUILT").is_ok() {
        let dir = env::var("RAYCA_CUDA_LIB_DIR").expect("feature `prebuilt` needs RAYCA_CUDA_LIB_DIR");
        println!("cargo:rustc-link-search=native={dir}");
        println!("cargo:rustc-link-lib=dylib=rayca_cuda");
        println!("cargo:rerun-if-env-changed=RAYCA_CUDA_LIB_DIR");
        return;
    }
    let manifest = PathBuf::from(env::var("CARGO_MANIFEST_DIR").expect("CARGO_MANIFEST_DIR"));
    let src = env::var("RAYCA_CUDA_SRC").map(PathBuf::from).unwrap_or_else(|_| manifest.join("../rayca_b200/csrc"));
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".to_string());
    let lib = out_dir.join("librayca_cuda.so");

    let mut cmd = Command::new(&nvcc);
    cmd.args(["-shared", "-o"]).arg(&lib);
    cmd.args(["-std=c++17", "-O3", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a", "--fmad=false", "-x", "cu"]);
    cmd.args(["-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math,-fopenmp,-O2", "-Xptxas", "-O3"]);
    for s in SOURCES {
        cmd.arg(src.join(s));
        println!("cargo:rerun-if-changed={}", src.join(s).display());
    }
    for h in HEADERS {
        println!("cargo:rerun-if-changed={}", src.join(h).display());
    }
    println!("cargo:rerun-if-changed={}", src.join("../../include/rayca_cuda.h").display());
    println!("cargo:rerun-if-env-changed=NVCC");
    println!("cargo:rerun-if-env-changed=RAYCA_CUDA_SRC");
    cmd.args(["-lgomp", "-lz"]);
    let status = cmd.status().unwrap_or_else(|e| panic!("cannot run {nvcc}: {e} (there is no CPU fallback: the crate needs the CUDA toolkit)"));
    if !status.success() {
        panic!("nvcc failed building librayca_cuda.so for sm_100a");
    }
    println!("cargo:rustc-link-search=native={}", out_dir.display());
    println!("cargo:rustc-link-lib=dylib=rayca_cuda");
    // the runtime loader must find the library next to the test binaries too
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", out_dir.display());
}
