//! Builds libminipath_b200.so from the CUDA / C++ sources of this repository with the flags of its Makefile
//! (sm_100a only, FMA contraction off on host and device so that the explicit `mul_add` sites are the only fused ones).
//! Paths: MINIPATH_B200_ROOT (default: two directories up = the repository root), NVCC, CXX.
use std::{env, path::PathBuf, process::Command};

fn run(cmd: &mut Command) {
    let status = cmd.status().unwrap_or_else(|e| panic!("cannot start {:?}: {e}", cmd));
    assert!(status.success(), "{:?} failed", cmd);
}

fn main() {
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    if env::var("CARGO_FEATURE_PREBUILT").is_ok() {
        let dir = env::var("MINIPATH_B200_LIB_DIR").expect("feature `prebuilt` needs MINIPATH_B200_LIB_DIR");
        println!("cargo:rustc-link-search=native={dir}");
        println!("cargo:rustc-link-lib=dylib=minipath_b200");
        return;
    }
    let root = env::var("MINIPATH_B200_ROOT").map(PathBuf::from).unwrap_or_else(|_| {
        PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../..")
    });
    let cs = root.join("minipath_b200/csrc");
    let inc = root.join("include");
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".into());
    let cxx = env::var("CXX").unwrap_or_else(|_| "g++".into());
    let host = ["host_bvh", "host_camera", "host_render", "host_image", "host_group"];
    for name in host {
        run(Command::new(&cxx)
            .args(["-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-c"])
            .arg(format!("-I{}", inc.display()))
            .arg("-I/usr/local/cuda/include") // host_group.cpp: cuda_runtime_api.h (events, streams)
            .arg(cs.join(format!("{name}.cpp")))
            .arg("-o")
            .arg(out.join(format!("{name}.o"))));
    }
    for unit in ["context", "build_device"] {
        run(Command::new(&nvcc)
            .args(["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false",
                   "-prec-div=true", "-prec-sqrt=true", "-Xcompiler", "-fPIC,-ffp-contract=off", "-c"])
            .arg(format!("-I{}", inc.display()))
            .arg(cs.join(format!("{unit}.cu")))
            .arg("-o")
            .arg(out.join(format!("{unit}.o"))));
    }
    let mut link = Command::new(&nvcc);
    link.args(["-shared", "-cudart", "static", "-o"]).arg(out.join("libminipath_b200.so"));
    for name in host {
        link.arg(out.join(format!("{name}.o")));
    }
    link.arg(out.join("context.o")).arg(out.join("build_device.o")).args(["-lpthread", "-lrt", "-ldl"]);
    run(&mut link);
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=minipath_b200");
    println!("cargo:rerun-if-changed={}", cs.display());
    println!("cargo:rerun-if-changed={}", inc.display());
    println!("cargo:rerun-if-env-changed=NVCC");
    println!("cargo:rerun-if-env-changed=MINIPATH_B200_ROOT");
}
