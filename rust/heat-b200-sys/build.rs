// UNCOMPILED IN THIS ENVIRONMENT (no cargo / rustc in the image).
// heat-b200-sys/build.rs — builds the CUDA library with nvcc for sm_100a (no other arch, no CPU fallback), exactly as
// heat_b200/csrc/Makefile does: one translation unit per lanes-per-wall tier and kernel family.
use std::{env, path::PathBuf, process::Command};

fn main() {
    let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../../heat_b200/csrc");
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".into());
    let flags = [
        "-O3", "-std=c++17", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a",
        "-Xcompiler", "-fPIC,-ffp-contract=off", "-I../../include", "-I.",
    ];
    let mut objs = vec![];
    let mut cc = |src: &str, obj: &str, extra: &[&str]| {
        let o = out.join(obj);
        let st = Command::new(&nvcc)
            .current_dir(&root)
            .args(flags)
            .args(extra)
            .args(["-c", "-o", o.to_str().unwrap(), src])
            .status()
            .expect("nvcc");
        assert!(st.success(), "nvcc failed on {src}");
        objs.push(o);
    };
    cc("hb_capi.cu", "hb_capi.o", &[]);
    cc("hb_host.cc", "hb_host.o", &[]);
    cc("hb_aux.cu", "hb_aux.o", &[]);
    for tier in [1, 2, 4, 8, 16, 32] {
        // block-tile kernels (hb_face_kernel / hb_march_kernel) and warp-tile kernels (hb_wtile_kernel)
        cc("hb_march_tier.cu", &format!("hb_march_{tier}.o"), &[&format!("-DHB_TIER={tier}")]);
        cc("hb_wtile_tier.cu", &format!("hb_wtile_{tier}.o"), &[&format!("-DHB_TIER={tier}")]);
    }
    let lib = out.join("libheat_b200.so");
    let st = Command::new(&nvcc)
        .args(["-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o"])
        .arg(&lib)
        .args(&objs)
        .args(["-lcudart", "-ldl"]) // NCCL is resolved at run time (dlopen of libnccl.so.2)
        .status()
        .expect("nvcc link");
    assert!(st.success());
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=heat_b200");
    println!("cargo:rerun-if-changed=../../heat_b200/csrc");
    println!("cargo:rerun-if-changed=../../include/heat_b200.h");
}
