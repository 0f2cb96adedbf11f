// build.rs of rtgpu-sys — builds librtgpu.so with nvcc (sm_100a) from the backend's sources and links it.
// UNCOMPILED (no Rust toolchain where it was written); the recipe mirrors ray_tracing_in_one_weekend_b200/build.py,
// which IS what builds and tests the library.
//
//   RTGPU_LIB_DIR=<dir holding a prebuilt librtgpu.so>   link that one, build nothing
//   RTGPU_SRC_DIR=<checkout of the backend>              (default: two levels above this crate) build from csrc/
//   NVCC=<path to nvcc>                                   (default /usr/local/cuda/bin/nvcc)
use std::env;
use std::path::{Path, PathBuf};
use std::process::Command;

fn run(cmd: &mut Command) {
    let status = cmd.status().unwrap_or_else(|e| panic!("cannot run {:?}: {}", cmd, e));
    assert!(status.success(), "{:?} failed", cmd);
}

fn main() {
    println!("cargo:rerun-if-env-changed=RTGPU_LIB_DIR");
    println!("cargo:rerun-if-env-changed=RTGPU_SRC_DIR");
    if let Ok(dir) = env::var("RTGPU_LIB_DIR") {
        println!("cargo:rustc-link-search=native={dir}");
        println!("cargo:rustc-link-lib=dylib=rtgpu");
        return;
    }
    let manifest = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap());
    let src_root = env::var("RTGPU_SRC_DIR").map(PathBuf::from).unwrap_or_else(|_| manifest.join("../../.."));
    let csrc = src_root.join("ray_tracing_in_one_weekend_b200/csrc");
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let nvcc = env::var("NVCC").unwrap_or_else(|_| "/usr/local/cuda/bin/nvcc".into());
    println!("cargo:rerun-if-changed={}", csrc.display());

    // host translation units: every float operation rounded individually, in the reference's operation order
    let host_flags = ["-std=c++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC"];
    // CUDA translation units: sm_100a only, no FMA contraction, IEEE division / square root, no flush-to-zero
    let cuda_flags = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false",
                      "-prec-div=true", "-prec-sqrt=true", "-ftz=false", "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math"];
    let mut objs: Vec<PathBuf> = Vec::new();
    let obj = |name: &str| -> PathBuf { out.join(format!("{name}.o")) };
    for s in ["scene_api", "demo_scene", "bvh_build", "png_writer"] {
        let o = obj(s);
        run(Command::new("g++").args(host_flags).arg("-c").arg(csrc.join(format!("{s}.cpp"))).arg("-o").arg(&o));
        objs.push(o);
    }
    for s in ["rtgpu", "rt_comm", "rt_static", "gpu_bvh"] {
        let o = obj(s);
        run(Command::new(&nvcc).args(cuda_flags).arg("-c").arg(csrc.join(format!("{s}.cu"))).arg("-o").arg(&o).current_dir(&csrc));
        objs.push(o);
    }
    // one instantiation of the CTA-wavefront kernel per scene feature set (csrc/cta_inst.cu)
    let mut cta = |f: u32, extra: &[&str], suffix: &str| {
        let o = obj(&format!("cta_f{f}{suffix}"));
        run(Command::new(&nvcc).args(cuda_flags).arg(format!("-DRT_CTA_F={f}")).args(extra).arg("-c").arg(csrc.join("cta_inst.cu"))
            .arg("-o").arg(&o).current_dir(&csrc));
        objs.push(o);
    };
    for f in [0, 4, 8, 9, 11, 15] {
        cta(f, &[], "");
    }
    for f in [0, 4] {
        cta(f, &["-DRT_CTA_CLS_MASK=0x3f", "-DRT_CTA_SUFFIX=_m3f"], "_m3f");
    }
    for f in [0, 8, 9, 11] {
        cta(f, &["-DRT_CTA_FLAT=1", "-DRT_CTA_SUFFIX=_flat"], "_flat");
    }
    let lib = out.join("librtgpu.so");
    run(Command::new(&nvcc).arg("-shared").arg("-o").arg(&lib).args(&objs)
        .args(["-gencode", "arch=compute_100a,code=sm_100a", "-lz", "-ldl", "-Xlinker"])
        .arg(format!("--version-script={}", Path::new(&csrc).join("exports.map").display())));
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=rtgpu");
}
