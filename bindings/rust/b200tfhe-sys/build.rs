// build.rs -- links libb200tfhe.so; with the feature `build-from-source` it first compiles the CUDA sources for sm_100a with nvcc
// (the same command line as tfhe-rs-odd_b200/csrc/Makefile).  The library needs only cudart at link time: cuTensorMapEncodeTiled is
// resolved at run time through cudaGetDriverEntryPoint.
use std::env;
use std::path::PathBuf;
use std::process::Command;

const SOURCES: &[&str] = &[
    "api", "blind_rotate", "blind_rotate_big", "blind_rotate_tmem", "blind_rotate_lat", "blind_rotate_x16", "blind_rotate_pair", "blind_rotate_mb",
    "keyswitch", "keyswitch_tc", "misc_kernels", "wave_kernels", "gadget_kernels", "wopbs_kernels", "cmux_tree", "noalloc",
    "external_product", "multi_gpu",
];

fn main() {
    let manifest = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap());
    let repo = manifest.join("../../..");
    println!("cargo:rerun-if-env-changed=B200TFHE_LIB_DIR");
    println!("cargo:rerun-if-changed={}", repo.join("include/b200tfhe.h").display());

    let lib_dir = if env::var("CARGO_FEATURE_BUILD_FROM_SOURCE").is_ok() {
        let out = PathBuf::from(env::var("OUT_DIR").unwrap());
        let csrc = repo.join("tfhe-rs-odd_b200/csrc");
        let nvcc = env::var("NVCC").unwrap_or_else(|_| "nvcc".into());
        let mut cmd = Command::new(nvcc);
        cmd.args(["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-shared"]);
        cmd.arg("-I").arg(repo.join("include"));
        for s in SOURCES {
            let f = csrc.join(format!("{s}.cu"));
            println!("cargo:rerun-if-changed={}", f.display());
            cmd.arg(f);
        }
        cmd.arg("-o").arg(out.join("libb200tfhe.so")).arg("-lcudart");
        let status = cmd.status().expect("nvcc not found: set NVCC or build without the feature `build-from-source`");
        assert!(status.success(), "nvcc failed");
        out
    } else {
        env::var("B200TFHE_LIB_DIR").map(PathBuf::from).unwrap_or_else(|_| repo.join("tfhe-rs-odd_b200"))
    };
    println!("cargo:rustc-link-search=native={}", lib_dir.display());
    println!("cargo:rustc-link-lib=dylib=b200tfhe");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", lib_dir.display());
}
