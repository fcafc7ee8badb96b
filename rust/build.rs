// UNVERIFIED: written without a Rust toolchain (none exists in the build image); kept in step with INTEGRATION.md
// by tools/sync_rust.py.  See INTEGRATION.md for where each piece goes in the reference tree.
// build.rs -- compiles the CUDA sources for sm_100a and links them.  UNVERIFIED: no cargo in the build image.
use std::{env, path::PathBuf, process::Command};
fn main() {
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let csrc = PathBuf::from("wrs_b200/csrc");           // this repo's whisper_rs_b200/csrc + include/
    let mut objs = vec![];
    for f in ["ctx.cu", "logmel.cu", "decode.cu", "decode_host.cu", "kv_reorder.cu", "no_speech.cu"] {
        let o = out.join(f.replace(".cu", ".o"));
        let ok = Command::new("nvcc")
            .args(["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
                   "-Xcompiler", "-fPIC", "-I", "wrs_b200/include", "-c"])
            .arg(csrc.join(f)).arg("-o").arg(&o).status().unwrap().success();
        assert!(ok, "nvcc failed on {f}");
        objs.push(o);
    }
    let lib = out.join("libwrs_b200.a");
    assert!(Command::new("ar").arg("crs").arg(&lib).args(&objs).status().unwrap().success());
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=static=wrs_b200");
    println!("cargo:rustc-link-lib=dylib=cudart");
    println!("cargo:rustc-link-lib=dylib=stdc++");
    println!("cargo:rerun-if-changed=wrs_b200");
}
