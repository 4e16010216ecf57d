// build.rs — compiles the sm_100a backend with nvcc (through the in-tree Makefile) and links libtinyr.so.
// UNBUILT ON THIS BOX (no Rust toolchain); the same Makefile is what __graft_entry__.build() runs.
use std::path::PathBuf;
use std::process::Command;

fn main() {
    let root = PathBuf::from(env!("CARGO_MANIFEST_DIR")).join("..");
    let csrc = root.join("tiny_renderer_b200").join("csrc");
    // nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false ... (see csrc/Makefile; -fmad=false is required
    // for bit parity with the reference's unfused f32 arithmetic)
    let status = Command::new("make").arg("-C").arg(&csrc).arg("all").status().expect("failed to run make");
    assert!(status.success(), "building libtinyr.so failed");
    let libdir = root.join("tiny_renderer_b200");
    println!("cargo:rustc-link-search=native={}", libdir.display());
    println!("cargo:rustc-link-lib=dylib=tinyr");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", libdir.display());
    println!("cargo:rerun-if-changed={}", csrc.display());
    println!("cargo:rerun-if-changed={}", root.join("include").join("tinyr.h").display());
}
