// Builds libposeidon_b200.so with this repository's build.sh (nvcc -gencode arch=compute_100a,code=sm_100a)
// and links it.  POSEIDON_B200_DIR defaults to the repository root three levels up.
use std::{env, path::PathBuf, process::Command};

fn main() {
    let root = env::var("POSEIDON_B200_DIR")
        .map(PathBuf::from)
        .unwrap_or_else(|_| PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../../.."));
    let status = Command::new("bash").arg(root.join("build.sh")).status().expect("failed to run build.sh");
    assert!(status.success(), "poseidon-b200 build failed (needs nvcc with sm_100a support)");
    let lib = root.join("sponge_b200/lib");
    println!("cargo:rustc-link-search=native={}", lib.display());
    println!("cargo:rustc-link-lib=dylib=poseidon_b200");
    println!("cargo:rustc-link-arg=-Wl,-rpath,{}", lib.display());
    println!("cargo:rerun-if-changed={}", root.join("include/poseidon_b200.h").display());
    println!("cargo:rerun-if-env-changed=POSEIDON_B200_DIR");
}
