// Builds libb2stark.so with the same command openvm_b200/build.py runs, or links a prebuilt one when B2STARK_LIB_DIR is set.
use std::{env, path::PathBuf, process::Command};

fn main() {
    println!("cargo:rerun-if-env-changed=B2STARK_LIB_DIR");
    if let Ok(dir) = env::var("B2STARK_LIB_DIR") {
        println!("cargo:rustc-link-search=native={dir}");
        println!("cargo:rustc-link-lib=dylib=b2stark");
        return;
    }
    let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../..");
    let csrc = root.join("openvm_b200/csrc");
    let out = PathBuf::from(env::var("OUT_DIR").unwrap());
    let mut objects = Vec::new();
    for entry in std::fs::read_dir(&csrc).expect("csrc") {
        let path = entry.unwrap().path();
        let ext = path.extension().and_then(|e| e.to_str()).unwrap_or("");
        if ext != "cu" && ext != "cpp" {
            continue;
        }
        println!("cargo:rerun-if-changed={}", path.display());
        let obj = out.join(path.file_stem().unwrap()).with_extension("o");
        let mut cmd = Command::new("nvcc");
        if ext == "cu" {
            cmd.args(["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
                      "--expt-relaxed-constexpr"]);
        } else {
            cmd.args(["-O3", "-std=c++17", "-Xcompiler", "-fPIC,-mavx2"]);
        }
        let status = cmd.arg("-c").arg(&path).arg("-o").arg(&obj).status().expect("nvcc");
        assert!(status.success(), "nvcc failed on {}", path.display());
        objects.push(obj);
    }
    let lib = out.join("libb2stark.so");
    let status = Command::new("nvcc").arg("-shared").args(&objects).args(["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart", "-o"]).arg(&lib).status().expect("nvcc link");
    assert!(status.success(), "link failed");
    println!("cargo:rustc-link-search=native={}", out.display());
    println!("cargo:rustc-link-lib=dylib=b2stark");
}
