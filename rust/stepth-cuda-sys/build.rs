// build.rs -- compiles stepth_b200/csrc/stepth_cuda.cu for sm_100a with nvcc (through the `cc` crate's CUDA mode) and
// links it; with `--features bindgen` the declarations are generated from include/stepth_cuda.h.
// NOTE: written as source only -- the build image has no Rust toolchain.
use std::{env, path::PathBuf};

fn main() {
    let root = PathBuf::from(env::var("CARGO_MANIFEST_DIR").unwrap()).join("../..");
    let csrc = root.join("stepth_b200/csrc");
    println!("cargo:rerun-if-changed={}", csrc.display());
    println!("cargo:rerun-if-changed={}", root.join("include/stepth_cuda.h").display());

    if let Ok(dir) = env::var("STEPTH_CUDA_LIB_DIR") {
        // use a prebuilt libstepth_cuda.so (python -c "import __graft_entry__ as g; g.build()")
        println!("cargo:rustc-link-search=native={dir}");
        println!("cargo:rustc-link-lib=dylib=stepth_cuda");
    } else {
        cc::Build::new()
            .cuda(true)
            .cudart("static")
            .flag("-gencode").flag("arch=compute_100a,code=sm_100a")   // B200 only: no multi-arch fatbin, no PTX fallback
            .flag("-lineinfo").flag("-O3").flag("-std=c++17")
            .flag("-Xcompiler").flag("-ffp-contract=off,-fno-fast-math") // the f32 resample must round like Rust (no FMA)
            .include(root.join("include"))
            .file(csrc.join("stepth_cuda.cu"))
            .compile("stepth_cuda");
        println!("cargo:rustc-link-lib=dylib=stdc++");
    }

    #[cfg(feature = "bindgen")]
    {
        let out = PathBuf::from(env::var("OUT_DIR").unwrap()).join("bindings.rs");
        bindgen::Builder::default()
            .header(root.join("include/stepth_cuda.h").to_str().unwrap())
            .allowlist_function("stepth_.*")
            .allowlist_type("stepth_.*")
            .allowlist_var("STEPTH_.*")
            .generate()
            .expect("bindgen failed on include/stepth_cuda.h")
            .write_to_file(out)
            .unwrap();
    }
}
