//! `nvllm-kernels`: the Rust side of the drop-in boundary.  build.rs compiles the CUDA sources for sm_100a
//! and links libnvllm.so; ffi_gen.rs is the complete generated `extern "C"` surface, ffi.rs the safe wrappers.
//! COMPILE-UNVERIFIED: the build image has no Rust toolchain.
pub mod ffi;
pub mod ffi_gen;
