//! Raw bindings to `libbitnet_b200.so` -- one declaration per symbol of `include/bitnet_b200.h`, in the header's
//! order. Source only: the build image has no Rust toolchain, so this file has never been through rustc;
//! `tests/test_abi.py::test_rust_bindings_name_every_header_symbol` keeps the names in step with the header
//! and the library's export table.
//!
//! The six `bitnet_cpp_*` functions are the ones BitNet-rs already binds for the microsoft/BitNet C++ bridge
//! (crates/bitnet-kernels/src/ffi/bridge.rs:17-39): same names, arguments, 0 / -1 + thread-local message.
#![allow(non_camel_case_types)]

use libc::{c_char, c_float, c_int, c_void, size_t};

#[repr(C)]
pub struct bitnet_b200_ctx {
    _private: [u8; 0],
}
#[repr(C)]
pub struct bitnet_b200_weight {
    _private: [u8; 0],
}
#[repr(C)]
pub struct bitnet_b200_model {
    _private: [u8; 0],
}

pub const BITNET_B200_LAYOUT_QK256: c_int = 0;
pub const BITNET_B200_LAYOUT_BITNET32_F16: c_int = 1;
pub const BITNET_B200_SCALE_ABS_CLAMP: c_int = 0;
pub const BITNET_B200_SCALE_RAW: c_int = 1;
pub const BITNET_B200_EPI_STORE: c_int = 0;
pub const BITNET_B200_EPI_RESIDUAL: c_int = 1;
pub const BITNET_B200_EPI_SWIGLU: c_int = 2;

#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct bitnet_b200_model_config {
    pub hidden: c_int,
    pub ffn: c_int,
    pub n_layers: c_int,
    pub n_heads: c_int,
    pub n_kv_heads: c_int,
    pub head_dim: c_int,
    pub vocab: c_int,
    pub max_ctx: c_int,
    pub batch: c_int,
    pub rms_eps: c_float,
    pub layout: c_int,
}

/// One cross-GPU exchange point of the fused tensor-parallel path (include/bitnet_b200.h, Part 5).
#[repr(C)]
#[derive(Clone, Copy)]
pub struct bitnet_b200_tp_xchg {
    pub peer_data: [*const c_void; 8],
    pub peer_flags: [*mut u32; 8],
    pub counters: *mut u32,
    pub n_ranks: c_int,
    pub rank: c_int,
    pub slot: c_int,
    pub phase: c_int,
}

extern "C" {
    // ---- Part 1: the reference's own bridge (ffi/bridge.rs:17-39) ------------------------------------
    pub fn bitnet_cpp_init() -> c_int;
    pub fn bitnet_cpp_cleanup();
    pub fn bitnet_cpp_is_available() -> c_int;
    pub fn bitnet_cpp_matmul_i2s(a: *const i8, b: *const u8, c: *mut c_float, m: c_int, n: c_int, k: c_int) -> c_int;
    pub fn bitnet_cpp_quantize(
        input: *const c_float, input_len: c_int, output: *mut u8, output_len: c_int, scales: *mut c_float,
        scales_len: c_int, qtype: c_int,
    ) -> c_int;
    pub fn bitnet_cpp_get_last_error() -> *const c_char;

    // ---- Part 2: gemv_qk256 and the legacy entry points (host slices) --------------------------------
    pub fn bitnet_b200_gemv_qk256(
        qs: *const u8, qs_len: size_t, x: *const c_float, x_len: size_t, y: *mut c_float, y_len: size_t,
        rows: size_t, cols: size_t, row_stride_bytes: size_t,
    ) -> c_int;
    pub fn bitnet_b200_qk256_gemv(
        output: *mut c_float, out_len: size_t, rows: size_t, cols: size_t, packed: *const u8, packed_len: size_t,
        scales: *const c_float, scales_len: size_t, activations: *const c_float, act_len: size_t,
    ) -> c_int;
    pub fn bitnet_b200_qk256_gemv_scalar(
        output: *mut c_float, out_len: size_t, rows: size_t, cols: size_t, packed: *const u8, packed_len: size_t,
        scales: *const c_float, scales_len: size_t, activations: *const c_float, act_len: size_t,
    ) -> c_int;

    // ---- Part 3: device-resident API (all data pointers are DEVICE pointers unless noted) ------------
    pub fn bitnet_b200_ctx_create(device: c_int, stream: *mut c_void, out: *mut *mut bitnet_b200_ctx) -> c_int;
    pub fn bitnet_b200_ctx_set_stream(ctx: *mut bitnet_b200_ctx, stream: *mut c_void) -> c_int;
    pub fn bitnet_b200_ctx_synchronize(ctx: *mut bitnet_b200_ctx) -> c_int;
    pub fn bitnet_b200_ctx_destroy(ctx: *mut bitnet_b200_ctx);
    pub fn bitnet_b200_weight_create(
        ctx: *mut bitnet_b200_ctx, host_bytes: *const c_void, nbytes: size_t, layout: c_int, rows: size_t,
        cols: size_t, lut: *const i8, scale_policy: c_int, out: *mut *mut bitnet_b200_weight,
    ) -> c_int;
    pub fn bitnet_b200_weight_destroy(w: *mut bitnet_b200_weight);
    pub fn bitnet_b200_weight_rows(w: *const bitnet_b200_weight) -> size_t;
    pub fn bitnet_b200_weight_cols(w: *const bitnet_b200_weight) -> size_t;
    pub fn bitnet_b200_weight_nbytes(w: *const bitnet_b200_weight) -> size_t;
    pub fn bitnet_b200_actquant_i8(
        ctx: *mut bitnet_b200_ctx, x: *const c_float, m: c_int, k: c_int, q: *mut i8, scale: *mut c_float,
    ) -> c_int;
    pub fn bitnet_b200_rmsnorm_actquant(
        ctx: *mut bitnet_b200_ctx, x: *const c_float, gamma: *const c_float, eps: c_float, m: c_int, k: c_int,
        q: *mut i8, scale: *mut c_float, normed: *mut c_float,
    ) -> c_int;
    pub fn bitnet_b200_linear_i8(
        ctx: *mut bitnet_b200_ctx, w: *const bitnet_b200_weight, q: *const i8, act_scale: *const c_float, m: c_int,
        y: *mut c_float, acc_out: *mut i32,
    ) -> c_int;
    pub fn bitnet_b200_linear_i8_swiglu(
        ctx: *mut bitnet_b200_ctx, w: *const bitnet_b200_weight, q: *const i8, act_scale: *const c_float, m: c_int,
        h: *mut c_float,
    ) -> c_int;
    /// No counterpart in the reference: linear_i8 whose split-K scratch is
    /// a static device array of the library (4 MiB + 32 KiB per GPU); allocates nothing; overlapping calls on a GPU race.
    pub fn bitnet_b200_linear_i8_static_scratch(
        ctx: *mut bitnet_b200_ctx, w: *const bitnet_b200_weight, q: *const i8, act_scale: *const c_float, m: c_int,
        y: *mut c_float,
    ) -> c_int;
    pub fn bitnet_b200_fused_linear(
        ctx: *mut bitnet_b200_ctx, w: *const bitnet_b200_weight, x: *const c_float, gamma: *const c_float,
        eps: c_float, m: c_int, epilogue: c_int, y: *mut c_float, residual: *const c_float, q_out: *mut i8,
        scale_out: *mut c_float,
    ) -> c_int;
    pub fn bitnet_b200_linear_f32(
        ctx: *mut bitnet_b200_ctx, w: *const bitnet_b200_weight, x: *const c_float, m: c_int, y: *mut c_float,
    ) -> c_int;
    /// host slices (fills launch_i2s_matmul, crates/bitnet-kernels/src/cuda/quantized_matmul.rs:282-303)
    pub fn bitnet_b200_i2s_matmul_f32(
        act: *const c_float, packed: *const u8, scales: *const c_float, out: *mut c_float, m: c_int, n: c_int,
        k: c_int, block_size: c_int,
    ) -> c_int;

    // ---- Part 4: decode harness ----------------------------------------------------------------------
    pub fn bitnet_b200_model_create(
        ctx: *mut bitnet_b200_ctx, cfg: *const bitnet_b200_model_config, out: *mut *mut bitnet_b200_model,
    ) -> c_int;
    pub fn bitnet_b200_model_destroy(m: *mut bitnet_b200_model);
    pub fn bitnet_b200_model_set_layer(
        m: *mut bitnet_b200_model, layer: c_int, wq: *const c_void, wk: *const c_void, wv: *const c_void,
        wo: *const c_void, wgate: *const c_void, wup: *const c_void, wdown: *const c_void,
        attn_norm: *const c_float, ffn_norm: *const c_float,
    ) -> c_int;
    pub fn bitnet_b200_model_set_embedding(
        m: *mut bitnet_b200_model, emb_f16: *const u16, final_norm: *const c_float,
    ) -> c_int;
    pub fn bitnet_b200_model_set_rope(m: *mut bitnet_b200_model, cos_t: *const c_float, sin_t: *const c_float) -> c_int;
    pub fn bitnet_b200_model_reset(m: *mut bitnet_b200_model) -> c_int;
    pub fn bitnet_b200_model_decode(
        m: *mut bitnet_b200_model, tokens: *const i32, next_tokens: *mut i32, logits_out: *mut c_float,
    ) -> c_int;
    pub fn bitnet_b200_model_decode_device(m: *mut bitnet_b200_model, n_steps: c_int) -> c_int;
    pub fn bitnet_b200_model_set_tokens(m: *mut bitnet_b200_model, tokens: *const i32) -> c_int;
    pub fn bitnet_b200_model_get_tokens(m: *mut bitnet_b200_model, next_tokens: *mut i32) -> c_int;
    pub fn bitnet_b200_model_time_linears(
        m: *mut bitnet_b200_model, reps: c_int, ms_per_pass: *mut c_float, launches_per_pass: *mut c_int,
    ) -> c_int;
    pub fn bitnet_b200_model_launches_per_step(m: *const bitnet_b200_model) -> u64;
    pub fn bitnet_b200_model_weight_bytes(m: *const bitnet_b200_model) -> u64;

    // ---- Part 5: tensor-parallel building blocks (tensor_parallel.rs:123-290 semantics) ---------------
    pub fn bitnet_b200_row_absmax(
        ctx: *mut bitnet_b200_ctx, x: *const c_float, m: c_int, k: c_int, swiglu: c_int, h: *mut c_float,
        absmax: *mut c_float,
    ) -> c_int;
    pub fn bitnet_b200_actquant_i8_absmax(
        ctx: *mut bitnet_b200_ctx, x: *const c_float, absmax: *const c_float, m: c_int, k: c_int, q: *mut i8,
        scale: *mut c_float,
    ) -> c_int;
    pub fn bitnet_b200_residual_acc(
        ctx: *mut bitnet_b200_ctx, acc: *const i32, act_scale: *const c_float, x: *mut c_float, m: c_int, n: c_int,
    ) -> c_int;
    pub fn bitnet_b200_rope_attn_decode(
        ctx: *mut bitnet_b200_ctx, qkv: *mut c_float, cos_t: *const c_float, sin_t: *const c_float,
        kc: *mut c_float, vc: *mut c_float, pos_dev: *const i32, b: c_int, nh: c_int, nkv: c_int, d: c_int,
        max_ctx: c_int, att: *mut c_float,
    ) -> c_int;

    pub fn bitnet_b200_tp_xchg_sum_rmsnorm_actquant(
        ctx: *mut bitnet_b200_ctx, xc: *const bitnet_b200_tp_xchg, act_scale: *const c_float, x: *mut c_float,
        gamma: *const c_float, eps: c_float, m: c_int, n: c_int, q: *mut i8, scale: *mut c_float, normed: *mut c_float,
    ) -> c_int;
    pub fn bitnet_b200_tp_xchg_max_actquant(
        ctx: *mut bitnet_b200_ctx, xc: *const bitnet_b200_tp_xchg, x: *const c_float, m: c_int, k: c_int, q: *mut i8,
        scale: *mut c_float,
    ) -> c_int;

    pub fn bitnet_b200_version() -> *const c_char;
    pub fn bitnet_b200_launch_count() -> u64;
    pub fn bitnet_b200_last_linear_kernel() -> *const c_char;
}

/// The library's thread-local error text (`bitnet_cpp_get_last_error`), empty when none.
pub fn last_error() -> String {
    // SAFETY: the library returns a NUL-terminated thread-local buffer or NULL.
    unsafe {
        let p = bitnet_cpp_get_last_error();
        if p.is_null() {
            String::new()
        } else {
            std::ffi::CStr::from_ptr(p).to_string_lossy().into_owned()
        }
    }
}

#[cfg(feature = "provider")]
pub mod b200;
