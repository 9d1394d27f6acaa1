/*
 * oracle.h -- CPU restatement of the BitNet-rs I2_S ternary-linear hot path.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under oracle/ is part of the shipped
 * product: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs may load this library, and only as the checker or as
 * the timed CPU baseline. The product (libbitnet_b200.so) never links it.
 *
 * Parity status: PINNED. Every function below restates one reference function
 * (file:line cited at its definition in oracle.cpp, paths relative to the
 * BitNet-rs tree) and is checked in tests/test_oracle_golden.py against every
 * known-answer vector the reference's own tests hold for it,
 * plus the reference's own C file ggml-quants.c compiled to oracle/_ref/.
 * The *composed* int8 linear (act-quant -> packed matmul -> scale) has no
 * reference test; it is defined here as the composition of pinned pieces.
 */
#ifndef BITNET_ORACLE_H
#define BITNET_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* oracle_last_error(void);
int oracle_has_avx2(void);
int oracle_has_avx512(void);

/* ---- KernelProvider::matmul_i2s (unpacked u8 codes, [k x n] row-major) ---- */
int ref_matmul_i2s_fallback(const int8_t* a, size_t a_len, const uint8_t* b, size_t b_len,
                            float* c, size_t c_len, size_t m, size_t n, size_t k);
int ref_matmul_i2s_avx2(const int8_t* a, size_t a_len, const uint8_t* b, size_t b_len,
                        float* c, size_t c_len, size_t m, size_t n, size_t k);
int ref_matmul_i2s_avx512(const int8_t* a, size_t a_len, const uint8_t* b, size_t b_len,
                          float* c, size_t c_len, size_t m, size_t n, size_t k);

/* ---- KernelProvider::quantize (weight quantisers; ORs into `out`) ---- */
/* qtype: 0 = I2_S (32), 1 = TL1 (64), 2 = TL2 (128) */
int ref_quantize(const float* in, size_t in_len, uint8_t* out, size_t out_len,
                 float* scales, size_t scales_len, int qtype);

/* ---- per-token absmax int8 activation quantiser ---- */
int ref_quantize_symmetric_i8(const float* in, size_t len, int bits, int8_t* out, float* scale);
void ref_dequantize_symmetric_i8(const int8_t* in, size_t len, float scale, float* out);

/* ---- QK256 ---- */
float ref_code_to_f32(uint8_t code);
void ref_unpack_qk256_block(const uint8_t* qs64, uint8_t* codes256);
float ref_gemv_qk256_row(const uint8_t* qs_row, const float* x, size_t cols);
int ref_gemv_qk256_scalar(const uint8_t* qs, size_t qs_len, const float* x, size_t x_len,
                          float* y, size_t y_len, size_t rows, size_t cols, size_t stride);
int ref_gemv_qk256_avx2(const uint8_t* qs, size_t qs_len, const float* x, size_t x_len,
                        float* y, size_t y_len, size_t rows, size_t cols, size_t stride);
int ref_gemv_qk256(const uint8_t* qs, size_t qs_len, const float* x, size_t x_len,
                   float* y, size_t y_len, size_t rows, size_t cols, size_t stride);
/* rows partitioned over `threads` std::threads; each row computed by the dispatching kernel */
int ref_gemv_qk256_mt(const uint8_t* qs, size_t qs_len, const float* x, size_t x_len,
                      float* y, size_t y_len, size_t rows, size_t cols, size_t stride, int threads);
int ref_i2s_qk256_new_check(size_t rows, size_t cols, size_t qs_len, size_t* row_stride_bytes);
/* legacy bench-only kernel (map D, per-block f32 scales) */
int ref_qk256_gemv_scalar_legacy(float* out, size_t rows, size_t cols, const uint8_t* packed,
                                 size_t packed_len, const float* scales, size_t scales_len,
                                 const float* act);

/* ---- BitNet32-F16 (and 64/128/256-block inline-f16 variants) ---- */
float ref_f16_to_f32(uint16_t bits);
uint16_t ref_f32_to_f16(float f);
void ref_i2s_dequant_block(float* dst, const uint8_t* qbits, size_t n, uint16_t scale_bits,
                           int inv, float k);
size_t ref_i2s_expected_bytes(size_t rows, size_t cols, size_t block);
size_t ref_i2s_infer_block_size(size_t bytes, size_t rows, size_t cols); /* 0 = None */
int ref_dequantize_i2s_with_block(const uint8_t* bytes, size_t bytes_len, size_t rows, size_t cols,
                                  size_t block, int inv, float k, float* out);
int ref_dequantize_i2s(const uint8_t* bytes, size_t bytes_len, size_t rows, size_t cols,
                       int inv, float k, float* out);

/* ---- layout detection ---- */
size_t ref_qk256_tolerance_bytes(size_t expected);
/* returns 0 BitNet32F16, 1 Split32WithSibling, 2 GgmlQk256NoScale, -1 no flavour */
int ref_detect_i2s_flavor(size_t available, int has_scale_sibling, size_t nelems, int strict);

/* ---- packed + scaled matmul (map C) ---- */
uint8_t ref_pack_i2s(const int8_t vals[4]);
int ref_i2s_matmul_f32(const float* act, size_t act_len, const uint8_t* w, size_t w_len,
                       const float* scales, size_t scales_len, float* out, size_t out_len,
                       size_t m, size_t n, size_t k, size_t block_size);
int ref_i2s_matmul_blocked(const float* act, size_t act_len, const uint8_t* w, size_t w_len,
                           const float* scales, size_t scales_len, float* out, size_t out_len,
                           size_t m, size_t n, size_t k, size_t block_size);
int ref_dequantize_and_matmul(const float* act, size_t act_len, const uint8_t* w, size_t w_len,
                              const float* scales, size_t scales_len, float* out, size_t out_len,
                              size_t m, size_t n, size_t k, size_t block_size);

/* ---- map B helpers ---- */
void ref_unpack_2bit_values(const uint8_t* packed, size_t packed_len, size_t output_len, int8_t* out);
void ref_pack_2bit_values(const int8_t* vals, size_t len, uint8_t* out);

/* ---- RMSNorm ---- */
int ref_rms_norm(const float* in, size_t in_len, const float* gamma, size_t norm_size, float eps,
                 float* out);
int ref_fused_rmsnorm_linear(const float* in, size_t n, const float* weight, size_t w_len,
                             const float* gamma, float eps, float* out);

/* ---- composed int8 linears (definition = composition of the pinned pieces) ---- */
/* acc[m,n] = (int32) ref_gemv_qk256(qs, (f32)q[m,:])[n] when lut == map A, else the
 * integer LUT dot; y[m,n] = act_scale[m] * (f32)acc[m,n]. */
int oracle_linear_qk256_i8(const uint8_t* qs, size_t rows, size_t cols, size_t stride,
                           const int8_t* q, const float* act_scale, size_t m,
                           const int8_t lut[4], int32_t* acc, float* y);
/* per-32-block integer dots; y = act_scale * sum_blk s[n,blk] * (f32)acc[m,n,blk], block order,
 * fp32; s = clamp(|f16|,1e-3,1e3) (scale_policy 0) or raw f16 (scale_policy 1).
 * acc_blk may be NULL; when given it is [m, rows, cols/32]. */
int oracle_linear_bitnet32_i8(const uint8_t* bytes, size_t rows, size_t cols,
                              const int8_t* q, const float* act_scale, size_t m,
                              const int8_t lut[4], int scale_policy, int32_t* acc_blk, float* y);
/* dequantise-then-f32 reference for the 1e-3 check: sum_k W[n,k] * (act_scale*q[m,k]) in f64 */
int oracle_linear_bitnet32_dequant(const uint8_t* bytes, size_t rows, size_t cols,
                                   const int8_t* q, const float* act_scale, size_t m, float* y);

/* ---- shared synthetic data generator: splitmix64 ---- */
void oracle_fill_bytes(uint8_t* dst, size_t n, uint64_t seed);
void oracle_fill_normal(float* dst, size_t n, uint64_t seed, float sigma);
void oracle_fill_uniform(float* dst, size_t n, uint64_t seed, float lo, float hi);

/* ---- 2B-4T-shaped decoder built from the pieces above (model.cpp) ---- */
typedef struct {
    int hidden, ffn, n_layers, n_heads, n_kv_heads, head_dim, vocab, max_ctx;
    float rms_eps;
    int layout; /* 0 = QK256, 1 = BitNet32-F16 */
} oracle_model_config;
void* oracle_model_create(const oracle_model_config* cfg, int n_seqs);
void oracle_model_destroy(void* m);
int oracle_model_set_layer(void* m, int layer, const uint8_t* wq, const uint8_t* wk,
                           const uint8_t* wv, const uint8_t* wo, const uint8_t* wgate,
                           const uint8_t* wup, const uint8_t* wdown, const float* attn_norm,
                           const float* ffn_norm);
int oracle_model_set_embedding(void* m, const uint16_t* emb_f16, const float* final_norm);
int oracle_model_set_rope(void* m, const float* cos_t, const float* sin_t);
void oracle_model_reset(void* m);
int oracle_model_step(void* m, int seq, int token, float* logits_out, float* hidden_out,
                      int threads);

#ifdef __cplusplus
}
#endif
#endif
