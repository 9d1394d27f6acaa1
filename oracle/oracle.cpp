/*
 * oracle.cpp -- CPU restatement of the BitNet-rs I2_S ternary-linear hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.h). Parity status: PINNED against the
 * reference's own known-answer tests (tests/test_oracle_golden.py) and against
 * the reference's ggml-quants.c built into oracle/_ref/ (tests/test_oracle_ref.py).
 *
 * Every function cites the reference file:line it follows; paths are relative to
 * the BitNet-rs tree. Scalar code is compiled with -ffp-contract=off so that
 * `acc += w * x` rounds twice, as the Rust source does; the SIMD restatements use
 * the same intrinsics, in the same order, as the Rust `std::arch` code.
 */
#include "oracle.h"

#include <immintrin.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

static thread_local std::string g_err;

static int fail(const std::string& s) {
    g_err = s;
    return -1;
}

extern "C" const char* oracle_last_error(void) { return g_err.empty() ? nullptr : g_err.c_str(); }
extern "C" int oracle_has_avx2(void) {
    return __builtin_cpu_supports("avx2") && __builtin_cpu_supports("fma");
}
extern "C" int oracle_has_avx512(void) {
    return __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw");
}

static inline size_t div_ceil(size_t a, size_t b) { return (a + b - 1) / b; }

/* ===================================================================== *
 * KernelProvider::matmul_i2s                                             *
 * ===================================================================== */

static int check_matmul_dims(size_t a_len, size_t b_len, size_t c_len, size_t m, size_t n,
                             size_t k) {
    /* crates/bitnet-kernels/src/cpu/fallback.rs:48-63 (same checks in x86.rs:24-42,188-206) */
    char buf[160];
    if (a_len != m * k) {
        snprintf(buf, sizeof buf, "Matrix A dimension mismatch: expected %zu, got %zu", m * k, a_len);
        return fail(buf);
    }
    if (b_len != k * n) {
        snprintf(buf, sizeof buf, "Matrix B dimension mismatch: expected %zu, got %zu", k * n, b_len);
        return fail(buf);
    }
    if (c_len != m * n) {
        snprintf(buf, sizeof buf, "Matrix C dimension mismatch: expected %zu, got %zu", m * n, c_len);
        return fail(buf);
    }
    return 0;
}

/* crates/bitnet-kernels/src/cpu/fallback.rs:39-83 */
extern "C" int ref_matmul_i2s_fallback(const int8_t* a, size_t a_len, const uint8_t* b,
                                       size_t b_len, float* c, size_t c_len, size_t m, size_t n,
                                       size_t k) {
    if (check_matmul_dims(a_len, b_len, c_len, m, n, k)) return -1;
    std::fill(c, c + c_len, 0.0f);
    for (size_t i = 0; i < m; ++i) {
        for (size_t j = 0; j < n; ++j) {
            float sum = 0.0f;
            for (size_t l = 0; l < k; ++l) {
                float a_val = (float)a[i * k + l];
                float b_val = (float)b[l * n + j];
                sum += a_val * b_val;
            }
            c[i * n + j] = sum;
        }
    }
    return 0;
}

/* crates/bitnet-kernels/src/cpu/x86.rs:417-517 (8x8 tiles, K step 32) */
__attribute__((target("avx2"))) static void matmul_i2s_avx2_impl(const int8_t* a,
                                                                 const uint8_t* b, float* c,
                                                                 size_t m, size_t n, size_t k) {
    std::fill(c, c + m * n, 0.0f);
    const size_t BM = 8, BN = 8, BK = 32;
    for (size_t i = 0; i < m; i += BM) {
        for (size_t j = 0; j < n; j += BN) {
            __m256 acc[8][8];
            for (auto& r : acc)
                for (auto& v : r) v = _mm256_setzero_ps();
            for (size_t l = 0; l < k; l += BK) {
                size_t k_end = std::min(l + BK, k);
                size_t k_len = k_end - l;
                for (size_t ii = 0; ii < std::min(BM, m - i); ++ii) {
                    const int8_t* a_row = a + (i + ii) * k + l;
                    __m256i a_vec;
                    if (k_len >= 32) {
                        a_vec = _mm256_loadu_si256((const __m256i*)a_row);
                    } else {
                        int8_t temp[32] = {0};
                        memcpy(temp, a_row, k_len);
                        a_vec = _mm256_loadu_si256((const __m256i*)temp);
                    }
                    for (size_t jj = 0; jj < std::min(BN, n - j); ++jj) {
                        uint8_t b_col[32] = {0};
                        for (size_t kk = 0; kk < k_len; ++kk)
                            if (l + kk < k) b_col[kk] = b[(l + kk) * n + (j + jj)];
                        __m256i b_vec = _mm256_loadu_si256((const __m256i*)b_col);
                        __m128i a_lo128 = _mm256_castsi256_si128(a_vec);
                        __m128i a_hi128 = _mm256_extracti128_si256(a_vec, 1);
                        __m128i b_lo128 = _mm256_castsi256_si128(b_vec);
                        __m128i b_hi128 = _mm256_extracti128_si256(b_vec, 1);
                        __m256i a_lo = _mm256_cvtepi8_epi16(a_lo128);
                        __m256i a_hi = _mm256_cvtepi8_epi16(a_hi128);
                        __m256i b_lo = _mm256_cvtepu8_epi16(b_lo128);
                        __m256i b_hi = _mm256_cvtepu8_epi16(b_hi128);
                        __m256i prod_lo = _mm256_madd_epi16(a_lo, b_lo);
                        __m256i prod_hi = _mm256_madd_epi16(a_hi, b_hi);
                        __m256i sum = _mm256_add_epi32(prod_lo, prod_hi);
                        acc[ii][jj] = _mm256_add_ps(acc[ii][jj], _mm256_cvtepi32_ps(sum));
                    }
                }
            }
            for (size_t ii = 0; ii < std::min(BM, m - i); ++ii) {
                for (size_t jj = 0; jj < std::min(BN, n - j); ++jj) {
                    __m256 sv = acc[ii][jj];
                    __m128 hi = _mm256_extractf128_ps(sv, 1);
                    __m128 lo = _mm256_castps256_ps128(sv);
                    __m128 quad = _mm_add_ps(hi, lo);
                    __m128 dual = _mm_add_ps(quad, _mm_movehl_ps(quad, quad));
                    __m128 single = _mm_add_ss(dual, _mm_shuffle_ps(dual, dual, 0x55));
                    c[(i + ii) * n + (j + jj)] += _mm_cvtss_f32(single);
                }
            }
        }
    }
}

extern "C" int ref_matmul_i2s_avx2(const int8_t* a, size_t a_len, const uint8_t* b, size_t b_len,
                                   float* c, size_t c_len, size_t m, size_t n, size_t k) {
    if (check_matmul_dims(a_len, b_len, c_len, m, n, k)) return -1;
    if (!oracle_has_avx2()) return fail("AVX2 not available on this host");
    matmul_i2s_avx2_impl(a, b, c, m, n, k);
    return 0;
}

/* crates/bitnet-kernels/src/cpu/x86.rs:247-321 (16x16 tiles, K step 64) */
__attribute__((target("avx512f,avx512bw,avx2"))) static void matmul_i2s_avx512_impl(
    const int8_t* a, const uint8_t* b, float* c, size_t m, size_t n, size_t k) {
    std::fill(c, c + m * n, 0.0f);
    const size_t BM = 16, BN = 16, BK = 64;
    alignas(64) __m512 accv[16 * 16];
    for (size_t i = 0; i < m; i += BM) {
        for (size_t j = 0; j < n; j += BN) {
            for (auto& v : accv) v = _mm512_setzero_ps();
            for (size_t l = 0; l < k; l += BK) {
                size_t k_end = std::min(l + BK, k);
                size_t k_len = k_end - l;
                for (size_t ii = 0; ii < std::min(BM, m - i); ++ii) {
                    const int8_t* a_row = a + (i + ii) * k + l;
                    __m512i a_vec;
                    if (k_len >= 64) {
                        a_vec = _mm512_loadu_si512((const void*)a_row);
                    } else {
                        int8_t temp[64] = {0};
                        memcpy(temp, a_row, k_len);
                        a_vec = _mm512_loadu_si512((const void*)temp);
                    }
                    for (size_t jj = 0; jj < std::min(BN, n - j); ++jj) {
                        uint8_t b_col[64] = {0};
                        for (size_t kk = 0; kk < k_len; ++kk)
                            if (l + kk < k) b_col[kk] = b[(l + kk) * n + (j + jj)];
                        __m512i b_vec = _mm512_loadu_si512((const void*)b_col);
                        __m256i a_lo256 = _mm512_castsi512_si256(a_vec);
                        __m256i a_hi256 = _mm512_extracti64x4_epi64(a_vec, 1);
                        __m256i b_lo256 = _mm512_castsi512_si256(b_vec);
                        __m256i b_hi256 = _mm512_extracti64x4_epi64(b_vec, 1);
                        __m512i a_lo = _mm512_cvtepi8_epi16(a_lo256);
                        __m512i a_hi = _mm512_cvtepi8_epi16(a_hi256);
                        __m512i b_lo = _mm512_cvtepu8_epi16(b_lo256);
                        __m512i b_hi = _mm512_cvtepu8_epi16(b_hi256);
                        __m512i prod_lo = _mm512_madd_epi16(a_lo, b_lo);
                        __m512i prod_hi = _mm512_madd_epi16(a_hi, b_hi);
                        __m512i sum = _mm512_add_epi32(prod_lo, prod_hi);
                        accv[ii * BN + jj] =
                            _mm512_add_ps(accv[ii * BN + jj], _mm512_cvtepi32_ps(sum));
                    }
                }
            }
            for (size_t ii = 0; ii < std::min(BM, m - i); ++ii)
                for (size_t jj = 0; jj < std::min(BN, n - j); ++jj)
                    c[(i + ii) * n + (j + jj)] += _mm512_reduce_add_ps(accv[ii * BN + jj]);
        }
    }
}

extern "C" int ref_matmul_i2s_avx512(const int8_t* a, size_t a_len, const uint8_t* b,
                                     size_t b_len, float* c, size_t c_len, size_t m, size_t n,
                                     size_t k) {
    if (check_matmul_dims(a_len, b_len, c_len, m, n, k)) return -1;
    if (!oracle_has_avx512()) return fail("AVX-512 not available on this host");
    matmul_i2s_avx512_impl(a, b, c, m, n, k);
    return 0;
}

/* ===================================================================== *
 * KernelProvider::quantize (weight quantisers)                           *
 * crates/bitnet-kernels/src/cpu/fallback.rs:85-290                       *
 * (C++ twin: src/ffi/cpp_bridge.cpp:166-217; CUDA twin: bitnet_kernels.cu:99-196)
 * ===================================================================== */
extern "C" int ref_quantize(const float* in, size_t in_len, uint8_t* out, size_t out_len,
                            float* scales, size_t scales_len, int qtype) {
    if (qtype < 0 || qtype > 2) return fail("Invalid quantization type");
    const size_t BLOCK = qtype == 0 ? 32 : (qtype == 1 ? 64 : 128);
    const char* nm = qtype == 0 ? "I2_S" : (qtype == 1 ? "TL1" : "TL2");
    const size_t num_blocks = div_ceil(in_len, BLOCK);
    char buf[160];
    if (out_len < in_len / 4) {
        snprintf(buf, sizeof buf, "Output buffer too small for %s: expected %zu, got %zu", nm,
                 in_len / 4, out_len);
        return fail(buf);
    }
    if (scales_len < num_blocks) {
        snprintf(buf, sizeof buf, "Scales buffer too small: expected %zu, got %zu", num_blocks,
                 scales_len);
        return fail(buf);
    }
    static const float LUT1[4] = {-1.0f, -0.33f, 0.33f, 1.0f};
    static const float LUT2[4] = {-1.2f, -0.4f, 0.4f, 1.2f};
    for (size_t blk = 0; blk < num_blocks; ++blk) {
        size_t start = blk * BLOCK, end = std::min(start + BLOCK, in_len);
        float max_val = 0.0f;
        for (size_t i = start; i < end; ++i) max_val = fmaxf(max_val, fabsf(in[i]));
        float scale = max_val > 1e-8f ? max_val / 1.5f : 1.0f;
        scales[blk] = scale;
        for (size_t i = start; i < end; ++i) {
            float normalized = in[i] / scale;
            uint8_t q;
            if (qtype == 0) {
                q = normalized > 0.5f ? 1 : (normalized < -0.5f ? 3 : 0); /* map C */
            } else {
                const float* lut = qtype == 1 ? LUT1 : LUT2;
                size_t best = 0;
                float best_d = fabsf(normalized - lut[0]);
                for (size_t t = 1; t < 4; ++t) {
                    float d = fabsf(normalized - lut[t]);
                    if (d < best_d) {
                        best_d = d;
                        best = t;
                    }
                }
                q = (uint8_t)best;
            }
            size_t byte_idx = i / 4, bit_off = (i % 4) * 2;
            if (byte_idx < out_len) out[byte_idx] |= (uint8_t)(q << bit_off);
        }
    }
    return 0;
}

/* ===================================================================== *
 * per-token absmax int8 activation quantiser                             *
 * crates/bitnet-kernels/src/cpu/quantize.rs:32-55                        *
 * ===================================================================== */
extern "C" int ref_quantize_symmetric_i8(const float* in, size_t len, int bits, int8_t* out,
                                         float* scale_out) {
    if (bits < 2 || bits > 8) return fail("bits must be in [2, 8]");
    const float qmax = (float)((1 << (bits - 1)) - 1);
    float abs_max = 0.0f;
    for (size_t i = 0; i < len; ++i) abs_max = fmaxf(abs_max, fabsf(in[i])); /* f32::max drops NaN */
    if (abs_max == 0.0f) {
        memset(out, 0, len);
        *scale_out = 0.0f;
        return 0;
    }
    const float scale = abs_max / qmax;
    const float inv_scale = 1.0f / scale;
    for (size_t i = 0; i < len; ++i) {
        float r = roundf(in[i] * inv_scale); /* Rust f32::round: half away from zero */
        if (r != r) {                        /* NaN: clamp keeps NaN, `as i8` gives 0 */
            out[i] = 0;
            continue;
        }
        r = r < -qmax ? -qmax : (r > qmax ? qmax : r);
        out[i] = (int8_t)r;
    }
    *scale_out = scale;
    return 0;
}

extern "C" void ref_dequantize_symmetric_i8(const int8_t* in, size_t len, float scale, float* out) {
    for (size_t i = 0; i < len; ++i) out[i] = (float)in[i] * scale;
}

/* ===================================================================== *
 * QK256                                                                  *
 * ===================================================================== */
static const size_t QK256_BLOCK = 256, QK256_PACKED = 64;

/* crates/bitnet-quantization/src/i2s_qk256.rs:139-146 (map A; ggml-quants.c:62) */
extern "C" float ref_code_to_f32(uint8_t code) {
    static const float LUT[4] = {-2.0f, -1.0f, 1.0f, 2.0f};
    return LUT[code & 3];
}

/* i2s_qk256.rs:159-168 */
extern "C" void ref_unpack_qk256_block(const uint8_t* qs64, uint8_t* codes) {
    for (size_t i = 0; i < QK256_PACKED; ++i) {
        uint8_t b = qs64[i];
        codes[i * 4 + 0] = b & 0x03;
        codes[i * 4 + 1] = (b >> 2) & 0x03;
        codes[i * 4 + 2] = (b >> 4) & 0x03;
        codes[i * 4 + 3] = (b >> 6) & 0x03;
    }
}

/* i2s_qk256.rs:196-274 (BITNET_QUANT_SANITY probe omitted: diagnostics only) */
extern "C" float ref_gemv_qk256_row(const uint8_t* qs_row, const float* x, size_t cols) {
    const size_t blocks = div_ceil(cols, QK256_BLOCK);
    float acc = 0.0f;
    uint8_t codes[256];
    size_t col = 0;
    for (size_t blk = 0; blk < blocks; ++blk) {
        ref_unpack_qk256_block(qs_row + blk * QK256_PACKED, codes);
        size_t take = std::min(QK256_BLOCK, cols - col);
        for (size_t j = 0; j < take; ++j) {
            float w = ref_code_to_f32(codes[j]);
            acc += w * x[col + j];
        }
        col += take;
        if (col >= cols) break;
    }
    return acc;
}

static int check_gemv_dims(const char* tag, size_t qs_len, size_t x_len, size_t y_len,
                           size_t rows, size_t cols, size_t stride) {
    /* i2s_qk256.rs:301-312 / i2s_qk256_avx2.rs:264-274 */
    char buf[160];
    if (y_len != rows) {
        snprintf(buf, sizeof buf, "%s: y_out length %zu != rows %zu", tag, y_len, rows);
        return fail(buf);
    }
    if (x_len < cols) {
        snprintf(buf, sizeof buf, "%s: x length %zu < cols %zu", tag, x_len, cols);
        return fail(buf);
    }
    if (qs_len < rows * stride) {
        snprintf(buf, sizeof buf, "%s: data too short: %zu < %zu", tag, qs_len, rows * stride);
        return fail(buf);
    }
    return 0;
}

/* i2s_qk256.rs:293-321 */
extern "C" int ref_gemv_qk256_scalar(const uint8_t* qs, size_t qs_len, const float* x,
                                     size_t x_len, float* y, size_t y_len, size_t rows,
                                     size_t cols, size_t stride) {
    if (check_gemv_dims("I2S_QK256", qs_len, x_len, y_len, rows, cols, stride)) return -1;
    if (stride != div_ceil(cols, QK256_BLOCK) * QK256_PACKED)
        return fail("I2S_QK256: row bytes mismatch"); /* debug_assert at :203-210 */
    for (size_t r = 0; r < rows; ++r) y[r] = ref_gemv_qk256_row(qs + r * stride, x, cols);
    return 0;
}

/* i2s_qk256_avx2.rs:42-61 */
__attribute__((target("avx2,fma"))) static inline __m256 decode_8_weights_avx2(
    uint8_t byte0, uint8_t byte1, __m256i shifts, __m256i mask_03, __m256i two, __m256i one) {
    int packed = (int)byte0 | ((int)byte1 << 16);
    __m256i bc = _mm256_set1_epi32(packed);
    __m256i codes = _mm256_and_si256(_mm256_srlv_epi32(bc, shifts), mask_03);
    __m256i shifted = _mm256_sub_epi32(codes, two);
    __m256i corr = _mm256_and_si256(_mm256_srli_epi32(codes, 1), one);
    return _mm256_cvtepi32_ps(_mm256_add_epi32(shifted, corr));
}

/* i2s_qk256_avx2.rs:81-229 */
__attribute__((target("avx2,fma"))) static float gemv_qk256_row_avx2(const uint8_t* qs_row,
                                                                     const float* x,
                                                                     size_t cols) {
    const size_t blocks = div_ceil(cols, QK256_BLOCK);
    const __m256i shifts = _mm256_setr_epi32(0, 2, 4, 6, 16, 18, 20, 22);
    const __m256i mask_03 = _mm256_set1_epi32(3);
    const __m256i two = _mm256_set1_epi32(2);
    const __m256i one = _mm256_set1_epi32(1);
    __m256 acc0 = _mm256_setzero_ps(), acc1 = acc0, acc2 = acc0, acc3 = acc0;
    float scalar_acc = 0.0f;
    size_t col = 0;
    for (size_t bi = 0; bi < blocks; ++bi) {
        const uint8_t* blk = qs_row + bi * QK256_PACKED;
        size_t take = std::min(QK256_BLOCK, cols - col);
        if (bi + 1 < blocks) {
            _mm_prefetch((const char*)(blk + QK256_PACKED), _MM_HINT_T0);
            _mm_prefetch((const char*)(x + col + QK256_BLOCK), _MM_HINT_T0);
            _mm_prefetch((const char*)(x + col + QK256_BLOCK + 16), _MM_HINT_T0);
        }
        size_t j = 0;
        while (j + 32 <= take) {
            size_t pi = j / 4;
            __m256 w0 = decode_8_weights_avx2(blk[pi], blk[pi + 1], shifts, mask_03, two, one);
            __m256 w1 = decode_8_weights_avx2(blk[pi + 2], blk[pi + 3], shifts, mask_03, two, one);
            __m256 w2 = decode_8_weights_avx2(blk[pi + 4], blk[pi + 5], shifts, mask_03, two, one);
            __m256 w3 = decode_8_weights_avx2(blk[pi + 6], blk[pi + 7], shifts, mask_03, two, one);
            size_t xj = col + j;
            __m256 x0 = _mm256_loadu_ps(x + xj);
            __m256 x1 = _mm256_loadu_ps(x + xj + 8);
            __m256 x2 = _mm256_loadu_ps(x + xj + 16);
            __m256 x3 = _mm256_loadu_ps(x + xj + 24);
            acc0 = _mm256_fmadd_ps(w0, x0, acc0);
            acc1 = _mm256_fmadd_ps(w1, x1, acc1);
            acc2 = _mm256_fmadd_ps(w2, x2, acc2);
            acc3 = _mm256_fmadd_ps(w3, x3, acc3);
            j += 32;
        }
        while (j + 8 <= take) {
            size_t pi = j / 4;
            __m256 w = decode_8_weights_avx2(blk[pi], blk[pi + 1], shifts, mask_03, two, one);
            acc0 = _mm256_fmadd_ps(w, _mm256_loadu_ps(x + col + j), acc0);
            j += 8;
        }
        while (j < take) {
            uint8_t code = (blk[j / 4] >> ((j % 4) * 2)) & 3;
            float w = code == 0 ? -2.0f : (code == 1 ? -1.0f : (code == 2 ? 1.0f : 2.0f));
            scalar_acc += w * x[col + j];
            ++j;
        }
        col += take;
        if (col >= cols) break;
    }
    __m256 sum01 = _mm256_add_ps(acc0, acc1);
    __m256 sum23 = _mm256_add_ps(acc2, acc3);
    __m256 acc = _mm256_add_ps(sum01, sum23);
    __m128 hi = _mm256_extractf128_ps(acc, 1);
    __m128 lo = _mm256_castps256_ps128(acc);
    __m128 s128 = _mm_add_ps(hi, lo);
    __m128 s64 = _mm_hadd_ps(s128, s128);
    __m128 s32 = _mm_hadd_ps(s64, s64);
    return _mm_cvtss_f32(s32) + scalar_acc;
}

/* i2s_qk256_avx2.rs:254-295 */
extern "C" int ref_gemv_qk256_avx2(const uint8_t* qs, size_t qs_len, const float* x, size_t x_len,
                                   float* y, size_t y_len, size_t rows, size_t cols,
                                   size_t stride) {
    if (check_gemv_dims("AVX2", qs_len, x_len, y_len, rows, cols, stride)) return -1;
    if (!oracle_has_avx2()) return fail("AVX2 not available on this host");
    if (stride != div_ceil(cols, QK256_BLOCK) * QK256_PACKED)
        return fail("AVX2: row bytes mismatch");
    for (size_t r = 0; r < rows; ++r) {
        if (r + 1 < rows) _mm_prefetch((const char*)(qs + (r + 1) * stride), _MM_HINT_T0);
        y[r] = gemv_qk256_row_avx2(qs + r * stride, x, cols);
    }
    return 0;
}

/* i2s_qk256.rs:346-372: runtime dispatch AVX2 -> scalar */
extern "C" int ref_gemv_qk256(const uint8_t* qs, size_t qs_len, const float* x, size_t x_len,
                              float* y, size_t y_len, size_t rows, size_t cols, size_t stride) {
    if (oracle_has_avx2())
        return ref_gemv_qk256_avx2(qs, qs_len, x, x_len, y, y_len, rows, cols, stride);
    return ref_gemv_qk256_scalar(qs, qs_len, x, x_len, y, y_len, rows, cols, stride);
}

/* Rows partitioned over host threads. The reference's gemv is single-threaded
 * (SURVEY §2.3); this wrapper is the "all host cores" variant of the CPU baseline. */
extern "C" int ref_gemv_qk256_mt(const uint8_t* qs, size_t qs_len, const float* x, size_t x_len,
                                 float* y, size_t y_len, size_t rows, size_t cols, size_t stride,
                                 int threads) {
    if (check_gemv_dims("I2S_QK256", qs_len, x_len, y_len, rows, cols, stride)) return -1;
    if (threads <= 1) return ref_gemv_qk256(qs, qs_len, x, x_len, y, y_len, rows, cols, stride);
    std::vector<std::thread> pool;
    std::vector<int> rc((size_t)threads, 0);
    size_t per = div_ceil(rows, (size_t)threads);
    for (int t = 0; t < threads; ++t) {
        size_t r0 = std::min(rows, (size_t)t * per), r1 = std::min(rows, r0 + per);
        if (r0 == r1) continue;
        pool.emplace_back([=, &rc] {
            rc[(size_t)t] = ref_gemv_qk256(qs + r0 * stride, (r1 - r0) * stride, x, x_len, y + r0,
                                           r1 - r0, r1 - r0, cols, stride);
        });
    }
    for (auto& th : pool) th.join();
    for (int v : rc)
        if (v) return fail("worker failed");
    return 0;
}

/* I2SQk256NoScale::new, i2s_qk256.rs:85-107 (+-128 B tolerance) */
extern "C" int ref_i2s_qk256_new_check(size_t rows, size_t cols, size_t qs_len,
                                       size_t* row_stride_bytes) {
    size_t stride = div_ceil(cols, QK256_BLOCK) * QK256_PACKED;
    size_t expected = rows * stride;
    size_t diff = qs_len > expected ? qs_len - expected : expected - qs_len;
    if (diff > 128) {
        char buf[256];
        snprintf(buf, sizeof buf,
                 "I2SQk256NoScale: data size mismatch: got %zu bytes, expected %zu for %zux%zu matrix.",
                 qs_len, expected, rows, cols);
        return fail(buf);
    }
    if (row_stride_bytes) *row_stride_bytes = stride;
    return 0;
}

/* crates/bitnet-quantization/src/qk256_dispatch.rs:60-99 (map D, bench-only) */
extern "C" int ref_qk256_gemv_scalar_legacy(float* out, size_t rows, size_t cols,
                                            const uint8_t* packed, size_t packed_len,
                                            const float* scales, size_t scales_len,
                                            const float* act) {
    if (cols % 256 != 0) return fail("Cols must be multiple of QK256");
    size_t bpr = cols / 256;
    if (packed_len != rows * cols / 4) return fail("Packed weight size mismatch");
    if (scales_len != rows * bpr) return fail("Scales length mismatch");
    static const float MAPD[4] = {-1.0f, 0.0f, 1.0f, -1.0f};
    for (size_t r = 0; r < rows; ++r) {
        float row_sum = 0.0f;
        for (size_t b = 0; b < bpr; ++b) {
            size_t gb = r * bpr + b;
            float scale = scales[gb];
            size_t byte_off = gb * 256 / 4, act_off = b * 256;
            float block_sum = 0.0f;
            for (size_t e = 0; e < 256; ++e) {
                uint8_t two = (packed[byte_off + e / 4] >> ((e % 4) * 2)) & 3;
                block_sum += MAPD[two] * act[act_off + e];
            }
            row_sum += block_sum * scale;
        }
        out[r] = row_sum;
    }
    return 0;
}

/* ===================================================================== *
 * BitNet32-F16 decode -- crates/bitnet-models/src/quant/i2s.rs           *
 * ===================================================================== */

/* IEEE binary16 -> binary32, exact (half::f16::to_f32; checked against the
 * reference's ggml-quants.c fp16_to_fp32 in tests/test_oracle_ref.py). */
extern "C" float ref_f16_to_f32(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu;
    uint32_t mant = h & 0x3ffu;
    uint32_t f;
    if (exp == 0) {
        if (mant == 0) {
            f = sign;
        } else {
            int e = -1;
            do {
                ++e;
                mant <<= 1;
            } while ((mant & 0x400u) == 0);
            mant &= 0x3ffu;
            f = sign | (uint32_t)(127 - 15 - e) << 23 | (mant << 13);
        }
    } else if (exp == 31) {
        f = sign | 0x7f800000u | (mant << 13);
    } else {
        f = sign | ((exp + 127 - 15) << 23) | (mant << 13);
    }
    float r;
    memcpy(&r, &f, 4);
    return r;
}

/* binary32 -> binary16, round-to-nearest-even (half::f16::from_f32) */
extern "C" uint16_t ref_f32_to_f16(float fl) {
    uint32_t x;
    memcpy(&x, &fl, 4);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t exp = (x >> 23) & 0xffu;
    uint32_t man = x & 0x7fffffu;
    if (exp == 255) return (uint16_t)(sign | 0x7c00u | (man ? (0x200u | (man >> 13)) : 0));
    int32_t half_exp = (int32_t)exp - 127 + 15;
    if (half_exp >= 31) return (uint16_t)(sign | 0x7c00u);
    if (half_exp <= 0) {
        if (14 - half_exp > 24) return (uint16_t)sign;
        man |= 0x800000u;
        uint32_t shift = (uint32_t)(14 - half_exp);
        uint32_t half_man = man >> shift;
        uint32_t round_bit = 1u << (shift - 1);
        if ((man & round_bit) != 0 && (man & (3 * round_bit - 1)) != 0) half_man += 1;
        return (uint16_t)(sign | half_man);
    }
    uint32_t half_man = man >> 13;
    uint32_t round_bit = 0x1000u;
    uint16_t r = (uint16_t)(sign | ((uint32_t)half_exp << 10) | half_man);
    if ((man & round_bit) != 0 && (man & (3 * round_bit - 1)) != 0) r += 1;
    return r;
}

static const float I2S_SCALE_MIN = 1e-3f, I2S_SCALE_MAX = 1e3f; /* quant/i2s.rs:15-16 */

static inline float i2s_effective_scale(uint16_t scale_bits, int inv, float k) {
    /* quant/i2s.rs:76-94 (= :155-171 with cfg) */
    float s = fabsf(ref_f16_to_f32(scale_bits));
    if (inv) s = (s < 1e-8f) ? 1.0f : 1.0f / s;
    s *= k;
    if (s == s) s = s < I2S_SCALE_MIN ? I2S_SCALE_MIN : (s > I2S_SCALE_MAX ? I2S_SCALE_MAX : s);
    return s;
}

/* quant/i2s.rs:66-140 (i2s_dequant_block; I2SMapping::Sym) */
extern "C" void ref_i2s_dequant_block(float* dst, const uint8_t* qbits, size_t n,
                                      uint16_t scale_bits, int inv, float k) {
    static const float LUT[4] = {-2.0f, -1.0f, 1.0f, 2.0f};
    float s = i2s_effective_scale(scale_bits, inv, k);
    float sl[4] = {s * LUT[0], s * LUT[1], s * LUT[2], s * LUT[3]};
    size_t chunks = n / 4, rem = n % 4;
    for (size_t c = 0; c < chunks; ++c) {
        uint8_t b = qbits[c];
        dst[c * 4 + 0] = sl[b & 3];
        dst[c * 4 + 1] = sl[(b >> 2) & 3];
        dst[c * 4 + 2] = sl[(b >> 4) & 3];
        dst[c * 4 + 3] = sl[(b >> 6) & 3];
    }
    if (rem) {
        uint8_t b = qbits[chunks];
        for (size_t i = 0; i < rem; ++i) dst[chunks * 4 + i] = sl[(b >> (i * 2)) & 3];
    }
}

/* quant/i2s.rs:205-209 */
extern "C" size_t ref_i2s_expected_bytes(size_t rows, size_t cols, size_t block) {
    return rows * div_ceil(cols, block) * (div_ceil(block, 4) + 2);
}

/* quant/i2s.rs:211-214 */
extern "C" size_t ref_i2s_infer_block_size(size_t bytes, size_t rows, size_t cols) {
    const size_t cands[4] = {256, 128, 64, 32};
    for (size_t b : cands)
        if (ref_i2s_expected_bytes(rows, cols, b) == bytes) return b;
    return 0;
}

/* quant/i2s.rs:621-692 (dequantize_to_f32_with_block_and_cfg) */
extern "C" int ref_dequantize_i2s_with_block(const uint8_t* bytes, size_t bytes_len, size_t rows,
                                             size_t cols, size_t block, int inv, float k,
                                             float* out) {
    size_t bpr = div_ceil(cols, block);
    size_t per_block = div_ceil(block, 4) + 2;
    if (bytes_len != rows * bpr * per_block) {
        char buf[96];
        snprintf(buf, sizeof buf, "I2_S: internal size mismatch for block=%zu", block);
        return fail(buf);
    }
    std::fill(out, out + rows * cols, 0.0f);
    size_t off = 0;
    for (size_t r = 0; r < rows; ++r) {
        size_t c = 0;
        for (size_t b = 0; b < bpr; ++b) {
            size_t n = std::min(cols - c, block);
            if (n == 0) break;
            size_t qlen = (n + 3) / 4;
            if (off + qlen + 2 > bytes_len) return fail("I2_S: buffer bounds exceeded");
            const uint8_t* q = bytes + off;
            off += qlen;
            uint16_t sb = (uint16_t)(bytes[off] | (bytes[off + 1] << 8));
            off += 2;
            ref_i2s_dequant_block(out + r * cols + c, q, n, sb, inv, k);
            c += n;
            if (c >= cols) break;
        }
    }
    return 0;
}

/* quant/i2s.rs:591-619 (dequantize_to_f32_with_cfg). The partial-data zero-fill
 * fallback (:606-617) is a loader robustness path outside the hot path: -2. */
extern "C" int ref_dequantize_i2s(const uint8_t* bytes, size_t bytes_len, size_t rows, size_t cols,
                                  int inv, float k, float* out) {
    size_t block = 256;
    if (bytes_len != ref_i2s_expected_bytes(rows, cols, block)) {
        size_t b = ref_i2s_infer_block_size(bytes_len, rows, cols);
        if (b == 0) {
            fail("I2_S: byte length mismatch (partial-block fallback not restated)");
            return -2;
        }
        block = b;
    }
    return ref_dequantize_i2s_with_block(bytes, bytes_len, rows, cols, block, inv, k, out);
}

/* ===================================================================== *
 * layout detection                                                       *
 * ===================================================================== */

/* crates/bitnet-quantization/src/lib.rs:86-90 */
extern "C" size_t ref_qk256_tolerance_bytes(size_t expected) {
    double tol = (double)expected * 0.001;
    double c = std::ceil(tol);
    return (size_t)(c < 8.0 ? 8.0 : c);
}

/* crates/bitnet-models/src/formats/gguf/types.rs:868-1066 */
extern "C" int ref_detect_i2s_flavor(size_t available, int has_sibling, size_t nelems, int strict) {
    size_t blocks32 = div_ceil(nelems, 32), blocks256 = div_ceil(nelems, 256);
    size_t split_need = blocks32 * 8, inline_need = blocks32 * 10, qk256_need = blocks256 * 64;
    size_t tol = strict ? 8 : ref_qk256_tolerance_bytes(std::min(split_need, qk256_need));
    auto ad = [](size_t a, size_t b) { return a > b ? a - b : b - a; };
    size_t d_split = ad(available, split_need), d_inline = ad(available, inline_need),
           d_qk = ad(available, qk256_need);
    if (d_qk == 0) return 2;
    if (d_inline == 0) return 0;
    if (d_split == 0 && has_sibling) return 1;
    if (d_qk <= tol) return 2;
    if (has_sibling && d_split <= tol) return 1;
    if (d_inline <= tol) return 0;
    if (d_split <= tol) return 1;
    fail("I2_S: no valid flavor detected");
    return -1;
}

/* ===================================================================== *
 * packed + scaled matmul, map C                                          *
 * crates/bitnet-kernels/src/cpu/quantized_matmul.rs                      *
 * ===================================================================== */
static inline int8_t decode_i2s_c(uint8_t bits) { /* :19-27 */
    switch (bits & 3) {
        case 0: return 0;
        case 1: return 1;
        case 3: return -1;
        default: return 0;
    }
}

extern "C" uint8_t ref_pack_i2s(const int8_t vals[4]) { /* :30-41 */
    uint8_t byte = 0;
    for (int i = 0; i < 4; ++i) {
        uint8_t code = vals[i] == 1 ? 1 : (vals[i] == -1 ? 3 : 0);
        byte |= (uint8_t)(code << (i * 2));
    }
    return byte;
}

static int validate_matmul_args(size_t act_len, size_t w_len, size_t scales_len, size_t out_len,
                                size_t m, size_t n, size_t k, size_t bs) { /* :204-255 */
    char buf[160];
    if (bs == 0) return fail("block_size must be > 0");
    if (m == 0 || n == 0 || k == 0) {
        snprintf(buf, sizeof buf, "dimensions must be > 0: m=%zu, n=%zu, k=%zu", m, n, k);
        return fail(buf);
    }
    size_t packed_k = div_ceil(k, 4), nbk = div_ceil(k, bs);
    if (act_len < m * k) {
        snprintf(buf, sizeof buf, "activations too small: expected %zu, got %zu", m * k, act_len);
        return fail(buf);
    }
    if (w_len < packed_k * n) {
        snprintf(buf, sizeof buf, "weights_packed too small: expected %zu, got %zu", packed_k * n,
                 w_len);
        return fail(buf);
    }
    if (scales_len < n * nbk) {
        snprintf(buf, sizeof buf, "scales too small: expected %zu, got %zu", n * nbk, scales_len);
        return fail(buf);
    }
    if (out_len < m * n) {
        snprintf(buf, sizeof buf, "output too small: expected %zu, got %zu", m * n, out_len);
        return fail(buf);
    }
    return 0;
}

/* :57-96 */
extern "C" int ref_i2s_matmul_f32(const float* act, size_t act_len, const uint8_t* w, size_t w_len,
                                  const float* scales, size_t scales_len, float* out,
                                  size_t out_len, size_t m, size_t n, size_t k, size_t bs) {
    if (validate_matmul_args(act_len, w_len, scales_len, out_len, m, n, k, bs)) return -1;
    size_t packed_k = div_ceil(k, 4), nbk = div_ceil(k, bs);
    std::fill(out, out + out_len, 0.0f);
    for (size_t row = 0; row < m; ++row) {
        const float* a_row = act + row * k;
        for (size_t col = 0; col < n; ++col) {
            float acc = 0.0f;
            for (size_t blk = 0; blk < nbk; ++blk) {
                size_t s0 = blk * bs, s1 = std::min(s0 + bs, k);
                float scale = scales[col * nbk + blk];
                for (size_t idx = s0; idx < s1; ++idx) {
                    uint8_t bits = (w[col * packed_k + idx / 4] >> ((idx % 4) * 2)) & 3;
                    float wv = (float)decode_i2s_c(bits) * scale;
                    acc += a_row[idx] * wv;
                }
            }
            out[row * n + col] = acc;
        }
    }
    return 0;
}

/* :155-200 */
extern "C" int ref_i2s_matmul_blocked(const float* act, size_t act_len, const uint8_t* w,
                                      size_t w_len, const float* scales, size_t scales_len,
                                      float* out, size_t out_len, size_t m, size_t n, size_t k,
                                      size_t bs) {
    if (validate_matmul_args(act_len, w_len, scales_len, out_len, m, n, k, bs)) return -1;
    if (bs > 256) return fail("block_size > 256 unsupported (w_blk is [i8; 256])");
    size_t packed_k = div_ceil(k, 4), nbk = div_ceil(k, bs);
    std::fill(out, out + out_len, 0.0f);
    for (size_t blk = 0; blk < nbk; ++blk) {
        size_t s0 = blk * bs, s1 = std::min(s0 + bs, k);
        for (size_t col = 0; col < n; ++col) {
            float scale = scales[col * nbk + blk];
            int8_t w_blk[256] = {0};
            for (size_t idx = s0; idx < s1; ++idx)
                w_blk[idx - s0] =
                    decode_i2s_c((w[col * packed_k + idx / 4] >> ((idx % 4) * 2)) & 3);
            for (size_t row = 0; row < m; ++row) {
                float acc = 0.0f;
                const float* a_row = act + row * k + s0;
                for (size_t i = 0; i < s1 - s0; ++i) acc += a_row[i] * (float)w_blk[i];
                out[row * n + col] += acc * scale;
            }
        }
    }
    return 0;
}

/* :105-148 */
extern "C" int ref_dequantize_and_matmul(const float* act, size_t act_len, const uint8_t* w,
                                         size_t w_len, const float* scales, size_t scales_len,
                                         float* out, size_t out_len, size_t m, size_t n, size_t k,
                                         size_t bs) {
    if (validate_matmul_args(act_len, w_len, scales_len, out_len, m, n, k, bs)) return -1;
    size_t packed_k = div_ceil(k, 4), nbk = div_ceil(k, bs);
    std::vector<float> wf(k * n, 0.0f);
    for (size_t col = 0; col < n; ++col)
        for (size_t blk = 0; blk < nbk; ++blk) {
            size_t s0 = blk * bs, s1 = std::min(s0 + bs, k);
            float scale = scales[col * nbk + blk];
            for (size_t idx = s0; idx < s1; ++idx) {
                uint8_t bits = (w[col * packed_k + idx / 4] >> ((idx % 4) * 2)) & 3;
                wf[idx * n + col] = (float)decode_i2s_c(bits) * scale;
            }
        }
    std::fill(out, out + out_len, 0.0f);
    for (size_t row = 0; row < m; ++row)
        for (size_t col = 0; col < n; ++col) {
            float acc = 0.0f;
            for (size_t idx = 0; idx < k; ++idx) acc += act[row * k + idx] * wf[idx * n + col];
            out[row * n + col] = acc;
        }
    return 0;
}

/* ===================================================================== *
 * map B -- crates/bitnet-quantization/src/utils.rs:57-91                 *
 * ===================================================================== */
extern "C" void ref_pack_2bit_values(const int8_t* vals, size_t len, uint8_t* out) {
    for (size_t c = 0; c < div_ceil(len, 4); ++c) {
        uint8_t byte = 0;
        for (size_t i = 0; i < 4 && c * 4 + i < len; ++i) {
            int v = vals[c * 4 + i];
            v = v < -2 ? -2 : (v > 1 ? 1 : v);
            byte |= (uint8_t)((v + 2) << (i * 2));
        }
        out[c] = byte;
    }
}

extern "C" void ref_unpack_2bit_values(const uint8_t* packed, size_t packed_len, size_t output_len,
                                       int8_t* out) {
    size_t cnt = 0;
    for (size_t b = 0; b < packed_len; ++b)
        for (int i = 0; i < 4; ++i) {
            if (cnt >= output_len) break;
            out[cnt++] = (int8_t)(((packed[b] >> (i * 2)) & 3) - 2);
        }
}

/* ===================================================================== *
 * RMSNorm                                                                *
 * ===================================================================== */

/* crates/bitnet-kernels/src/cpu/layer_norm.rs:118-137,162-169 */
extern "C" int ref_rms_norm(const float* in, size_t in_len, const float* gamma, size_t norm_size,
                            float eps, float* out) {
    if (norm_size == 0 || in_len == 0) return fail("empty input");
    if (in_len % norm_size != 0) return fail("input length is not a multiple of norm_size");
    if (!(eps > 0.0f) || !std::isfinite(eps)) return fail("eps must be positive and finite");
    size_t batch = in_len / norm_size;
    for (size_t b = 0; b < batch; ++b) {
        const float* s = in + b * norm_size;
        float* o = out + b * norm_size;
        double sum = 0.0;
        for (size_t i = 0; i < norm_size; ++i) {
            double v = (double)s[i];
            sum += v * v;
        }
        float rms = (float)(sum / (double)norm_size);
        float inv_rms = 1.0f / sqrtf(rms + eps);
        for (size_t i = 0; i < norm_size; ++i) o[i] = s[i] * inv_rms * gamma[i];
    }
    return 0;
}

/* crates/bitnet-kernels/src/cpu/fusion.rs:122-155 */
extern "C" int ref_fused_rmsnorm_linear(const float* in, size_t n, const float* weight,
                                        size_t w_len, const float* gamma, float eps, float* out) {
    if (n == 0) return fail("empty input");
    if (w_len % n != 0) return fail("dimension mismatch");
    size_t out_dim = w_len / n;
    float sum_sq = 0.0f;
    for (size_t i = 0; i < n; ++i) sum_sq += in[i] * in[i];
    float rms = sqrtf(sum_sq / (float)n + eps);
    float inv_rms = 1.0f / rms;
    for (size_t o = 0; o < out_dim; ++o) {
        float acc = 0.0f;
        for (size_t i = 0; i < n; ++i) acc += weight[o * n + i] * (in[i] * gamma[i] * inv_rms);
        out[o] = acc;
    }
    return 0;
}

/* ===================================================================== *
 * composed int8 linears (see oracle.h)                                   *
 * ===================================================================== */
static bool is_map_a(const int8_t lut[4]) {
    return lut[0] == -2 && lut[1] == -1 && lut[2] == 1 && lut[3] == 2;
}

extern "C" int oracle_linear_qk256_i8(const uint8_t* qs, size_t rows, size_t cols, size_t stride,
                                      const int8_t* q, const float* act_scale, size_t m,
                                      const int8_t lut[4], int32_t* acc, float* y) {
    if (stride != div_ceil(cols, 256) * 64) return fail("row stride mismatch");
    std::vector<float> x(cols), yf(rows);
    for (size_t i = 0; i < m; ++i) {
        if (is_map_a(lut)) {
            /* the reference kernel itself: exact under |sum| < 2^24 */
            for (size_t c = 0; c < cols; ++c) x[c] = (float)q[i * cols + c];
            if (ref_gemv_qk256(qs, rows * stride, x.data(), cols, yf.data(), rows, rows, cols,
                               stride))
                return -1;
            for (size_t r = 0; r < rows; ++r) acc[i * rows + r] = (int32_t)yf[r];
        } else {
            for (size_t r = 0; r < rows; ++r) {
                int32_t s = 0;
                const uint8_t* row = qs + r * stride;
                for (size_t c = 0; c < cols; ++c)
                    s += (int32_t)lut[(row[c / 4] >> ((c % 4) * 2)) & 3] * (int32_t)q[i * cols + c];
                acc[i * rows + r] = s;
            }
        }
        if (y)
            for (size_t r = 0; r < rows; ++r) y[i * rows + r] = act_scale[i] * (float)acc[i * rows + r];
    }
    return 0;
}

extern "C" int oracle_linear_bitnet32_i8(const uint8_t* bytes, size_t rows, size_t cols,
                                         const int8_t* q, const float* act_scale, size_t m,
                                         const int8_t lut[4], int scale_policy, int32_t* acc_blk,
                                         float* y) {
    if (cols % 32 != 0) return fail("BitNet32-F16 linear requires cols % 32 == 0");
    size_t nb = cols / 32;
    for (size_t i = 0; i < m; ++i)
        for (size_t r = 0; r < rows; ++r) {
            float sum = 0.0f;
            for (size_t b = 0; b < nb; ++b) {
                const uint8_t* blk = bytes + (r * nb + b) * 10;
                int32_t s = 0;
                for (size_t e = 0; e < 32; ++e)
                    s += (int32_t)lut[(blk[e / 4] >> ((e % 4) * 2)) & 3] *
                         (int32_t)q[i * cols + b * 32 + e];
                uint16_t sb = (uint16_t)(blk[8] | (blk[9] << 8));
                float sc = scale_policy == 0 ? i2s_effective_scale(sb, 0, 1.0f) : ref_f16_to_f32(sb);
                if (acc_blk) acc_blk[(i * rows + r) * nb + b] = s;
                sum += sc * (float)s;
            }
            if (y) y[i * rows + r] = act_scale[i] * sum;
        }
    return 0;
}

extern "C" int oracle_linear_bitnet32_dequant(const uint8_t* bytes, size_t rows, size_t cols,
                                              const int8_t* q, const float* act_scale, size_t m,
                                              float* y) {
    std::vector<float> w(rows * cols);
    if (ref_dequantize_i2s_with_block(bytes, rows * (cols / 32) * 10, rows, cols, 32, 0, 1.0f,
                                      w.data()))
        return -1;
    for (size_t i = 0; i < m; ++i)
        for (size_t r = 0; r < rows; ++r) {
            double s = 0.0;
            for (size_t c = 0; c < cols; ++c)
                s += (double)w[r * cols + c] * (double)(act_scale[i] * (float)q[i * cols + c]);
            y[i * rows + r] = (float)s;
        }
    return 0;
}

/* ===================================================================== *
 * shared synthetic data: splitmix64, counter-based          *
 * ===================================================================== */
static inline uint64_t splitmix64_at(uint64_t seed, uint64_t i) {
    uint64_t z = seed + (i + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

extern "C" void oracle_fill_bytes(uint8_t* dst, size_t n, uint64_t seed) {
    for (size_t i = 0; i < n; i += 8) {
        uint64_t v = splitmix64_at(seed, i / 8);
        for (size_t j = 0; j < 8 && i + j < n; ++j) dst[i + j] = (uint8_t)(v >> (8 * j));
    }
}

static inline double u01(uint64_t v) { return (double)(v >> 11) * (1.0 / 9007199254740992.0); }

extern "C" void oracle_fill_uniform(float* dst, size_t n, uint64_t seed, float lo, float hi) {
    for (size_t i = 0; i < n; ++i)
        dst[i] = (float)((double)lo + ((double)hi - (double)lo) * u01(splitmix64_at(seed, i)));
}

extern "C" void oracle_fill_normal(float* dst, size_t n, uint64_t seed, float sigma) {
    for (size_t i = 0; i < n; i += 2) {
        double u1 = u01(splitmix64_at(seed, i)), u2 = u01(splitmix64_at(seed, i + 1));
        if (u1 < 1e-300) u1 = 1e-300;
        double r = std::sqrt(-2.0 * std::log(u1));
        dst[i] = (float)(sigma * r * std::cos(6.283185307179586 * u2));
        if (i + 1 < n) dst[i + 1] = (float)(sigma * r * std::sin(6.283185307179586 * u2));
    }
}
