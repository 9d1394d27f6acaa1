/*
 * bitnet_b200.h -- C ABI of libbitnet_b200.so: the B200 (sm_100a) implementation of
 * BitNet-rs's I2_S ternary linear layer (QK256 and BitNet32-F16 layouts) and of the
 * per-token absmax int8 activation quantiser that feeds it.
 *
 * Part 1 is a drop-in for the C bridge the reference already binds
 *   (crates/bitnet-kernels/src/ffi/bridge.rs:17-39, ffi/cpp_bridge.cpp:34-297):
 *   same six symbols, same argument meaning, same 0 / -1 + thread-local error string.
 * Part 2 is the C twin of the free function the live transformer calls
 *   (crates/bitnet-quantization/src/i2s_qk256.rs:346-372).
 * Part 3 is additive: device-resident weights and activations, no per-call copies.
 *
 * Conventions (all parts): plain pointers and sizes, no C++/torch types; the caller owns
 * every buffer for the duration of the call; every entry point is re-entrant and may be
 * called concurrently from arbitrary host threads; int return 0 = ok, -1 = error, message
 * via bitnet_cpp_get_last_error() (thread-local). There is NO CPU fallback: without a
 * CUDA device every compute entry point fails with "no CUDA device".
 *
 * Paths cited below are relative to the BitNet-rs tree.
 */
#ifndef BITNET_B200_H
#define BITNET_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------ *
 * Part 1 -- the reference's kernel bridge (KernelProvider seam)                          *
 * ------------------------------------------------------------------------------------ */

/* replaces ffi/cpp_bridge.cpp:34-53 (bound at ffi/bridge.rs:18). Reference-counted:
 * several FfiKernel instances may coexist (bridge.rs:165-174,257-263). 0 = ok. */
int bitnet_cpp_init(void);
/* replaces ffi/cpp_bridge.cpp:287-297 (bridge.rs:19) */
void bitnet_cpp_cleanup(void);
/* replaces ffi/cpp_bridge.cpp:59-65 (bridge.rs:20): non-zero iff an sm_100 device is usable */
int bitnet_cpp_is_available(void);
/* replaces ffi/cpp_bridge.cpp:73-120 (bridge.rs:21-28); semantics of
 * KernelProvider::matmul_i2s (crates/bitnet-kernels/src/lib.rs:42-50, cpu/fallback.rs:39-83):
 *   a: m*k int8 row-major; b: k*n uint8 row-major, ONE BYTE PER WEIGHT (any value 0..255);
 *   c: m*n f32 row-major, fully overwritten with sum_l a[i,l]*b[l,j].
 * Host pointers; synchronous (results in c on return). Accumulates in int32 and converts
 * once, which equals the reference's f32 accumulation whenever that is exact (|sum| < 2^24). */
int bitnet_cpp_matmul_i2s(const int8_t* a, const uint8_t* b, float* c, int m, int n, int k);
/* replaces ffi/cpp_bridge.cpp:127-278 (bridge.rs:29-37); semantics of
 * KernelProvider::quantize (lib.rs:51-57, cpu/fallback.rs:85-290): the WEIGHT quantisers.
 * qtype 0 = I2_S (32-blocks), 1 = TL1 (64), 2 = TL2 (128); scale = absmax/1.5 (1.0 if
 * absmax <= 1e-8); 2-bit codes are OR-ed into `out` LSB-first (caller zeroes it).
 * Needs out_len >= in_len/4 and scales_len >= ceil(in_len/block). Host pointers. */
int bitnet_cpp_quantize(const float* in, int in_len, uint8_t* out, int out_len, float* scales,
                        int scales_len, int qtype);
/* replaces ffi/cpp_bridge.cpp:280-282 (bridge.rs:38): NULL when no error is pending */
const char* bitnet_cpp_get_last_error(void);

/* ------------------------------------------------------------------------------------ *
 * Part 2 -- gemv_qk256 (the symbol the live decode/prefill loop calls)                   *
 * ------------------------------------------------------------------------------------ */

/* replaces bitnet_quantization::i2s_qk256::gemv_qk256 (i2s_qk256.rs:346-372; called from
 * crates/bitnet-transformer/src/lib.rs:683-700 and :923-940). Host slices:
 *   qs: rows*row_stride_bytes packed 2-bit codes, row_stride_bytes = ceil(cols/256)*64;
 *   x: >= cols f32; y: exactly rows f32. y[r] = sum_c LUT_A[code(r,c)] * x[c], no scale,
 *   LUT_A = {-2,-1,+1,+2}. Errors carry the reference's substrings "y_out length",
 *   "x length", "too short" (i2s_qk256.rs:301-312). */
int bitnet_b200_gemv_qk256(const uint8_t* qs, size_t qs_len, const float* x, size_t x_len,
                           float* y, size_t y_len, size_t rows, size_t cols,
                           size_t row_stride_bytes);

/* Legacy QK256 entry points kept by the reference for old call sites and benches
 * (crates/bitnet-quantization/src/qk256_dispatch.rs). Host slices; `packed` is rows*cols/4 bytes,
 * `scales` rows*(cols/256) f32. The reference asserts on a violated contract; here the same messages
 * ("Output length mismatch", "Activation length mismatch", "Cols must be multiple of QK256",
 * "Packed weight size mismatch", "Scales length mismatch") come back as errors.
 *   qk256_gemv        (:20-43): validates, then forwards to gemv_qk256 -- map A, `scales` NOT consumed;
 *   qk256_gemv_scalar (:45-99): map D {-1, 0, +1, -1}, out[r] = sum_blk scales[r,blk] * blocksum. */
int bitnet_b200_qk256_gemv(float* output, size_t out_len, size_t rows, size_t cols, const uint8_t* packed,
                           size_t packed_len, const float* scales, size_t scales_len, const float* activations,
                           size_t act_len);
int bitnet_b200_qk256_gemv_scalar(float* output, size_t out_len, size_t rows, size_t cols, const uint8_t* packed,
                                  size_t packed_len, const float* scales, size_t scales_len,
                                  const float* activations, size_t act_len);

/* ------------------------------------------------------------------------------------ *
 * Part 3 -- additive device-resident API (the fast path)                                 *
 * ------------------------------------------------------------------------------------ */

typedef struct bitnet_b200_ctx bitnet_b200_ctx;       /* device ordinal + stream */
typedef struct bitnet_b200_weight bitnet_b200_weight; /* one packed weight tensor in HBM */

enum { BITNET_B200_LAYOUT_QK256 = 0,        /* I2SFlavor::GgmlQk256NoScale  (types.rs:797-806) */
       BITNET_B200_LAYOUT_BITNET32_F16 = 1  /* I2SFlavor::BitNet32F16, 10 B per 32 weights      */ };
enum { BITNET_B200_SCALE_ABS_CLAMP = 0,     /* s = clamp(|f16|, 1e-3, 1e3)  (quant/i2s.rs:76-94) */
       BITNET_B200_SCALE_RAW = 1            /* s = f16 as stored                                 */ };

/* device selection: counterpart of CudaKernel::new_with_device (gpu/cuda.rs:62).
 * `stream` is the cudaStream_t (as void*) the context launches on; NULL is the legacy
 * default stream. The context never owns the stream. */
int bitnet_b200_ctx_create(int device, void* stream, bitnet_b200_ctx** out);
int bitnet_b200_ctx_set_stream(bitnet_b200_ctx* ctx, void* stream);
int bitnet_b200_ctx_synchronize(bitnet_b200_ctx* ctx);
void bitnet_b200_ctx_destroy(bitnet_b200_ctx* ctx);

/* Upload one weight tensor, once, from host bytes in the reference's on-disk layout:
 *   QK256: rows * ceil(cols/256)*64 bytes (I2SQk256NoScale, i2s_qk256.rs:65-107; +-128 B
 *          trailing padding tolerated as there);
 *   BITNET32_F16: rows * (cols/32) * 10 bytes, blocks along cols (quant/i2s.rs:621-692);
 *          cols % 32 == 0 required.
 * `lut` = code -> value map, 4 int8 (NULL = map A {-2,-1,1,2}). BitNet32-F16 is split
 * on the device, losslessly, into a 16-byte-aligned code plane + f16 scale plane. */
int bitnet_b200_weight_create(bitnet_b200_ctx* ctx, const void* host_bytes, size_t nbytes,
                              int layout, size_t rows, size_t cols, const int8_t* lut,
                              int scale_policy, bitnet_b200_weight** out);
void bitnet_b200_weight_destroy(bitnet_b200_weight* w);
size_t bitnet_b200_weight_rows(const bitnet_b200_weight* w);
size_t bitnet_b200_weight_cols(const bitnet_b200_weight* w);
/* packed bytes in the reference layout (the algorithmic weight bytes of one linear call) */
size_t bitnet_b200_weight_nbytes(const bitnet_b200_weight* w);

/* per-token absmax int8 activation quantiser, quantize_symmetric_i8(.., 8) per row
 * (crates/bitnet-kernels/src/cpu/quantize.rs:32-49), bit-exact: scale = absmax/127,
 * q = clamp(round_half_away(v * (1/scale)), -127, 127); all-zero row -> scale 0.
 * x: [M,K] f32, q: [M,K] int8, scale: [M] f32 -- all DEVICE pointers. */
int bitnet_b200_actquant_i8(bitnet_b200_ctx* ctx, const float* x, int M, int K, int8_t* q,
                            float* scale);
/* RMSNorm (cpu::rms_norm, crates/bitnet-kernels/src/cpu/layer_norm.rs:118-137: f64 mean of
 * squares, x * (1/sqrt(ms+eps)) * gamma) with the act-quant fused into its epilogue.
 * `normed` (DEVICE, [M,K] f32) may be NULL. */
int bitnet_b200_rmsnorm_actquant(bitnet_b200_ctx* ctx, const float* x, const float* gamma,
                                 float eps, int M, int K, int8_t* q, float* scale, float* normed);

/* The ternary linear: y[m,n] = act_scale[m] * sum_blk s[n,blk] * acc_i32[m,n,blk], with
 * acc = integer dot of LUT codes and int8 activations (QK256: one block, s = 1).
 * q: [M,K] int8, act_scale: [M] f32, y: [M,N] f32 -- DEVICE pointers. M <= 16 runs the
 * bandwidth-bound GEMV, M >= 64 the tcgen05 kind::i8 GEMM; between, whichever is faster.
 * acc_out (DEVICE, may be NULL): QK256 -> [M,N] int32; BitNet32-F16 -> [M,N,K/32] int32.
 * It exists so the bit-exact check can read the accumulators directly. */
int bitnet_b200_linear_i8(bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q,
                          const float* act_scale, int M, float* y, int32_t* acc_out);
/* linear_i8 with the SwiGLU glue in the GEMM's epilogue (batched decode / prefill, M >= 9, QK256): the weight holds
 * interleaved (gate, up) rows [2F, K]; h[m, i] = silu(y[m, 2i]) * y[m, 2i+1] with y = the linear_i8 result, bit for
 * bit, and y itself is never written (half the output bytes, and the following act-quant reads h once).
 * crates/bitnet-transformer/src/lib.rs:756-789. h: [M, F] f32, DEVICE. */
int bitnet_b200_linear_i8_swiglu(bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q,
                                 const float* act_scale, int M, float* h);
/* linear_i8 for callers that must not allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc). No counterpart in the reference.
 * Same result as bitnet_b200_linear_i8 bit for bit (the tcgen05 GEMM, QK256 weights, cols % 128 == 0, M >= 9, no
 * acc_out), but the split-K scratch lives in a file-scope device array of the library instead of cudaMalloc'ed
 * buffers of the context: 32 KiB of ticket counters + 4 MiB of int32 accumulation blocks, held by every GPU that loads
 * the library until the process exits. Limit: ceil(rows / 128) * (tokens, padded per token tile to 16) * 512 bytes
 * <= 4 MiB, e.g. 13824 x 64, 4096 x 256, 2560 x 320; larger problems are rejected with an error naming the bytes
 * needed. The call itself allocates nothing and clears the part of the array it may use before the launch (on the
 * context's stream). Every call on a GPU shares the array: calls that overlap on that GPU (two streams, two host
 * threads) race on it -- serialise them or use bitnet_b200_linear_i8. */
int bitnet_b200_linear_i8_static_scratch(bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q,
                                         const float* act_scale, int M, float* y);
/* The decode fast path as ONE launch (gemv_stream.cu): [RMSNorm ->] per-token absmax int8
 * quantise -> ternary linear -> epilogue. It is the composition of rmsnorm_actquant /
 * actquant_i8 and linear_i8 above, bit for bit, with the activation work done in the GEMV's
 * prologue ("act-quant fused with RMSNorm", here fused into the consumer so that no separate
 * launch sits on the token's critical path). DEVICE pointers.
 *   x [M,K] f32; gamma [K] f32 or NULL (no norm); eps as cpu::rms_norm;
 *   epilogue BITNET_B200_EPI_STORE    y[M,N]  = act_scale * acc (* block scales)
 *            BITNET_B200_EPI_RESIDUAL y[M,N]  = residual[M,N] + that (residual may alias y)
 *            BITNET_B200_EPI_SWIGLU   y[M,N/2]= silu(row 2i) * row 2i+1  (gate/up rows interleaved;
 *                                      crates/bitnet-transformer/src/lib.rs:756-789)
 *   q_out [M,K] int8 / scale_out [M] f32: optional copies of what the prologue quantised.
 * M <= 4 and K % 64 == 0 (BitNet32-F16: K % 256 == 0); other shapes are an error here --
 * call the two-step API. */
enum { BITNET_B200_EPI_STORE = 0, BITNET_B200_EPI_RESIDUAL = 1, BITNET_B200_EPI_SWIGLU = 2 };
int bitnet_b200_fused_linear(bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const float* x,
                             const float* gamma, float eps, int M, int epilogue, float* y,
                             const float* residual, int8_t* q_out, float* scale_out);
/* f32-activation GEMV on a device-resident QK256 weight (device twin of Part 2):
 * x: [M,K] f32, y: [M,N] f32, DEVICE pointers. */
int bitnet_b200_linear_f32(bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const float* x,
                           int M, float* y);

/* Host-slice fills of the reference's stubbed CUDA launchers:
 *   launch_qk256_gemv (crates/bitnet-kernels/src/cuda/qk256_gemv.rs:119-137)
 *   launch_i2s_matmul (crates/bitnet-kernels/src/cuda/quantized_matmul.rs:282-303):
 * out[m,n] = sum_blk scales[n*nblk+blk] * sum_{k in blk} act[m,k] * LUT_C[code(n,k)],
 * weights packed [n, ceil(k/4)] (map C {0,+1,0,-1}), f32 activations, block_size 32 or 256. */
int bitnet_b200_i2s_matmul_f32(const float* act, const uint8_t* packed, const float* scales,
                               float* out, int m, int n, int k, int block_size);

/* ------------------------------------------------------------------------------------ *
 * Part 4 -- decode harness ("next" rows of the scope table): a 2B-4T-shaped decoder     *
 * step on device, the ternary linears above plus the glue the reference runs on the     *
 * host through candle (crates/bitnet-transformer/src/lib.rs:275-543,751-798,1557-1630). *
 * ------------------------------------------------------------------------------------ */
typedef struct bitnet_b200_model bitnet_b200_model;
typedef struct {
    int hidden, ffn, n_layers, n_heads, n_kv_heads, head_dim, vocab, max_ctx, batch;
    float rms_eps;
    int layout; /* BITNET_B200_LAYOUT_* of every projection */
} bitnet_b200_model_config; /* cf. BitNetConfig, crates/bitnet-common/src/config.rs:25-46 */

int bitnet_b200_model_create(bitnet_b200_ctx* ctx, const bitnet_b200_model_config* cfg,
                             bitnet_b200_model** out);
void bitnet_b200_model_destroy(bitnet_b200_model* m);
/* host bytes in the reference layout, [out, in] rows: wq [NH*D,H], wk/wv [NKV*D,H],
 * wo [H,NH*D], wgate/wup [F,H], wdown [H,F]; norms are [H] f32 RMSNorm gains. */
int bitnet_b200_model_set_layer(bitnet_b200_model* m, int layer, const void* wq, const void* wk,
                                const void* wv, const void* wo, const void* wgate, const void* wup,
                                const void* wdown, const float* attn_norm, const float* ffn_norm);
/* tied embedding / LM head as IEEE f16 bits [vocab, hidden]; final RMSNorm gain [hidden] */
int bitnet_b200_model_set_embedding(bitnet_b200_model* m, const uint16_t* emb_f16,
                                    const float* final_norm);
/* RoPE tables [max_ctx, head_dim/2] (split-half pairing, transformer lib.rs:134-184) */
int bitnet_b200_model_set_rope(bitnet_b200_model* m, const float* cos_t, const float* sin_t);
int bitnet_b200_model_reset(bitnet_b200_model* m); /* position <- 0 */
/* one greedy step through HOST buffers: tokens[batch] in, next_tokens[batch] out (argmax,
 * lowest index on ties), logits_out [batch, vocab] optional; synchronous */
int bitnet_b200_model_decode(bitnet_b200_model* m, const int32_t* tokens, int32_t* next_tokens,
                             float* logits_out);
/* n greedy steps with the token fed back on device; asynchronous on the context's stream */
int bitnet_b200_model_decode_device(bitnet_b200_model* m, int n_steps);
int bitnet_b200_model_set_tokens(bitnet_b200_model* m, const int32_t* tokens);
int bitnet_b200_model_get_tokens(bitnet_b200_model* m, int32_t* next_tokens);
/* times only the ternary linears of one step (graph replay between events on the stream) */
int bitnet_b200_model_time_linears(bitnet_b200_model* m, int reps, float* ms_per_pass,
                                   int* launches_per_pass);
uint64_t bitnet_b200_model_launches_per_step(const bitnet_b200_model* m);
uint64_t bitnet_b200_model_weight_bytes(const bitnet_b200_model* m); /* packed ternary bytes */

/* ------------------------------------------------------------------------------------ *
 * Part 5 -- tensor-parallel building blocks. The reference has only  *
 * the in-process simulation of the semantics (partition_tensor / all_reduce Sum and Max, *
 * crates/bitnet-inference/src/tensor_parallel.rs:123-290); the exchange itself (NCCL      *
 * all-reduce over NVLink) is issued by the host between these calls, on the same stream.  *
 * q/k/v and gate/up are sharded by output rows (no exchange); o and down by input         *
 * columns: each rank quantises its K/tp slice with the GLOBAL per-token absmax            *
 * (row_absmax -> all-reduce MAX -> actquant_i8_absmax), so the int32 partials of          *
 * linear_i8(acc_out) add up (all-reduce SUM, int32) to the tp = 1 accumulator bit for bit *
 * and residual_acc applies the same one-multiply epilogue as tp = 1. DEVICE pointers.     *
 * ------------------------------------------------------------------------------------ */
/* absmax[m] = max_k |v[m,k]|; swiglu = 0: v = x [M,K]; swiglu = 1: x is [M,2K] with (gate, up)
 * interleaved, v = silu(gate) * up (crates/bitnet-transformer/src/lib.rs:756-789), also stored to h [M,K] */
int bitnet_b200_row_absmax(bitnet_b200_ctx* ctx, const float* x, int M, int K, int swiglu, float* h,
                           float* absmax);
/* quantize_symmetric_i8 (cpu/quantize.rs:32-49) per row with a GIVEN absmax[m] instead of the row's own */
int bitnet_b200_actquant_i8_absmax(bitnet_b200_ctx* ctx, const float* x, const float* absmax, int M, int K,
                                   int8_t* q, float* scale);
/* x[m,n] += act_scale[m] * float(acc[m,n]) -- the residual epilogue of linear_i8 on reduced accumulators */
int bitnet_b200_residual_acc(bitnet_b200_ctx* ctx, const int32_t* acc, const float* act_scale, float* x, int M,
                             int N);
/* decode glue on the rank's local heads: RoPE on q (in place) and k, KV append at *pos_dev, GQA attention
 * over 0..*pos_dev (crates/bitnet-transformer/src/lib.rs:374-540). qkv [B,(NH+2*NKV)*D], caches
 * [B,NKV,max_ctx,D], att [B,NH*D]; cos/sin [max_ctx, D/2] */
int bitnet_b200_rope_attn_decode(bitnet_b200_ctx* ctx, float* qkv, const float* cos_t, const float* sin_t,
                                 float* kc, float* vc, const int32_t* pos_dev, int B, int NH, int NKV, int D,
                                 int max_ctx, float* att);

/* ---- fused exchanges: the consumer kernel pulls its peers' buffers over NVLink itself (no NCCL call) --------
 * One descriptor per exchange point. All pointers are DEVICE pointers valid on THIS GPU: peer_data[r] / peer_flags[r]
 * are rank r's buffers mapped into this process (symmetric memory / CUDA IPC / the same GPU in the emulation);
 * index `rank` is the local one. peer_flags[r]: uint32 [4 slots][8 ranks], zero-initialised; counters: local uint32
 * [8], zero-initialised. slot in 0..3 names the exchange point (o absmax, o sum, down absmax, down sum).
 * phase: 1 = only tell the peers "my data of this exchange is ready", 2 = wait for every rank and do the work,
 * 3 = both (what a real rank passes; 1 then 2 lets several emulated ranks share one GPU without waiting on a
 * kernel that is not running yet). Protocol and reuse rules: bitnet-rs_b200/csrc/tp_xchg.cuh. */
typedef struct {
    const void* peer_data[8];
    uint32_t* peer_flags[8];
    uint32_t* counters;
    int n_ranks, rank, slot, phase;
} bitnet_b200_tp_xchg;
/* all-reduce(SUM, int32) + residual + RMSNorm + act-quant in ONE kernel, replacing
 *   ncclAllReduce(acc) ; residual_acc(acc, act_scale, x) ; rmsnorm_actquant(x, gamma, eps, q, scale, normed):
 * acc_total[m,n] = sum_r peer_data[r][m,n] (int32 [M,N] partial accumulators of a row-parallel linear_i8);
 * x[m,n] += act_scale[m] * float(acc_total) ; then cpu::rms_norm + quantize_symmetric_i8 of the row, bit for bit
 * what the three calls produce. Every rank computes all M rows (x stays replicated). N % 4 == 0, N <= 16384. */
int bitnet_b200_tp_xchg_sum_rmsnorm_actquant(bitnet_b200_ctx* ctx, const bitnet_b200_tp_xchg* xc,
                                             const float* act_scale, float* x, const float* gamma, float eps, int M,
                                             int N, int8_t* q, float* scale, float* normed);
/* all-reduce(MAX, f32) of the per-token absmax + quantise the local slice in ONE kernel, replacing
 *   ncclAllReduce(absmax, MAX) ; actquant_i8_absmax(x, absmax, q, scale):  peer_data[r] = rank r's absmax [M] f32. */
int bitnet_b200_tp_xchg_max_actquant(bitnet_b200_ctx* ctx, const bitnet_b200_tp_xchg* xc, const float* x, int M,
                                     int K, int8_t* q, float* scale);

/* library / build identification */
const char* bitnet_b200_version(void);
/* number of kernels this library has launched on the calling process (bench bookkeeping) */
uint64_t bitnet_b200_launch_count(void);
/* which kernel the calling thread's last bitnet_b200_linear_i8 ran: "gemm_small", "gemm_small_resident",
 * "gemm_tcgen05", "gemm_f16_tcgen05", "gemv_fast" or "gemv" (diagnostics: bench labels and routing tests) */
const char* bitnet_b200_last_linear_kernel(void);

#ifdef __cplusplus
}
#endif
#endif /* BITNET_B200_H */
