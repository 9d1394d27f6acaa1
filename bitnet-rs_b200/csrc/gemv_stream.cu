// gemv_stream.cu -- the decode-regime ternary GEMV (M = 1..4) as a weight-STREAMING kernel.
//
// Replaces, for device-resident packed weights, gemv_qk256 / gemv_qk256_row_avx2
// (crates/bitnet-quantization/src/i2s_qk256.rs:346-372, i2s_qk256_avx2.rs:81-229), the
// blocked/scaled template i2s_matmul_blocked (crates/bitnet-kernels/src/cpu/quantized_matmul.rs:
// 155-200) and the per-token quantiser quantize_symmetric_i8 (cpu/quantize.rs:32-49) that
// feeds them; fills the stubs launch_qk256_gemv / launch_rmsnorm
// (crates/bitnet-kernels/src/cuda/qk256_gemv.rs:119-137, cuda/rmsnorm.rs:29-80).
//
// Why it looks like this (B200-first): one 2B-4T projection is 1.6-8.8 MB, i.e. 0.25-1.4 us
// of HBM time -- less than a kernel boundary. So the kernel is built to keep HBM busy ACROSS
// kernel boundaries:
//   * a CTA owns a contiguous slice of weight rows (= one contiguous byte range, both
//     layouts keep rows contiguous) and fetches it with a handful of 1-D bulk copies (TMA,
//     cp.async.bulk -> mbarrier) issued by one thread BEFORE griddepcontrol.wait: with
//     programmatic dependent launch the next projection's weights stream into shared memory
//     while the previous kernel is still computing;
//   * the activation work that used to be separate launches is the prologue, recomputed by
//     every CTA from the L2-resident f32 row (10-27 KB): RMSNorm (f64 mean of squares,
//     cpu/layer_norm.rs:118-137) and/or per-token absmax int8 quantisation, bit-exact;
//   * thread t owns 16-byte chunk (t mod rc) of every row it visits (rc chunks per row), so
//     its 64 int8 activations stay in registers and shared-memory reads are conflict-free
//     128-bit; 2-bit codes go through a 4-entry int8 LUT with one PRMT per four weights
//     and into dp4a (int32 accumulators, exact);
//   * epilogues: scale, residual add, or SwiGLU glue (silu(gate)*up on interleaved rows,
//     crates/bitnet-transformer/src/lib.rs:756-789).
#include <type_traits>

#include "common.cuh"

namespace b200 {

enum { PRO_Q8 = 0, PRO_F32 = 1, PRO_RMS = 2 };
enum { EPI_STORE = 0, EPI_RESIDUAL = 1, EPI_SWIGLU = 2 };

struct StreamParams {
    const uint8_t* codes;    // [N, rc*16] device layout
    const __half* scales;    // [N, nblk32] raw f16 (BitNet32-F16) or nullptr
    const int8_t* q;         // PRO_Q8: [M, K]
    const float* act_scale;  // PRO_Q8: [M]
    const float* x;          // PRO_F32 / PRO_RMS: [M, K]
    const float* gamma;      // PRO_RMS: [K]
    float* y;                // [M, N]  (EPI_SWIGLU: [M, N/2])
    const float* residual;   // EPI_RESIDUAL: [M, N]
    int32_t* acc_out;        // optional
    int8_t* q_out;           // optional, written by CTA 0: the prologue's int8 row(s)
    float* scale_out;        // optional, CTA 0
    float* normed_out;       // optional, CTA 0 (PRO_RMS)
    float eps;
    uint32_t lut;
    int N, K, M;
    int rc, P, rows_per_cta, nblk32, clamp_scale, pro, epi;
};

__device__ __forceinline__ float apply_scale_policy(float sv, int clamp) {
    if (clamp) {  // quant/i2s.rs:76-94: |s| clamped to [1e-3, 1e3]; comparisons keep NaN
        sv = fabsf(sv);
        sv = sv < 1e-3f ? 1e-3f : sv;
        sv = sv > 1e3f ? 1e3f : sv;
    }
    return sv;
}

// q = clamp(round_half_away(v * inv), -127, 127); NaN -> 0 (Rust `as i8`)
__device__ __forceinline__ int quant_i8(float v, float inv) {
    float r = roundf(v * inv);
    if (r != r) return 0;
    r = fminf(fmaxf(r, -127.f), 127.f);
    return (int)r;
}

constexpr int NVMAX = 8;  // float4 per thread held by the fused prologue

template <int MT, bool SCALED, bool WACC>
__global__ void __launch_bounds__(MT >= 4 ? 512 : 1024) gemv_stream_kernel(const StreamParams p) {
    using acc_t = typename std::conditional<SCALED, float, int>::type;
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int nact = p.rc * p.P;  // threads that own a chunk
    const int stride = p.rc * 16;
    const int row_begin = blockIdx.x * p.rows_per_cta;
    const int nrows = min(p.rows_per_cta, p.N - row_begin);
    const int npass = (nrows + p.P - 1) / p.P;
    const int nch = npass < 4 ? npass : 4;
    const int ppc = (npass + nch - 1) / nch;

    // ---- shared memory carve-up ------------------------------------------------------------
    uint64_t* mbar = reinterpret_cast<uint64_t*>(smem);                       // [4]
    double* shd = reinterpret_cast<double*>(smem + 64);                        // [32][MT]
    float* shf = reinterpret_cast<float*>(smem + 64 + 32 * MT * 8);            // [32][MT]
    float* sc_s = shf + 32 * MT;                                               // [MT]
    size_t off = 64 + 32 * MT * 8 + 32 * MT * 4 + 64;
    off = (off + 127) & ~(size_t)127;
    unsigned char* wbuf = smem + off;
    off += (size_t)p.rows_per_cta * stride;
    unsigned char* sbuf = smem + off;
    if (SCALED) off += ((size_t)p.rows_per_cta * p.nblk32 * 2 + 15) & ~(size_t)15;
    int8_t* qbuf = reinterpret_cast<int8_t*>(smem + off);
    off += (size_t)MT * p.K;
    acc_t* red = reinterpret_cast<acc_t*>(smem + off);  // [MT][rows_per_cta][rc + 1]
    const int red_ld = p.rc + 1;

    // ---- weight prefetch: independent of the previous kernel -> before griddepcontrol.wait --
    if (tid == 0) {
        for (int c = 0; c < nch; ++c) mbar_init(&mbar[c], 1);
        mbar_fence_init();
        for (int c = 0; c < nch; ++c) {
            const int r0 = c * ppc * p.P;
            const int r1 = min(nrows, (c + 1) * ppc * p.P);
            if (r1 <= r0) {  // cannot happen for c < nch, kept for safety
                mbar_arrive(&mbar[c]);
                continue;
            }
            const uint32_t wb = (uint32_t)(r1 - r0) * stride;
            const uint32_t sb = SCALED ? (uint32_t)(r1 - r0) * p.nblk32 * 2 : 0u;
            mbar_expect_tx(&mbar[c], wb + sb);
            bulk_g2s(wbuf + (size_t)r0 * stride, p.codes + (size_t)(row_begin + r0) * stride, wb, &mbar[c]);
            if (SCALED)
                bulk_g2s(sbuf + (size_t)r0 * p.nblk32 * 2,
                         reinterpret_cast<const unsigned char*>(p.scales) + (size_t)(row_begin + r0) * p.nblk32 * 2,
                         sb, &mbar[c]);
        }
    }
    pdl_launch_dependents();
    pdl_wait();  // everything below may read what the previous kernel wrote
    __syncthreads();  // mbarrier inits visible to all waiters

    // ---- prologue: int8 activations into qbuf, per-token scale into sc_s -------------------
    if (p.pro == PRO_Q8) {
        const int nvec = MT * p.K / 16;
        for (int i = tid; i < nvec; i += blockDim.x) {
            const int m = (i * 16) / p.K;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (m < p.M) v = *reinterpret_cast<const uint4*>(p.q + (size_t)i * 16);
            reinterpret_cast<uint4*>(qbuf)[i] = v;
        }
        if (tid < MT) sc_s[tid] = tid < p.M ? p.act_scale[tid] : 0.f;
    } else {
        const int ngrp = p.K / 4;  // float4 groups per row
#pragma unroll
        for (int m = 0; m < MT; ++m) {  // one token at a time: 32 live registers, 2 barriers each
            float4 v[NVMAX];
#pragma unroll
            for (int j = 0; j < NVMAX; ++j) {
                const int g = tid + j * blockDim.x;
                v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (g < ngrp && m < p.M) v[j] = *reinterpret_cast<const float4*>(p.x + (size_t)m * p.K + 4 * g);
            }
            if (p.pro == PRO_RMS) {
                double ss = 0.0;
#pragma unroll
                for (int j = 0; j < NVMAX; ++j)
                    ss += (double)v[j].x * (double)v[j].x + (double)v[j].y * (double)v[j].y +
                          (double)v[j].z * (double)v[j].z + (double)v[j].w * (double)v[j].w;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
                if (lane == 0) shd[warp * MT + m] = ss;
                __syncthreads();
                double tot = 0.0;
                for (int wv = 0; wv < nwarps; ++wv) tot += shd[wv * MT + m];
                const float ms = (float)(tot / (double)p.K);
                const float inv_rms = 1.0f / sqrtf(ms + p.eps);
#pragma unroll
                for (int j = 0; j < NVMAX; ++j) {
                    const int g = tid + j * blockDim.x;
                    if (g < ngrp) {
                        const float4 gm = *reinterpret_cast<const float4*>(p.gamma + 4 * g);
                        v[j].x = v[j].x * inv_rms * gm.x;
                        v[j].y = v[j].y * inv_rms * gm.y;
                        v[j].z = v[j].z * inv_rms * gm.z;
                        v[j].w = v[j].w * inv_rms * gm.w;
                    }
                }
            }
            float amax = 0.f;
#pragma unroll
            for (int j = 0; j < NVMAX; ++j)
                amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v[j].x), fabsf(v[j].y)), fmaxf(fabsf(v[j].z), fabsf(v[j].w))));
            amax = warp_max(amax);
            if (lane == 0) shf[warp * MT + m] = amax;
            __syncthreads();
            amax = 0.f;
            for (int wv = 0; wv < nwarps; ++wv) amax = fmaxf(amax, shf[wv * MT + m]);
            const float sc = amax == 0.f ? 0.f : amax / 127.f;
            const float inv = amax == 0.f ? 0.f : 1.0f / sc;
            if (tid == 0) {
                sc_s[m] = sc;
                if (blockIdx.x == 0 && p.scale_out && m < p.M) p.scale_out[m] = sc;
            }
#pragma unroll
            for (int j = 0; j < NVMAX; ++j) {
                const int g = tid + j * blockDim.x;
                if (g < ngrp) {
                    uint32_t wq = 0;
                    if (amax != 0.f)
                        wq = (uint32_t)(uint8_t)quant_i8(v[j].x, inv) | ((uint32_t)(uint8_t)quant_i8(v[j].y, inv) << 8) |
                             ((uint32_t)(uint8_t)quant_i8(v[j].z, inv) << 16) | ((uint32_t)(uint8_t)quant_i8(v[j].w, inv) << 24);
                    reinterpret_cast<uint32_t*>(qbuf + (size_t)m * p.K)[g] = wq;
                    if (blockIdx.x == 0 && m < p.M) {
                        if (p.q_out) reinterpret_cast<uint32_t*>(p.q_out + (size_t)m * p.K)[g] = wq;
                        if (p.normed_out) *reinterpret_cast<float4*>(p.normed_out + (size_t)m * p.K + 4 * g) = v[j];
                    }
                }
            }
        }
    }
    __syncthreads();

    // ---- this thread's 64 activations per token: registers for the whole kernel -------------
    const bool active = tid < nact;
    const int c = active ? tid % p.rc : 0;
    const int slot = active ? tid / p.rc : 0;
    uint4 act[MT][4];
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int w4 = 0; w4 < 4; ++w4)
            act[m][w4] = active ? *reinterpret_cast<const uint4*>(qbuf + (size_t)m * p.K + c * 64 + w4 * 16)
                                : make_uint4(0, 0, 0, 0);

    // ---- main loop: one 16-byte chunk (64 weights) per thread per pass ----------------------
    for (int i = 0; i < npass; ++i) {
        if (i % ppc == 0) mbar_wait(&mbar[i / ppc], 0);
        const int rl = i * p.P + slot;
        if (active && rl < nrows) {
            const uint4 wv = *reinterpret_cast<const uint4*>(wbuf + ((size_t)i * nact + tid) * 16);
            float s0 = 0.f, s1 = 0.f;
            if (SCALED) {
                const __half2 hs = *reinterpret_cast<const __half2*>(sbuf + ((size_t)i * nact + tid) * 4);
                s0 = apply_scale_policy(__low2float(hs), p.clamp_scale);
                s1 = apply_scale_policy(__high2float(hs), p.clamp_scale);
            }
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                int d01 = dot16(wv.x, p.lut, act[m][0], 0);
                d01 = dot16(wv.y, p.lut, act[m][1], d01);
                int d23 = dot16(wv.z, p.lut, act[m][2], 0);
                d23 = dot16(wv.w, p.lut, act[m][3], d23);
                acc_t r;
                if (SCALED) {
                    r = (acc_t)(s0 * (float)d01 + s1 * (float)d23);
                    if (WACC && m < p.M) {
                        int32_t* ao = p.acc_out + ((size_t)m * p.N + row_begin + rl) * p.nblk32 + 2 * c;
                        ao[0] = d01;
                        ao[1] = d23;
                    }
                } else {
                    r = (acc_t)(d01 + d23);
                }
                red[((size_t)m * p.rows_per_cta + rl) * red_ld + c] = r;
            }
        }
    }
    __syncthreads();

    // ---- per-row reduction over the rc chunk partials + epilogue (one warp per row pair) ----
    const int npairs = (nrows + 1) / 2;
    for (int pr = warp; pr < npairs; pr += nwarps) {
#pragma unroll
        for (int m = 0; m < MT; ++m) {
            acc_t t0 = 0, t1 = 0;
            const acc_t* r0p = red + ((size_t)m * p.rows_per_cta + 2 * pr) * red_ld;
            const bool has1 = 2 * pr + 1 < nrows;
            for (int cc = lane; cc < p.rc; cc += 32) {
                t0 += r0p[cc];
                if (has1) t1 += r0p[red_ld + cc];
            }
            t0 = warp_sum(t0);
            t1 = warp_sum(t1);
            if (lane == 0 && m < p.M) {
                const float as = sc_s[m];
                const float v0 = __fmul_rn(as, (float)t0), v1 = __fmul_rn(as, (float)t1);  // no FMA contraction
                const int row = row_begin + 2 * pr;
                if (p.epi == EPI_SWIGLU) {
                    // rows (2i, 2i+1) = (gate_i, up_i): h_i = silu(gate) * up
                    p.y[(size_t)m * (p.N / 2) + (row >> 1)] = silu_mul(v0, v1);
                } else {
                    const size_t oi = (size_t)m * p.N + row;
                    if (p.epi == EPI_RESIDUAL) {
                        p.y[oi] = __fadd_rn(p.residual[oi], v0);
                        if (has1) p.y[oi + 1] = __fadd_rn(p.residual[oi + 1], v1);
                    } else {
                        p.y[oi] = v0;
                        if (has1) p.y[oi + 1] = v1;
                    }
                    if (WACC && !SCALED) {
                        p.acc_out[oi] = (int32_t)t0;
                        if (has1) p.acc_out[oi + 1] = (int32_t)t1;
                    }
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
struct StreamPlan {
    int rc, P, threads, rows_per_cta, grid;
    size_t smem;
};

static size_t stream_smem(int MT, bool scaled, int rows, int rc, int nblk32, int K) {
    size_t off = 64 + 32 * (size_t)MT * 8 + 32 * (size_t)MT * 4 + 64;
    off = (off + 127) & ~(size_t)127;
    off += (size_t)rows * rc * 16;
    if (scaled) off += ((size_t)rows * nblk32 * 2 + 15) & ~(size_t)15;
    off += (size_t)MT * K;
    off += (size_t)MT * rows * (rc + 1) * 4;
    return off;
}

int gemv_stream_rows_override = 0;  // test/bench knob (0 = heuristic)

// returns false when the shape must take the generic kernels (gemv.cu)
static bool plan_stream(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M, int MT, bool need_f32_prologue,
                        StreamPlan* pl) {
    if (M < 1 || M > 4) return false;
    if (w->code_stride % 16 != 0 || w->cols % 64 != 0) return false;
    const bool scaled = w->layout == BITNET_B200_LAYOUT_BITNET32_F16;
    if (scaled && (w->cols % 256 != 0)) return false;  // scale rows must be 16-byte multiples for the bulk copy
    const int rc = (int)(w->code_stride / 16);
    const int max_threads = MT >= 4 ? 512 : 1024;
    if (rc > max_threads) return false;
    int P = (336 + rc / 2) / rc;
    if (P < 1) P = 1;
    while (rc * P > max_threads) --P;
    const int threads = (rc * P + 31) / 32 * 32;
    if (need_f32_prologue && (w->cols / 4 > (size_t)NVMAX * threads)) return false;
    // rows per CTA: one slice per SM when it fits in ~64 KB, else more CTAs per SM
    int R = (int)div_ceil(w->rows, (size_t)ctx->num_sms);
    const size_t row_bytes = w->code_stride + (scaled ? w->nblk32 * 2 : 0);
    const int Rmax = (int)(65536 / row_bytes);
    if (R > Rmax) {
        const int k = (int)div_ceil((size_t)R, (size_t)Rmax);
        R = (int)div_ceil(w->rows, (size_t)ctx->num_sms * k);
    }
    if (gemv_stream_rows_override > 0) R = gemv_stream_rows_override;
    R = (R + 1) & ~1;
    if (R < 2) R = 2;
    pl->rc = rc;
    pl->P = P;
    pl->threads = threads;
    pl->rows_per_cta = R;
    pl->grid = (int)div_ceil(w->rows, (size_t)R);
    pl->smem = stream_smem(MT, scaled, R, rc, (int)w->nblk32, (int)w->cols);
    return pl->smem <= 200 * 1024;
}

bool gemv_stream_supported(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M, bool f32_prologue) {
    StreamPlan pl;
    const int MT = M == 1 ? 1 : (M == 2 ? 2 : 4);
    return plan_stream(ctx, w, M, MT, f32_prologue, &pl);
}

template <int MT, bool SCALED, bool WACC>
static int launch_stream_inst(const StreamParams& p, const StreamPlan& pl, cudaStream_t st, bool pdl) {
    static bool attr_set[16] = {};  // per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 16 && !attr_set[dev]) {
        B200_CUDA(cudaFuncSetAttribute(gemv_stream_kernel<MT, SCALED, WACC>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set[dev] = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)pl.grid);
    cfg.blockDim = dim3((unsigned)pl.threads);
    cfg.dynamicSmemBytes = pl.smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    B200_CUDA(cudaLaunchKernelEx(&cfg, gemv_stream_kernel<MT, SCALED, WACC>, p));
    count_launch();
    return 0;
}

// Fused streaming GEMV. Returns 1 when the shape is not supported (caller falls back to the
// generic kernels), 0 on success, -1 on error.
//   pro = PRO_Q8: (q, act_scale) are the int8 rows; PRO_F32 / PRO_RMS: x (and gamma, eps) are f32 rows
//   epi = EPI_STORE / EPI_RESIDUAL (residual may alias y) / EPI_SWIGLU (y is [M, N/2])
int launch_gemv_stream(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M, int pro, const int8_t* q,
                       const float* act_scale, const float* x, const float* gamma, float eps, int epi, float* y,
                       const float* residual, int32_t* acc_out, int8_t* q_out, float* scale_out, float* normed_out,
                       bool pdl) {
    const int MT = M == 1 ? 1 : (M == 2 ? 2 : 4);
    StreamPlan pl;
    if (!plan_stream(ctx, w, M, MT, pro != PRO_Q8, &pl)) return 1;
    if (pro == PRO_Q8 && (reinterpret_cast<uintptr_t>(q) & 15)) return 1;
    if (pro != PRO_Q8 && ((reinterpret_cast<uintptr_t>(x) & 15) || (gamma && (reinterpret_cast<uintptr_t>(gamma) & 15)))) return 1;
    if (epi == EPI_SWIGLU && (w->rows & 1)) return 1;
    if (q_out && (reinterpret_cast<uintptr_t>(q_out) & 3)) return 1;
    if (normed_out && (reinterpret_cast<uintptr_t>(normed_out) & 15)) return 1;
    const bool scaled = w->layout == BITNET_B200_LAYOUT_BITNET32_F16;
    StreamParams p{};
    p.codes = w->codes;
    p.scales = w->scales;
    p.q = q;
    p.act_scale = act_scale;
    p.x = x;
    p.gamma = gamma;
    p.y = y;
    p.residual = residual;
    p.acc_out = acc_out;
    p.q_out = q_out;
    p.scale_out = scale_out;
    p.normed_out = normed_out;
    p.eps = eps;
    p.lut = w->lut_word;
    p.N = (int)w->rows;
    p.K = (int)w->cols;
    p.M = M;
    p.rc = pl.rc;
    p.P = pl.P;
    p.rows_per_cta = pl.rows_per_cta;
    p.nblk32 = (int)w->nblk32;
    p.clamp_scale = w->scale_policy == BITNET_B200_SCALE_ABS_CLAMP ? 1 : 0;
    p.pro = pro;
    p.epi = epi;
    const bool wacc = acc_out != nullptr;
    cudaStream_t st = ctx->stream;
#define B200_STREAM_CASE(MTv)                                                                               \
    if (MT == MTv) {                                                                                        \
        if (scaled) return wacc ? launch_stream_inst<MTv, true, true>(p, pl, st, pdl)                       \
                                : launch_stream_inst<MTv, true, false>(p, pl, st, pdl);                     \
        return wacc ? launch_stream_inst<MTv, false, true>(p, pl, st, pdl)                                  \
                    : launch_stream_inst<MTv, false, false>(p, pl, st, pdl);                                \
    }
    B200_STREAM_CASE(1)
    B200_STREAM_CASE(2)
    B200_STREAM_CASE(4)
#undef B200_STREAM_CASE
    return fail("gemv_stream: unreachable");
}

}  // namespace b200
