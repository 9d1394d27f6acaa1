// gemv.cu -- decode-regime ternary GEMV (M = 1..16), HBM-bandwidth bound.
//
// Replaces, for device-resident packed weights, the reference's per-token GEMV
//   gemv_qk256 / gemv_qk256_row_avx2 (crates/bitnet-quantization/src/i2s_qk256.rs:346-372,
//   i2s_qk256_avx2.rs:81-229) and the blocked/scaled template i2s_matmul_blocked
//   (crates/bitnet-kernels/src/cpu/quantized_matmul.rs:155-200); fills the stub
//   launch_qk256_gemv (crates/bitnet-kernels/src/cuda/qk256_gemv.rs:119-137).
//
// This file holds the GENERIC kernels (any K, any LUT, M up to 8 per pass); the decode fast
// path for the model shapes is gemv_stream.cu.
// Mapping (B200-first, not the reference's row-per-thread plan): thread t of a CTA owns the
// 16-weight word t of EVERY row, so its sixteen int8 activations live in registers for the
// whole kernel and a warp's weight load is one fully coalesced 128-byte line per row. Each
// thread keeps R rows of loads in flight (R*4 B x 1024 threads/SM ~ 32 KB/SM outstanding,
// what Little's law asks for at ~6.5 TB/s). 2-bit codes go through a 4-entry int8 LUT with
// one PRMT per four weights and into dp4a; row sums are reduced with a transposing
// butterfly (V-1 + log2(32/V) shuffles for V values instead of 5*V).
#include "common.cuh"

namespace b200 {

struct GemvParams {
    const uint8_t* codes;    // [N, code_stride]
    const __half* scales;    // [N, nblk32] raw f16 (SCALED) or nullptr
    const int8_t* q;         // [M, K] int8 activations
    const float* act_scale;  // [M]
    float* y;                // [M, N]
    const float* residual;   // optional [M, N]: y = residual + act_scale * acc (may alias y)
    int32_t* acc_out;        // optional: QK256 [M,N]; BitNet32 [M,N,nblk32]
    uint32_t lut;            // 4 x int8
    int N, K, M;
    int nwords;              // code_stride / 4
    int code_stride;
    int nblk32;
    int rows_per_cta;
    int clamp_scale;         // BITNET_B200_SCALE_ABS_CLAMP -> 1
};

__device__ __forceinline__ uint4 load_act16(const int8_t* q, int K, int m, int p) {
    // sixteen int8 activations of token m at k = 16p .. 16p+15, zero beyond K
    const int k0 = p * 16;
    const int8_t* src = q + (size_t)m * K + k0;
    if (k0 + 16 <= K && ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
        return *reinterpret_cast<const uint4*>(src);
    }
    uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        if (k0 + j < K) w[j >> 2] |= (uint32_t)(uint8_t)src[j] << (8 * (j & 3));
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

// MT tokens, R rows in flight, SCALED = per-32 f16 block scales, WACC = also write int32 accs
template <int MT, int R, bool SCALED, bool WACC>
__global__ void __launch_bounds__(1024) gemv_i8_kernel(const GemvParams p) {
    using acc_t = typename std::conditional<SCALED, float, int>::type;
    constexpr int V = R * MT;
    static_assert(V <= 32 && (V & (V - 1)) == 0, "R*MT must be a power of two <= 32");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    acc_t* red = reinterpret_cast<acc_t*>(smem_raw);  // [2][nwarps][V]

    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int nwarps = blockDim.x >> 5;
    const bool active = tid < p.nwords;

    // activations: registers for the whole kernel (natural k order; the code plane was
    // bit-permuted at upload so that dot16 decodes in that order)
    uint4 act[MT];
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        act[m] = make_uint4(0, 0, 0, 0);
        if (active && m < p.M) act[m] = load_act16(p.q, p.K, m, tid);
    }

    const int row_begin = blockIdx.x * p.rows_per_cta;
    const int row_end = min(p.N, row_begin + p.rows_per_cta);
    int buf = 0;
    for (int r0 = row_begin; r0 < row_end; r0 += R, buf ^= 1) {
        uint32_t w[R];
        float s[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int row = r0 + r;
            w[r] = 0;
            s[r] = 0.f;
            if (active && row < row_end) {
                w[r] = ldg_nc_u32(p.codes + (size_t)row * p.code_stride + (size_t)tid * 4);
                if (SCALED) {
                    uint16_t hb = ldg_nc_u16(p.scales + (size_t)row * p.nblk32 + (tid >> 1));
                    float sv = __half2float(__ushort_as_half(hb));
                    if (p.clamp_scale) {
                        sv = fabsf(sv);
                        sv = sv < 1e-3f ? 1e-3f : sv;  // comparisons keep NaN, as f32::clamp does
                        sv = sv > 1e3f ? 1e3f : sv;
                    }
                    s[r] = sv;
                }
            }
        }
        acc_t v[V];
#pragma unroll
        for (int r = 0; r < R; ++r) {
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                const int d = dot16(w[r], p.lut, act[m], 0);
                if (SCALED) {
                    if (WACC) {
                        // per-32 block accumulator = this half-block + the neighbouring lane's
                        const int blk = d + __shfl_xor_sync(0xffffffffu, d, 1);
                        const int row = r0 + r;
                        if (active && !(tid & 1) && row < row_end && m < p.M)
                            p.acc_out[((size_t)m * p.N + row) * p.nblk32 + (tid >> 1)] = blk;
                    }
                    v[r * MT + m] = (acc_t)((float)d * s[r]);
                } else {
                    v[r * MT + m] = (acc_t)d;
                }
            }
        }
        warp_reduce_transpose<V, acc_t>(v);
        constexpr int SHIFT = 5 - Log2<V>::value;
        if ((lane & ((1 << SHIFT) - 1)) == 0) red[(buf * nwarps + warp) * V + (lane >> SHIFT)] = v[0];
        __syncthreads();
        if (tid < V) {
            acc_t tot = 0;
            for (int wv = 0; wv < nwarps; ++wv) tot += red[(buf * nwarps + wv) * V + tid];
            const int r = tid / MT, m = tid % MT;
            const int row = r0 + r;
            if (row < row_end && m < p.M) {
                const float val = p.act_scale[m] * (float)tot;
                const size_t oi = (size_t)m * p.N + row;
                p.y[oi] = p.residual ? p.residual[oi] + val : val;
                if (WACC && !SCALED) p.acc_out[(size_t)m * p.N + row] = (int32_t)tot;
            }
        }
    }
}

// f32-activation GEMV over QK256 codes: device twin of gemv_qk256 (i2s_qk256.rs:346-372).
template <int R>
__global__ void __launch_bounds__(1024) gemv_f32_kernel(const uint8_t* __restrict__ codes,
                                                        const float* __restrict__ x,
                                                        float* __restrict__ y, int N, int K,
                                                        int nwords, int code_stride,
                                                        int rows_per_cta, float l0, float l1,
                                                        float l2, float l3) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* red = reinterpret_cast<float*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const bool active = tid < nwords;
    const int m = blockIdx.y;
    float xv[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const int k = tid * 16 + j;
        xv[j] = (active && k < K) ? x[(size_t)m * K + k] : 0.f;
    }
    const int row_begin = blockIdx.x * rows_per_cta;
    const int row_end = min(N, row_begin + rows_per_cta);
    int buf = 0;
    for (int r0 = row_begin; r0 < row_end; r0 += R, buf ^= 1) {
        uint32_t w[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int row = r0 + r;
            w[r] = (active && row < row_end)
                       ? ldg_nc_u32(codes + (size_t)row * code_stride + (size_t)tid * 4)
                       : 0u;
        }
        float v[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            float a0 = 0.f, a1 = 0.f;
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
                const uint32_t c0 = code_of(w[r], j), c1 = code_of(w[r], j + 1);
                const float w0 = c0 & 2u ? (c0 & 1u ? l3 : l2) : (c0 & 1u ? l1 : l0);
                const float w1 = c1 & 2u ? (c1 & 1u ? l3 : l2) : (c1 & 1u ? l1 : l0);
                a0 = fmaf(w0, xv[j], a0);
                a1 = fmaf(w1, xv[j + 1], a1);
            }
            v[r] = a0 + a1;
        }
        warp_reduce_transpose<R, float>(v);
        constexpr int SHIFT = 5 - Log2<R>::value;
        if ((lane & ((1 << SHIFT) - 1)) == 0) red[(buf * nwarps + warp) * R + (lane >> SHIFT)] = v[0];
        __syncthreads();
        if (tid < R) {
            float tot = 0.f;
            for (int wv = 0; wv < nwarps; ++wv) tot += red[(buf * nwarps + wv) * R + tid];
            const int row = r0 + tid;
            if (row < row_end) y[(size_t)m * N + row] = tot;
        }
    }
}

template <int MT, int R>
static int launch_gemv_mt(const GemvParams& p, int threads, int grid, size_t smem, bool scaled,
                          bool wacc, cudaStream_t st) {
    if (scaled) {
        if (wacc) gemv_i8_kernel<MT, R, true, true><<<grid, threads, smem, st>>>(p);
        else gemv_i8_kernel<MT, R, true, false><<<grid, threads, smem, st>>>(p);
    } else {
        if (wacc) gemv_i8_kernel<MT, R, false, true><<<grid, threads, smem, st>>>(p);
        else gemv_i8_kernel<MT, R, false, false><<<grid, threads, smem, st>>>(p);
    }
    B200_LAUNCH_CHECK("gemv_i8_kernel");
    return 0;
}

int gemv_rows_per_cta_override = 0;  // test/bench knob (0 = heuristic)

// y[m, n] for m in [0, M): loops token chunks of at most 8.
int launch_gemv_i8(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q,
                   const float* act_scale, int M, float* y, int32_t* acc_out, const float* residual) {
    const int nwords = (int)(w->code_stride / 4);
    if (nwords > 1024)
        return fail("linear_i8 GEMV: cols > 16384 not supported by the decode kernel");
    const bool scaled = w->layout == BITNET_B200_LAYOUT_BITNET32_F16;
    const int threads = (int)div_ceil((size_t)nwords, 32) * 32;
    for (int m0 = 0; m0 < M; m0 += 8) {
        const int mc = M - m0 < 8 ? M - m0 : 8;
        const int MT = mc == 1 ? 1 : (mc == 2 ? 2 : (mc <= 4 ? 4 : 8));
        const int R = MT <= 2 ? 8 : (MT == 4 ? 4 : 2);
        GemvParams p;
        p.codes = w->codes;
        p.scales = w->scales;
        p.q = q + (size_t)m0 * w->cols;
        p.act_scale = act_scale + m0;
        p.y = y + (size_t)m0 * w->rows;
        p.residual = residual ? residual + (size_t)m0 * w->rows : nullptr;
        p.acc_out = acc_out ? acc_out + (size_t)m0 * w->rows * (scaled ? w->nblk32 : 1) : nullptr;
        p.lut = w->lut_word;
        p.N = (int)w->rows;
        p.K = (int)w->cols;
        p.M = mc;
        p.nwords = nwords;
        p.code_stride = (int)w->code_stride;
        p.nblk32 = (int)w->nblk32;
        p.clamp_scale = w->scale_policy == BITNET_B200_SCALE_ABS_CLAMP ? 1 : 0;
        // one R-group per CTA keeps every SM busy on the small decode shapes; large N gets
        // several groups per CTA so the grid stays within a few waves.
        int rpc = R;
        const long max_ctas = (long)ctx->num_sms * 16;
        while ((long)div_ceil((size_t)p.N, (size_t)rpc) > max_ctas) rpc += R;
        if (gemv_rows_per_cta_override > 0) rpc = (int)div_ceil(gemv_rows_per_cta_override, R) * R;
        p.rows_per_cta = rpc;
        const int grid = (int)div_ceil((size_t)p.N, (size_t)rpc);
        const size_t smem = 2 * (size_t)(threads / 32) * (R * MT) * 4;
        const bool wacc = acc_out != nullptr;
        int rc;
        switch (MT) {
            case 1: rc = launch_gemv_mt<1, 8>(p, threads, grid, smem, scaled, wacc, ctx->stream); break;
            case 2: rc = launch_gemv_mt<2, 8>(p, threads, grid, smem, scaled, wacc, ctx->stream); break;
            case 4: rc = launch_gemv_mt<4, 4>(p, threads, grid, smem, scaled, wacc, ctx->stream); break;
            default: rc = launch_gemv_mt<8, 2>(p, threads, grid, smem, scaled, wacc, ctx->stream); break;
        }
        if (rc) return rc;
    }
    return 0;
}

int launch_gemv_f32(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const float* x, int M,
                    float* y) {
    if (w->layout != BITNET_B200_LAYOUT_QK256)
        return fail("linear_f32: only the QK256 layout has an f32-activation path (gemv_qk256)");
    const int nwords = (int)(w->code_stride / 4);
    if (nwords > 1024) return fail("linear_f32: cols > 16384 not supported");
    const int threads = (int)div_ceil((size_t)nwords, 32) * 32;
    constexpr int R = 8;
    int rpc = R;
    const long max_ctas = (long)ctx->num_sms * 16;
    while ((long)div_ceil(w->rows, (size_t)rpc) > max_ctas) rpc += R;
    dim3 grid((unsigned)div_ceil(w->rows, (size_t)rpc), (unsigned)M);
    const size_t smem = 2 * (size_t)(threads / 32) * R * 4;
    gemv_f32_kernel<R><<<grid, threads, smem, ctx->stream>>>(
        w->codes, x, y, (int)w->rows, (int)w->cols, nwords, (int)w->code_stride, rpc,
        (float)w->lut[0], (float)w->lut[1], (float)w->lut[2], (float)w->lut[3]);
    B200_LAUNCH_CHECK("gemv_f32_kernel");
    return 0;
}

}  // namespace b200
