// gemv_mma.cu -- decode-regime ternary GEMV for the QK256 layout, M = 1..8 tokens, HBM-bound.
//
// Replaces gemv_qk256 / gemv_qk256_row_avx2 (crates/bitnet-quantization/src/i2s_qk256.rs:346-372,
// i2s_qk256_avx2.rs:81-229) for device-resident weights, fused with the activation quantiser
// that feeds it (act_prologue.cuh); fills launch_qk256_gemv
// (crates/bitnet-kernels/src/cuda/qk256_gemv.rs:119-137).
//
// Shape of the kernel (B200-first):
//   * a CTA owns T tiles of 16 consecutive weight rows = one contiguous byte range; thread 0
//     fetches it with one 1-D bulk copy (TMA) per tile BEFORE griddepcontrol.wait, so under
//     programmatic dependent launch the weights of projection n+1 stream into shared memory
//     while projection n is still computing (a 2B-4T projection is only 0.25-1.4 us of HBM);
//   * the inner product runs on the tensor pipe, not on dp4a: the 2-bit codes of a 32-bit
//     word go through the 4-entry int8 LUT with four PRMTs, and each PRMT result IS one
//     A-fragment register of mma.sync.m16n8k32 (16 weight rows x 32 k, int8); the B fragment
//     holds up to 8 tokens of int8 activations. The MMA's k index is only a summation index,
//     so lane t simply owns 64 consecutive k of every 256-k step (one 128-bit shared load per
//     row) and the activations are read with the same permutation. int32 accumulation over
//     all of K stays in the C fragment: no per-chunk partials, bit-exact;
//   * K is split over the CTA's warps (one 256-k step at a time round-robin); partial C
//     fragments meet in shared memory once per tile;
//   * epilogues: scale / residual add / SwiGLU glue, as in gemv_stream.cu.
// This is the legacy tensor path on purpose: at M <= 8 the op is bandwidth-bound and the
// point of the MMA is to take the 16 dp4a + reduction per 64 weights off the issue slots.
// The M >= 16 regime is gemm_tcgen05.cu.
#include "act_prologue.cuh"

namespace b200 {

struct MmaParams {
    const uint8_t* codes;  // [N, stride] device layout
    PrologueArgs pa;
    float* y;               // [M, N] (EPI_SWIGLU: [M, N/2])
    const float* residual;  // EPI_RESIDUAL
    int32_t* acc_out;       // optional [M, N]
    uint32_t lut;
    int N, K, M, stride, T, nS, q_ld, epi;
};

__device__ __forceinline__ void mma_s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                       uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__global__ void __launch_bounds__(512) gemv_mma_kernel(const MmaParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    uint64_t* mbar = reinterpret_cast<uint64_t*>(smem);            // [8]
    double* shd = reinterpret_cast<double*>(smem + 64);             // [32]
    float* shf = reinterpret_cast<float*>(smem + 64 + 256);         // [32]
    float* sc_s = reinterpret_cast<float*>(smem + 64 + 256 + 128);  // [8]
    unsigned char* wbuf = smem + 512;
    int8_t* qbuf = reinterpret_cast<int8_t*>(wbuf + (size_t)p.T * 16 * p.stride);
    int4* part = reinterpret_cast<int4*>(qbuf + (size_t)p.M * p.q_ld);  // [T][nwarps][32]

    const int row_begin = blockIdx.x * p.T * 16;

    // ---- weight prefetch: independent of the previous kernel -> before griddepcontrol.wait --
    if (tid == 0) {
        for (int i = 0; i < p.T; ++i) mbar_init(&mbar[i], 1);
        mbar_fence_init();
        for (int i = 0; i < p.T; ++i) {
            const int r0 = row_begin + i * 16;
            const int nr = min(16, p.N - r0);
            if (nr <= 0) {
                mbar_arrive(&mbar[i]);
                continue;
            }
            const uint32_t bytes = (uint32_t)nr * p.stride;
            mbar_expect_tx(&mbar[i], bytes);
            bulk_g2s(wbuf + (size_t)i * 16 * p.stride, p.codes + (size_t)r0 * p.stride, bytes, &mbar[i]);
        }
    }
    pdl_launch_dependents();
    pdl_wait();       // everything below may read what the previous kernel wrote
    __syncthreads();  // mbarrier inits visible to all waiters

    act_prologue(p.pa, qbuf, p.q_ld, sc_s, shd, shf);

    // ---- main loop ------------------------------------------------------------------------
    const bool tok = g < p.M;
    const int8_t* qrow = qbuf + g * p.q_ld + 64 * t;
    for (int i = 0; i < p.T; ++i) {
        mbar_wait(&mbar[i], 0);
        int c[4] = {0, 0, 0, 0};
        const unsigned char* wr = wbuf + ((size_t)i * 16 + g) * p.stride + 16 * t;
        for (int S = warp; S < p.nS; S += nwarps) {
            const uint4 wg = *reinterpret_cast<const uint4*>(wr + S * 64);
            const uint4 wh = *reinterpret_cast<const uint4*>(wr + 8 * p.stride + S * 64);
            const uint32_t wgs[4] = {wg.x, wg.y, wg.z, wg.w};
            const uint32_t whs[4] = {wh.x, wh.y, wh.z, wh.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint4 b = make_uint4(0, 0, 0, 0);
                if (tok) b = *reinterpret_cast<const uint4*>(qrow + S * 256 + 16 * j);
                const uint4 dg = decode16(wgs[j], p.lut);
                const uint4 dh = decode16(whs[j], p.lut);
                mma_s8(c, dg.x, dh.x, dg.y, dh.y, b.x, b.y);
                mma_s8(c, dg.z, dh.z, dg.w, dh.w, b.z, b.w);
            }
        }
        part[((size_t)i * nwarps + warp) * 32 + lane] = make_int4(c[0], c[1], c[2], c[3]);
    }
    __syncthreads();

    // ---- cross-warp K reduction + epilogue: warp i finishes tile i ---------------------------
    for (int i = warp; i < p.T; i += nwarps) {
        int4 s = make_int4(0, 0, 0, 0);
        for (int w = 0; w < nwarps; ++w) {
            const int4 v = part[((size_t)i * nwarps + w) * 32 + lane];
            s.x += v.x;
            s.y += v.y;
            s.z += v.z;
            s.w += v.w;
        }
        // lane (g, t): s.x (row g, token 2t), s.y (row g, 2t+1), s.z (row g+8, 2t), s.w (row g+8, 2t+1)
        const int rA = row_begin + i * 16 + g, rB = rA + 8;
        const int m0 = 2 * t, m1 = 2 * t + 1;
        const float a0 = m0 < p.M ? sc_s[m0] : 0.f, a1 = m1 < p.M ? sc_s[m1] : 0.f;
        const float vA0 = __fmul_rn(a0, (float)s.x), vA1 = __fmul_rn(a1, (float)s.y);
        const float vB0 = __fmul_rn(a0, (float)s.z), vB1 = __fmul_rn(a1, (float)s.w);
        if (p.epi == EPI_SWIGLU) {
            // rows (2r, 2r+1) = (gate_r, up_r) sit in lanes g and g^1 = lane ^ 4
            const float uA0 = __shfl_xor_sync(0xffffffffu, vA0, 4), uA1 = __shfl_xor_sync(0xffffffffu, vA1, 4);
            const float uB0 = __shfl_xor_sync(0xffffffffu, vB0, 4), uB1 = __shfl_xor_sync(0xffffffffu, vB1, 4);
            if (!(g & 1)) {
                const int half = p.N / 2;
                if (rA < p.N) {
                    if (m0 < p.M) p.y[(size_t)m0 * half + (rA >> 1)] = silu_mul(vA0, uA0);
                    if (m1 < p.M) p.y[(size_t)m1 * half + (rA >> 1)] = silu_mul(vA1, uA1);
                }
                if (rB < p.N) {
                    if (m0 < p.M) p.y[(size_t)m0 * half + (rB >> 1)] = silu_mul(vB0, uB0);
                    if (m1 < p.M) p.y[(size_t)m1 * half + (rB >> 1)] = silu_mul(vB1, uB1);
                }
            }
        } else {
            const bool res = p.epi == EPI_RESIDUAL;
#define B200_MMA_STORE(row, m, val, acc)                                              \
    if ((row) < p.N && (m) < p.M) {                                                   \
        const size_t oi = (size_t)(m)*p.N + (row);                                    \
        p.y[oi] = res ? __fadd_rn(p.residual[oi], (val)) : (val);                     \
        if (p.acc_out) p.acc_out[oi] = (acc);                                         \
    }
            B200_MMA_STORE(rA, m0, vA0, s.x)
            B200_MMA_STORE(rA, m1, vA1, s.y)
            B200_MMA_STORE(rB, m0, vB0, s.z)
            B200_MMA_STORE(rB, m1, vB1, s.w)
#undef B200_MMA_STORE
        }
    }
}

// ---------------------------------------------------------------------------------------------
struct MmaPlan {
    int T, nS, threads, grid, q_ld;
    size_t smem;
};

int gemv_mma_tiles_override = 0;  // test/bench knob (0 = heuristic)

static bool plan_mma(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M, bool f32_prologue, MmaPlan* pl) {
    if (w->layout != BITNET_B200_LAYOUT_QK256) return false;
    if (M < 1 || M > 8) return false;
    if (w->cols % 256 != 0 || w->code_stride != w->cols / 4) return false;
    const int nS = (int)(w->cols / 256);
    const int nw = nS < 16 ? nS : 16;
    const int threads = nw * 32;
    if (f32_prologue && w->cols / 4 > (size_t)NVMAX * threads) return false;
    const size_t tiles = div_ceil(w->rows, 16);
    int T = (int)(tiles / (2 * (size_t)ctx->num_sms));
    if (gemv_mma_tiles_override > 0) T = gemv_mma_tiles_override;
    if (T < 1) T = 1;
    if (T > 8) T = 8;
    while (T > 1 && (size_t)T * 16 * w->code_stride > 96 * 1024) --T;
    pl->T = T;
    pl->nS = nS;
    pl->threads = threads;
    pl->grid = (int)div_ceil(tiles, (size_t)T);
    pl->q_ld = (int)w->cols + 64;  // consecutive tokens land in different bank halves
    pl->smem = 512 + (size_t)T * 16 * w->code_stride + (size_t)M * pl->q_ld + (size_t)T * nw * 32 * 16;
    return pl->smem <= 200 * 1024;
}

bool gemv_mma_supported(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M, bool f32_prologue) {
    MmaPlan pl;
    return plan_mma(ctx, w, M, f32_prologue, &pl);
}

// Same contract as launch_gemv_stream (1 = shape not covered, 0 = ok, -1 = error), M <= 8.
int launch_gemv_mma(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M, int pro, const int8_t* q,
                    const float* act_scale, const float* x, const float* gamma, float eps, int epi, float* y,
                    const float* residual, int32_t* acc_out, int8_t* q_out, float* scale_out, float* normed_out,
                    bool pdl) {
    MmaPlan pl;
    if (!plan_mma(ctx, w, M, pro != PRO_Q8, &pl)) return 1;
    if (pro == PRO_Q8 && (reinterpret_cast<uintptr_t>(q) & 15)) return 1;
    if (pro != PRO_Q8 && ((reinterpret_cast<uintptr_t>(x) & 15) || (gamma && (reinterpret_cast<uintptr_t>(gamma) & 15)))) return 1;
    if (epi == EPI_SWIGLU && (w->rows & 1)) return 1;
    if (q_out && (reinterpret_cast<uintptr_t>(q_out) & 3)) return 1;
    if (normed_out && (reinterpret_cast<uintptr_t>(normed_out) & 15)) return 1;
    MmaParams p{};
    p.codes = w->codes;
    p.pa.q = q;
    p.pa.act_scale = act_scale;
    p.pa.x = x;
    p.pa.gamma = gamma;
    p.pa.q_out = q_out;
    p.pa.scale_out = scale_out;
    p.pa.normed_out = normed_out;
    p.pa.eps = eps;
    p.pa.pro = pro;
    p.pa.K = (int)w->cols;
    p.pa.M = M;
    p.y = y;
    p.residual = residual;
    p.acc_out = acc_out;
    p.lut = w->lut_word;
    p.N = (int)w->rows;
    p.K = (int)w->cols;
    p.M = M;
    p.stride = (int)w->code_stride;
    p.T = pl.T;
    p.nS = pl.nS;
    p.q_ld = pl.q_ld;
    p.epi = epi;
    static bool attr_set[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 16 && !attr_set[dev]) {
        B200_CUDA(cudaFuncSetAttribute(gemv_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set[dev] = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)pl.grid);
    cfg.blockDim = dim3((unsigned)pl.threads);
    cfg.dynamicSmemBytes = pl.smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    B200_CUDA(cudaLaunchKernelEx(&cfg, gemv_mma_kernel, p));
    count_launch();
    return 0;
}

// ---- the decode fast path: tensor-pipe GEMV (QK256, M <= 8), else the dp4a streaming GEMV
// (both layouts, M <= 4); 1 = neither covers the shape ----------------------------------------
int launch_gemv_stream(const bitnet_b200_ctx*, const bitnet_b200_weight*, int M, int pro, const int8_t* q,
                       const float* act_scale, const float* x, const float* gamma, float eps, int epi, float* y,
                       const float* residual, int32_t* acc_out, int8_t* q_out, float* scale_out, float* normed_out,
                       bool pdl);
bool gemv_stream_supported(const bitnet_b200_ctx*, const bitnet_b200_weight*, int M, bool f32_prologue);

static bool mma_enabled() {
    static const bool v = getenv("BITNET_B200_NO_MMA") == nullptr;
    return v;
}

bool gemv_fast_supported(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M, bool f32_prologue) {
    return (mma_enabled() && gemv_mma_supported(ctx, w, M, f32_prologue)) || gemv_stream_supported(ctx, w, M, f32_prologue);
}

int launch_gemv_fast(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M, int pro, const int8_t* q,
                     const float* act_scale, const float* x, const float* gamma, float eps, int epi, float* y,
                     const float* residual, int32_t* acc_out, int8_t* q_out, float* scale_out, float* normed_out,
                     bool pdl) {
    // per-32-block accumulators (BitNet32-F16 acc_out) only exist in the dp4a kernels
    if (mma_enabled()) {
        const int rc = launch_gemv_mma(ctx, w, M, pro, q, act_scale, x, gamma, eps, epi, y, residual, acc_out, q_out,
                                       scale_out, normed_out, pdl);
        if (rc != 1) return rc;
    }
    return launch_gemv_stream(ctx, w, M, pro, q, act_scale, x, gamma, eps, epi, y, residual, acc_out, q_out, scale_out,
                              normed_out, pdl);
}

}  // namespace b200
