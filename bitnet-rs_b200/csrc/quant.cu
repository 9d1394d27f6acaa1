// quant.cu -- activation quantiser (standalone and fused into the RMSNorm epilogue), the
// reference's weight quantisers, the unpacked-byte matmul of the KernelProvider seam, and the
// one-time BitNet32-F16 plane split done at upload.
#include "common.cuh"
#include "tp_xchg.cuh"

namespace b200 {

// ---------------------------------------------------------------------------------------
// block-wide reductions (blockDim.x multiple of 32, <= 1024)
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ float block_max(float v, float* sh) {
    v = warp_max(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    float r = lane < nw ? sh[lane] : 0.f;
    r = warp_max(r);
    __syncthreads();
    return r;
}

__device__ __forceinline__ double block_sum_f64(double v, double* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    double r = lane < nw ? sh[lane] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
    __syncthreads();
    return r;
}

// q = clamp(round_half_away(v * inv), -127, 127) as i8; NaN -> 0 (Rust `as i8`)
__device__ __forceinline__ int quant1(float v, float inv) {
    float r = roundf(v * inv);
    if (r != r) return 0;
    r = fminf(fmaxf(r, -127.f), 127.f);
    return (int)r;
}

// ---------------------------------------------------------------------------------------
// per-token absmax int8: quantize_symmetric_i8(row, 8), cpu/quantize.rs:32-49.
// One CTA per token row. MODE selects what is fused in front of the quantiser:
//   0  nothing (x is the row)
//   1  RMSNorm, cpu/layer_norm.rs:118-137: x * (1/sqrt(mean_f64(x^2) + eps)) * gamma
//   2  SwiGLU glue, transformer lib.rs:756-789: silu(gate) * up with (gate, up) interleaved
//      in the input row (2K floats)
// The fast kernel keeps the row in registers (16 values per thread, K <= 16 * blockDim);
// the generic kernel re-reads it and takes any K.
// ---------------------------------------------------------------------------------------
template <int MODE>
__device__ __forceinline__ float fused_value(const float* __restrict__ xr,
                                             const float* __restrict__ gamma, float inv_rms, int k) {
    if (MODE == 2) {
        const float g = xr[2 * k], u = xr[2 * k + 1];
        return silu_mul(g, u);
    }
    float v = xr[k];
    if (MODE == 1) v = v * inv_rms * gamma[k];
    return v;
}

template <int MODE>
__global__ void __launch_bounds__(1024) actquant_fast_kernel(const float* __restrict__ x,
                                                             const float* __restrict__ gamma,
                                                             float eps, int K,
                                                             int8_t* __restrict__ q,
                                                             float* __restrict__ scale,
                                                             float* __restrict__ normed) {
    __shared__ float shf[32];
    __shared__ double shd[32];
    const int row = blockIdx.x, tid = threadIdx.x, nt = blockDim.x;
    const float* xr = x + (size_t)row * (MODE == 2 ? 2 * K : K);
    float v[16];
    // thread owns float4 groups g = tid + j*nt (j < 4): elements 4g .. 4g+3
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int k0 = (tid + j * nt) * 4;
#pragma unroll
        for (int e = 0; e < 4; ++e) v[j * 4 + e] = 0.f;
        if (MODE == 2) {
            if (k0 + 4 <= K) {
                const float4 a = *reinterpret_cast<const float4*>(xr + 2 * k0);
                const float4 b = *reinterpret_cast<const float4*>(xr + 2 * k0 + 4);
                v[j * 4 + 0] = silu_mul(a.x, a.y);
                v[j * 4 + 1] = silu_mul(a.z, a.w);
                v[j * 4 + 2] = silu_mul(b.x, b.y);
                v[j * 4 + 3] = silu_mul(b.z, b.w);
            } else {
                for (int e = 0; e < 4; ++e)
                    if (k0 + e < K) v[j * 4 + e] = fused_value<2>(xr, gamma, 1.f, k0 + e);
            }
        } else {
            if (k0 + 4 <= K) {
                const float4 a = *reinterpret_cast<const float4*>(xr + k0);
                v[j * 4 + 0] = a.x; v[j * 4 + 1] = a.y; v[j * 4 + 2] = a.z; v[j * 4 + 3] = a.w;
            } else {
                for (int e = 0; e < 4; ++e)
                    if (k0 + e < K) v[j * 4 + e] = xr[k0 + e];
            }
        }
    }
    if (MODE == 1) {
        double ss = 0.0;
#pragma unroll
        for (int i = 0; i < 16; ++i) ss += (double)v[i] * (double)v[i];
        ss = block_sum_f64(ss, shd);
        const float ms = (float)(ss / (double)K);
        const float inv_rms = 1.0f / sqrtf(ms + eps);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int k0 = (tid + j * nt) * 4;
            if (k0 + 4 <= K) {
                const float4 g = *reinterpret_cast<const float4*>(gamma + k0);
                v[j * 4 + 0] = v[j * 4 + 0] * inv_rms * g.x;
                v[j * 4 + 1] = v[j * 4 + 1] * inv_rms * g.y;
                v[j * 4 + 2] = v[j * 4 + 2] * inv_rms * g.z;
                v[j * 4 + 3] = v[j * 4 + 3] * inv_rms * g.w;
            } else {
                for (int e = 0; e < 4; ++e)
                    if (k0 + e < K) v[j * 4 + e] = v[j * 4 + e] * inv_rms * gamma[k0 + e];
            }
        }
    }
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) amax = fmaxf(amax, fabsf(v[i]));
    amax = block_max(amax, shf);
    const float sc = amax == 0.f ? 0.f : amax / 127.f;
    const float inv = amax == 0.f ? 0.f : 1.0f / sc;
    if (tid == 0) scale[row] = sc;
    int8_t* qr = q + (size_t)row * K;
    float* nr = normed ? normed + (size_t)row * K : nullptr;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int k0 = (tid + j * nt) * 4;
        if (k0 >= K) continue;
        int b[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) b[e] = amax == 0.f ? 0 : quant1(v[j * 4 + e], inv);
        if (k0 + 4 <= K) {
            *reinterpret_cast<uint32_t*>(qr + k0) = (uint32_t)(uint8_t)b[0] | ((uint32_t)(uint8_t)b[1] << 8) |
                                                    ((uint32_t)(uint8_t)b[2] << 16) | ((uint32_t)(uint8_t)b[3] << 24);
            if (nr) *reinterpret_cast<float4*>(nr + k0) = make_float4(v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
        } else {
            for (int e = 0; e < 4; ++e)
                if (k0 + e < K) {
                    qr[k0 + e] = (int8_t)b[e];
                    if (nr) nr[k0 + e] = v[j * 4 + e];
                }
        }
    }
}

// ---------------------------------------------------------------------------------------
// Tensor-parallel fused exchange (tp_xchg.cuh): all-reduce(SUM, int32) of the row-parallel linear's partial
// accumulators + residual + RMSNorm + act-quant, one CTA per token row. The row statistics and the quantiser are
// those of actquant_fast_kernel<1> with the same thread -> element map and the same reduction trees, so the int8
// row and its scale equal what  all-reduce ; residual_acc ; rmsnorm_actquant  produce, bit for bit.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) tp_xchg_sum_norm_quant_kernel(const __grid_constant__ bitnet_b200_tp_xchg xc,
                                                                      const float* __restrict__ act_scale, float* __restrict__ x,
                                                                      const float* __restrict__ gamma, float eps, int K, int nts,
                                                                      int8_t* __restrict__ q, float* __restrict__ scale,
                                                                      float* __restrict__ normed) {
    extern __shared__ __align__(16) float xrow[];  // [K] the updated residual row
    __shared__ float shf[32];
    __shared__ double shd[32];
    const uint32_t epoch = xchg_begin(xc);
    if (!(xc.phase & 2)) return;
    const int row = blockIdx.x, tid = threadIdx.x;
    float* xr = x + (size_t)row * K;
    const float as = act_scale[row];
    // Phase 1, all threads: pull the peers' int32 partials, every rank's load of a 16-byte group in flight at once
    // (NVLink latency is paid once per group, not once per rank), add, apply the residual epilogue, keep the row in
    // shared memory. Integer sums: the order over ranks is free. (Pulling with one bulk copy per rank into a staging
    // area instead was measured at tp = 8: 3.89 ms per step against 3.66 ms with these loads -- the exchange is bound by
    // the signal / wait latency between the GPUs, not by how the 8 MB travel.)
    for (int i = tid; i < (K >> 2); i += blockDim.x) {  // K % 4 == 0 (launcher)
        int4 pv[TP_MAX_RANKS];
#pragma unroll
        for (int r = 0; r < TP_MAX_RANKS; ++r)
            if (r < xc.n_ranks) pv[r] = ld_sys_s32x4(static_cast<const int32_t*>(xc.peer_data[r]) + (size_t)row * K + 4 * i);
        int4 a = make_int4(0, 0, 0, 0);
#pragma unroll
        for (int r = 0; r < TP_MAX_RANKS; ++r)
            if (r < xc.n_ranks) { a.x += pv[r].x; a.y += pv[r].y; a.z += pv[r].z; a.w += pv[r].w; }
        float4 xv = *reinterpret_cast<const float4*>(xr + 4 * i);
        xv.x = __fadd_rn(xv.x, __fmul_rn(as, (float)a.x));  // as tp_residual_acc_kernel / EPI_RESIDUAL
        xv.y = __fadd_rn(xv.y, __fmul_rn(as, (float)a.y));
        xv.z = __fadd_rn(xv.z, __fmul_rn(as, (float)a.z));
        xv.w = __fadd_rn(xv.w, __fmul_rn(as, (float)a.w));
        *reinterpret_cast<float4*>(xr + 4 * i) = xv;
        *reinterpret_cast<float4*>(xrow + 4 * i) = xv;
    }
    __syncthreads();
    // Phase 2: the row statistics and the quantiser with the thread -> element map and the reduction trees of
    // actquant_fast_kernel<1> at nts threads; threads beyond nts hold zeros, which change neither the f64 sum nor
    // the max (x + 0.0 == x, values are magnitudes), so the bits are those of the three-kernel form.
    const int nt = nts;
    const bool act = tid < nts;
    float v[16];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int k0 = (tid + j * nt) * 4;
#pragma unroll
        for (int e = 0; e < 4; ++e) v[j * 4 + e] = 0.f;
        if (act && k0 + 4 <= K) {
            const float4 xv = *reinterpret_cast<const float4*>(xrow + k0);
            v[j * 4 + 0] = xv.x; v[j * 4 + 1] = xv.y; v[j * 4 + 2] = xv.z; v[j * 4 + 3] = xv.w;
        }
    }
    double ss = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) ss += (double)v[i] * (double)v[i];
    ss = block_sum_f64(ss, shd);
    const float ms = (float)(ss / (double)K);
    const float inv_rms = 1.0f / sqrtf(ms + eps);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int k0 = (tid + j * nt) * 4;
        if (act && k0 + 4 <= K) {
            const float4 g = *reinterpret_cast<const float4*>(gamma + k0);
            v[j * 4 + 0] = v[j * 4 + 0] * inv_rms * g.x;
            v[j * 4 + 1] = v[j * 4 + 1] * inv_rms * g.y;
            v[j * 4 + 2] = v[j * 4 + 2] * inv_rms * g.z;
            v[j * 4 + 3] = v[j * 4 + 3] * inv_rms * g.w;
        }
    }
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) amax = fmaxf(amax, fabsf(v[i]));
    amax = block_max(amax, shf);
    const float sc = amax == 0.f ? 0.f : amax / 127.f;
    const float inv = amax == 0.f ? 0.f : 1.0f / sc;
    if (tid == 0) scale[row] = sc;
    int8_t* qr = q + (size_t)row * K;
    float* nr = normed ? normed + (size_t)row * K : nullptr;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int k0 = (tid + j * nt) * 4;
        if (!act || k0 + 4 > K) continue;
        int b[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) b[e] = amax == 0.f ? 0 : quant1(v[j * 4 + e], inv);
        *reinterpret_cast<uint32_t*>(qr + k0) = (uint32_t)(uint8_t)b[0] | ((uint32_t)(uint8_t)b[1] << 8) |
                                                ((uint32_t)(uint8_t)b[2] << 16) | ((uint32_t)(uint8_t)b[3] << 24);
        if (nr) *reinterpret_cast<float4*>(nr + k0) = make_float4(v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
    }
    xchg_end(xc, epoch);
}

int launch_tp_xchg_sum_norm_quant(const bitnet_b200_ctx* ctx, const bitnet_b200_tp_xchg* xc, const float* act_scale, float* x,
                                  const float* gamma, float eps, int M, int K, int8_t* q, float* scale, float* normed) {
    if (M <= 0 || K <= 0 || K % 4 || K > 16384) return fail("tp_xchg_sum: M > 0, K % 4 == 0 and K <= 16384 required");
    if (xc->n_ranks < 1 || xc->n_ranks > TP_MAX_RANKS || xc->rank < 0 || xc->rank >= xc->n_ranks || xc->slot < 0 ||
        xc->slot >= TP_SLOTS || !(xc->phase & 3))
        return fail("tp_xchg: bad descriptor (ranks 1..8, slot 0..3, phase 1..3)");
    int nts = (int)div_ceil((size_t)K, 16);   // the thread count of actquant_fast_kernel: same element map
    nts = (int)div_ceil((size_t)nts, 32) * 32;
    if (nts < 64) nts = 64;
    int threads = (int)div_ceil((size_t)K / 4, 32) * 32;  // one 16-byte group per thread in the pull phase
    threads = threads < nts ? nts : threads > 1024 ? 1024 : threads;
    static bool attr_set[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 16 && !attr_set[dev]) {
        B200_CUDA(cudaFuncSetAttribute(tp_xchg_sum_norm_quant_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
        attr_set[dev] = true;
    }
    const int grid = (xc->phase & 2) ? M : 1;
    tp_xchg_sum_norm_quant_kernel<<<grid, threads, (size_t)K * 4, ctx->stream>>>(*xc, act_scale, x, gamma, eps, K, nts, q, scale,
                                                                                normed);
    B200_LAUNCH_CHECK("tp_xchg_sum_norm_quant_kernel");
    return 0;
}

// all-reduce(MAX) of the per-token absmax + quantize_symmetric_i8 of the local slice with it (tp_quant_absmax_kernel)
__global__ void __launch_bounds__(1024) tp_xchg_max_quant_kernel(const __grid_constant__ bitnet_b200_tp_xchg xc,
                                                                const float* __restrict__ x, int K, int8_t* __restrict__ q,
                                                                float* __restrict__ scale) {
    const uint32_t epoch = xchg_begin(xc);
    if (!(xc.phase & 2)) return;
    const int row = blockIdx.x;
    __shared__ float pm[TP_MAX_RANKS];
    if ((int)threadIdx.x < xc.n_ranks)  // one peer per thread: one NVLink round trip, not n_ranks in a chain
        pm[threadIdx.x] = ld_sys_f32(static_cast<const float*>(xc.peer_data[threadIdx.x]) + row);
    __syncthreads();
    float amax = 0.f;
    for (int r = 0; r < xc.n_ranks; ++r) amax = fmaxf(amax, pm[r]);
    const float sc = amax == 0.f ? 0.f : amax / 127.f;
    const float inv = amax == 0.f ? 0.f : 1.0f / sc;
    if (threadIdx.x == 0) scale[row] = sc;
    const float* xr = x + (size_t)row * K;
    int8_t* qr = q + (size_t)row * K;
    if (K % 4 == 0 && !(reinterpret_cast<uintptr_t>(x) & 15) && !(reinterpret_cast<uintptr_t>(q) & 3)) {
        // 16-byte loads, four in flight per thread (tp_ops.cu: tp_quant_absmax_vec_kernel); same bits per element
        const float4* x4 = reinterpret_cast<const float4*>(xr);
        uint32_t* q4 = reinterpret_cast<uint32_t*>(qr);
        const int n4 = K >> 2;
#pragma unroll 4
        for (int i = threadIdx.x; i < n4; i += blockDim.x) {
            const float4 v = x4[i];
            uint32_t o = 0;
            if (amax != 0.f)
                o = (uint32_t)(uint8_t)quant1(v.x, inv) | ((uint32_t)(uint8_t)quant1(v.y, inv) << 8) |
                    ((uint32_t)(uint8_t)quant1(v.z, inv) << 16) | ((uint32_t)(uint8_t)quant1(v.w, inv) << 24);
            q4[i] = o;
        }
    } else {
        for (int k = threadIdx.x; k < K; k += blockDim.x) qr[k] = (int8_t)(amax == 0.f ? 0 : quant1(xr[k], inv));
    }
    xchg_end(xc, epoch);
}

int launch_tp_xchg_max_quant(const bitnet_b200_ctx* ctx, const bitnet_b200_tp_xchg* xc, const float* x, int M, int K, int8_t* q,
                             float* scale) {
    if (M <= 0 || K <= 0) return fail("tp_xchg_max: M and K must be > 0");
    if (xc->n_ranks < 1 || xc->n_ranks > TP_MAX_RANKS || xc->rank < 0 || xc->rank >= xc->n_ranks || xc->slot < 0 ||
        xc->slot >= TP_SLOTS || !(xc->phase & 3))
        return fail("tp_xchg: bad descriptor (ranks 1..8, slot 0..3, phase 1..3)");
    int threads = (K / 4 + 31) / 32 * 32;  // one 16-byte item per thread where the row allows it
    threads = threads < 64 ? 64 : threads > 1024 ? 1024 : threads;
    tp_xchg_max_quant_kernel<<<(xc->phase & 2) ? M : 1, threads, 0, ctx->stream>>>(*xc, x, K, q, scale);
    B200_LAUNCH_CHECK("tp_xchg_max_quant_kernel");
    return 0;
}

template <int MODE>
__global__ void __launch_bounds__(256) actquant_generic_kernel(const float* __restrict__ x,
                                                               const float* __restrict__ gamma,
                                                               float eps, int K,
                                                               int8_t* __restrict__ q,
                                                               float* __restrict__ scale,
                                                               float* __restrict__ normed) {
    __shared__ float shf[32];
    __shared__ double shd[32];
    const int row = blockIdx.x, tid = threadIdx.x;
    const float* xr = x + (size_t)row * (MODE == 2 ? 2 * K : K);
    float inv_rms = 1.f;
    if (MODE == 1) {
        double ss = 0.0;
        for (int k = tid; k < K; k += blockDim.x) {
            const double v = (double)xr[k];
            ss += v * v;
        }
        ss = block_sum_f64(ss, shd);
        const float ms = (float)(ss / (double)K);
        inv_rms = 1.0f / sqrtf(ms + eps);
    }
    float amax = 0.f;
    for (int k = tid; k < K; k += blockDim.x) amax = fmaxf(amax, fabsf(fused_value<MODE>(xr, gamma, inv_rms, k)));
    amax = block_max(amax, shf);
    const float sc = amax == 0.f ? 0.f : amax / 127.f;
    const float inv = amax == 0.f ? 0.f : 1.0f / sc;
    if (tid == 0) scale[row] = sc;
    int8_t* qr = q + (size_t)row * K;
    float* nr = normed ? normed + (size_t)row * K : nullptr;
    for (int k = tid; k < K; k += blockDim.x) {
        const float v = fused_value<MODE>(xr, gamma, inv_rms, k);
        qr[k] = (int8_t)(amax == 0.f ? 0 : quant1(v, inv));
        if (nr) nr[k] = v;
    }
}

// mode: 0 plain, 1 rmsnorm, 2 silu(gate)*up on interleaved input
int launch_actquant(const bitnet_b200_ctx* ctx, const float* x, const float* gamma, float eps,
                    int M, int K, int8_t* q, float* scale, float* normed, int mode) {
    if (M <= 0 || K <= 0) return fail("actquant: M and K must be > 0");
    const bool aligned = (K % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0) &&
                         ((reinterpret_cast<uintptr_t>(q) & 3) == 0) &&
                         (!normed || (reinterpret_cast<uintptr_t>(normed) & 15) == 0) &&
                         (!gamma || (reinterpret_cast<uintptr_t>(gamma) & 15) == 0);
    if (aligned && K <= 16384) {
        int threads = (int)div_ceil((size_t)K, 16);
        threads = (int)div_ceil((size_t)threads, 32) * 32;
        if (threads < 64) threads = 64;
        switch (mode) {
            case 0: actquant_fast_kernel<0><<<M, threads, 0, ctx->stream>>>(x, gamma, eps, K, q, scale, normed); break;
            case 1: actquant_fast_kernel<1><<<M, threads, 0, ctx->stream>>>(x, gamma, eps, K, q, scale, normed); break;
            default: actquant_fast_kernel<2><<<M, threads, 0, ctx->stream>>>(x, gamma, eps, K, q, scale, normed); break;
        }
    } else {
        switch (mode) {
            case 0: actquant_generic_kernel<0><<<M, 256, 0, ctx->stream>>>(x, gamma, eps, K, q, scale, normed); break;
            case 1: actquant_generic_kernel<1><<<M, 256, 0, ctx->stream>>>(x, gamma, eps, K, q, scale, normed); break;
            default: actquant_generic_kernel<2><<<M, 256, 0, ctx->stream>>>(x, gamma, eps, K, q, scale, normed); break;
        }
    }
    B200_LAUNCH_CHECK("actquant_kernel");
    return 0;
}

// ---------------------------------------------------------------------------------------
// KernelProvider::quantize -- weight quantisers, cpu/fallback.rs:102-290
// one warp per block of BLOCK elements; ORs 2-bit codes into `out` (caller zeroes)
// ---------------------------------------------------------------------------------------
__global__ void quantize_weights_kernel(const float* __restrict__ in, int len, uint8_t* out,
                                        int out_len, float* __restrict__ scales, int block,
                                        int qtype) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int nblocks = (len + block - 1) / block;
    if (warp >= nblocks) return;
    const int start = warp * block, end = min(start + block, len);
    float mx = 0.f;
    for (int i = start + lane; i < end; i += 32) mx = fmaxf(mx, fabsf(in[i]));
    mx = warp_max(mx);
    const float sc = mx > 1e-8f ? mx / 1.5f : 1.0f;
    if (lane == 0) scales[warp] = sc;
    // each lane owns whole output bytes (4 consecutive elements) so the OR needs no atomics
    for (int b = start / 4 + lane; b * 4 < end; b += 32) {
        uint32_t byte = 0;
        for (int j = 0; j < 4; ++j) {
            const int i = b * 4 + j;
            if (i < start || i >= end) continue;
            const float nv = in[i] / sc;
            uint32_t code;
            if (qtype == 0) {
                code = nv > 0.5f ? 1u : (nv < -0.5f ? 3u : 0u);
            } else {
                const float l0 = qtype == 1 ? -1.0f : -1.2f, l1 = qtype == 1 ? -0.33f : -0.4f;
                const float l2 = qtype == 1 ? 0.33f : 0.4f, l3 = qtype == 1 ? 1.0f : 1.2f;
                float bd = fabsf(nv - l0);
                code = 0;
                float d = fabsf(nv - l1);
                if (d < bd) { bd = d; code = 1; }
                d = fabsf(nv - l2);
                if (d < bd) { bd = d; code = 2; }
                d = fabsf(nv - l3);
                if (d < bd) { bd = d; code = 3; }
            }
            byte |= code << (2 * j);
        }
        if (b < out_len) out[b] |= (uint8_t)byte;
    }
}

int launch_quantize_weights(cudaStream_t st, const float* in, int len, uint8_t* out, int out_len,
                            float* scales, int qtype) {
    const int block = qtype == 0 ? 32 : (qtype == 1 ? 64 : 128);
    const int nblocks = (len + block - 1) / block;
    const int threads = 256, wpb = threads / 32;
    quantize_weights_kernel<<<(nblocks + wpb - 1) / wpb, threads, 0, st>>>(in, len, out, out_len,
                                                                           scales, block, qtype);
    B200_LAUNCH_CHECK("quantize_weights_kernel");
    return 0;
}

// ---------------------------------------------------------------------------------------
// KernelProvider::matmul_i2s on UNPACKED bytes: C[m,n] = sum_l a[m,l] * b[l,n]
// (cpu/fallback.rs:39-83). b is [k, n] row-major, one byte per weight, any value 0..255.
// Each thread owns 4 adjacent columns; a 4x4 byte block of b is transposed with PRMT so
// that dp4a (signed a x unsigned b) consumes it. Up to 8 rows of a per pass, staged in smem.
// ---------------------------------------------------------------------------------------
constexpr int MM_ROWS = 8;
__global__ void __launch_bounds__(128) matmul_u8_kernel(const int8_t* __restrict__ a,
                                                        const uint8_t* __restrict__ b,
                                                        float* __restrict__ c, int m, int n,
                                                        int k) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* as = reinterpret_cast<uint32_t*>(smem_raw);  // [MM_ROWS][k4] words, zero padded
    const int k4 = (k + 3) / 4;
    const int m0 = blockIdx.y * MM_ROWS;
    const int mr = min(MM_ROWS, m - m0);
    for (int i = threadIdx.x; i < MM_ROWS * k4; i += blockDim.x) {
        const int r = i / k4, kw = i % k4;
        uint32_t wv = 0;
        if (r < mr)
            for (int j = 0; j < 4; ++j) {
                const int kk = kw * 4 + j;
                if (kk < k) wv |= (uint32_t)(uint8_t)a[(size_t)(m0 + r) * k + kk] << (8 * j);
            }
        as[i] = wv;
    }
    __syncthreads();
    const int col0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (col0 >= n) return;
    const bool full = (col0 + 4 <= n) && (n % 4 == 0);
    int acc[MM_ROWS][4];
#pragma unroll
    for (int r = 0; r < MM_ROWS; ++r)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[r][j] = 0;
    for (int kw = 0; kw < k4; ++kw) {
        uint32_t rowsw[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int kk = kw * 4 + j;
            uint32_t v = 0;
            if (kk < k) {
                const uint8_t* src = b + (size_t)kk * n + col0;
                if (full) {
                    v = *reinterpret_cast<const uint32_t*>(src);
                } else {
                    for (int t = 0; t < 4; ++t)
                        if (col0 + t < n) v |= (uint32_t)src[t] << (8 * t);
                }
            }
            rowsw[j] = v;
        }
        // transpose the 4x4 byte block: colw[t] = bytes t of rows 0..3
        const uint32_t t01l = prmt(rowsw[0], rowsw[1], 0x5140u);  // r0b0 r1b0 r0b1 r1b1
        const uint32_t t01h = prmt(rowsw[0], rowsw[1], 0x7362u);  // r0b2 r1b2 r0b3 r1b3
        const uint32_t t23l = prmt(rowsw[2], rowsw[3], 0x5140u);
        const uint32_t t23h = prmt(rowsw[2], rowsw[3], 0x7362u);
        uint32_t colw[4];
        colw[0] = prmt(t01l, t23l, 0x5410u);
        colw[1] = prmt(t01l, t23l, 0x7632u);
        colw[2] = prmt(t01h, t23h, 0x5410u);
        colw[3] = prmt(t01h, t23h, 0x7632u);
#pragma unroll
        for (int r = 0; r < MM_ROWS; ++r) {
            const uint32_t av = as[r * k4 + kw];
#pragma unroll
            for (int t = 0; t < 4; ++t) acc[r][t] = dp4a_su(av, colw[t], acc[r][t]);
        }
    }
    for (int r = 0; r < mr; ++r)
        for (int t = 0; t < 4; ++t)
            if (col0 + t < n) c[(size_t)(m0 + r) * n + col0 + t] = (float)acc[r][t];
}

int launch_matmul_u8(cudaStream_t st, const int8_t* a, const uint8_t* b, float* c, int m, int n,
                     int k) {
    const size_t smem = (size_t)MM_ROWS * ((k + 3) / 4) * 4;
    if (smem > 200 * 1024) return fail("matmul_i2s: k too large for the staging buffer");
    if (smem > 48 * 1024)
        B200_CUDA(cudaFuncSetAttribute(matmul_u8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    dim3 grid((unsigned)div_ceil((size_t)n, 4 * 128), (unsigned)div_ceil((size_t)m, MM_ROWS));
    matmul_u8_kernel<<<grid, 128, smem, st>>>(a, b, c, m, n, k);
    B200_LAUNCH_CHECK("matmul_u8_kernel");
    return 0;
}

// ---------------------------------------------------------------------------------------
// launch_i2s_matmul fill (cuda/quantized_matmul.rs:282-303; numerics of
// cpu/quantized_matmul.rs:57-96): f32 activations, map C packed [n, ceil(k/4)], f32 scales
// per (col, k-block). One CTA per output column, all m rows.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) i2s_matmul_f32_kernel(const float* __restrict__ act,
                                                             const uint8_t* __restrict__ packed,
                                                             const float* __restrict__ scales,
                                                             float* __restrict__ out, int m, int n,
                                                             int k, int bs) {
    __shared__ float sh[4];
    const int col = blockIdx.x, packed_k = (k + 3) / 4, nbk = (k + bs - 1) / bs;
    for (int row = 0; row < m; ++row) {
        float acc = 0.f;
        for (int idx = threadIdx.x; idx < k; idx += blockDim.x) {
            const uint32_t bits = (packed[(size_t)col * packed_k + idx / 4] >> ((idx % 4) * 2)) & 3u;
            const float wv = bits == 1u ? 1.f : (bits == 3u ? -1.f : 0.f);
            acc = fmaf(act[(size_t)row * k + idx], wv * scales[(size_t)col * nbk + idx / bs], acc);
        }
        acc = warp_sum(acc);
        if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) out[(size_t)row * n + col] = sh[0] + sh[1] + sh[2] + sh[3];
        __syncthreads();
    }
}

int launch_i2s_matmul_f32(cudaStream_t st, const float* act, const uint8_t* packed,
                          const float* scales, float* out, int m, int n, int k, int bs) {
    i2s_matmul_f32_kernel<<<n, 128, 0, st>>>(act, packed, scales, out, m, n, k, bs);
    B200_LAUNCH_CHECK("i2s_matmul_f32_kernel");
    return 0;
}

// ---------------------------------------------------------------------------------------
// Legacy bench-only kernel qk256_gemv_scalar (crates/bitnet-quantization/src/qk256_dispatch.rs:45-99):
// map D {-1, 0, +1, -1}, one f32 scale per (row, 256-block), f32 activations;
// out[r] = sum_blk scale[r, blk] * (sum_{e < 256} D[code(r, blk, e)] * x[blk * 256 + e]).
// One warp per row; a lane owns 8 consecutive elements (two bytes) of every block, the block sum is
// formed before the scale is applied, as in the reference (only the order inside a block differs).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) qk256_legacy_kernel(const uint8_t* __restrict__ packed, const float* __restrict__ scales,
                                                           const float* __restrict__ x, float* __restrict__ out, int rows,
                                                           int nblk) {
    const int row = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= rows) return;
    float row_sum = 0.f;
    for (int b = 0; b < nblk; ++b) {
        const size_t gb = (size_t)row * nblk + b;
        const uint16_t two = *reinterpret_cast<const uint16_t*>(packed + gb * 64 + 2 * lane);
        const float* xb = x + (size_t)b * 256 + 8 * lane;
        float bs = 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const uint32_t c = (two >> (2 * e)) & 3u;
            const float wv = c == 2u ? 1.f : (c == 1u ? 0.f : -1.f);
            bs = fmaf(wv, xb[e], bs);
        }
        bs = warp_sum(bs);
        row_sum = fmaf(bs, scales[gb], row_sum);
    }
    if (lane == 0) out[row] = row_sum;
}

int launch_qk256_legacy(cudaStream_t st, const uint8_t* packed, const float* scales, const float* x, float* out, int rows,
                        int nblk) {
    qk256_legacy_kernel<<<(unsigned)div_ceil((size_t)rows, 4), 128, 0, st>>>(packed, scales, x, out, rows, nblk);
    B200_LAUNCH_CHECK("qk256_legacy_kernel");
    return 0;
}

// ---------------------------------------------------------------------------------------
// BitNet32-F16 upload: split [rows, cols/32, 10 B] into a code plane [rows, cols/4] and an
// f16 scale plane [rows, cols/32] (lossless; 10-byte blocks cannot be loaded 16 B aligned)
// ---------------------------------------------------------------------------------------
__global__ void split_bitnet32_kernel(const uint8_t* __restrict__ src, size_t nblocks,
                                      uint8_t* __restrict__ codes, __half* __restrict__ scales,
                                      float* __restrict__ scales_f32, int clamp) {
    const size_t b = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nblocks) return;
    const uint8_t* s = src + b * 10;
    uint32_t lo = 0, hi = 0;
    for (int j = 0; j < 4; ++j) {
        lo |= (uint32_t)s[j] << (8 * j);
        hi |= (uint32_t)s[4 + j] << (8 * j);
    }
    reinterpret_cast<uint2*>(codes)[b] = make_uint2(repack_word(lo), repack_word(hi));
    const uint16_t hb = (uint16_t)(s[8] | (s[9] << 8));
    scales[b] = __ushort_as_half(hb);
    float sv = __half2float(__ushort_as_half(hb));
    if (clamp) {
        sv = fabsf(sv);
        sv = sv < 1e-3f ? 1e-3f : sv;
        sv = sv > 1e3f ? 1e3f : sv;
    }
    scales_f32[b] = sv;
}

int launch_split_bitnet32(cudaStream_t st, const uint8_t* src, size_t nblocks, uint8_t* codes,
                          __half* scales, float* scales_f32, int clamp) {
    const int threads = 256;
    split_bitnet32_kernel<<<(unsigned)div_ceil(nblocks, threads), threads, 0, st>>>(
        src, nblocks, codes, scales, scales_f32, clamp);
    B200_LAUNCH_CHECK("split_bitnet32_kernel");
    return 0;
}

// QK256 upload: bit-permute every 32-bit word of the code plane into the device layout
// (common.cuh, repack_word) -- lossless, same size, done once.
__global__ void repack_codes_kernel(uint32_t* __restrict__ words, size_t nwords) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nwords) words[i] = repack_word(words[i]);
}

int launch_repack_codes(cudaStream_t st, uint8_t* codes, size_t nbytes) {
    const size_t nwords = nbytes / 4;
    if (nwords == 0) return 0;
    repack_codes_kernel<<<(unsigned)div_ceil(nwords, 256), 256, 0, st>>>(reinterpret_cast<uint32_t*>(codes), nwords);
    B200_LAUNCH_CHECK("repack_codes_kernel");
    return 0;
}

}  // namespace b200
