// tp_ops.cu -- building blocks of the tensor-parallel ternary linear:
// Megatron-style sharding, q/k/v and gate/up by output rows (no exchange), o and down by input
// columns. For the row-parallel layers each GPU holds K/tp of the activation, so the per-token
// absmax int8 quantiser (crates/bitnet-kernels/src/cpu/quantize.rs:32-49) is split in two:
//   1. row_absmax      local max|x| per token        -> all-reduce(MAX) of M floats (NCCL, by the host)
//   2. quant_absmax    scale = absmax/127, q = clamp(round(v * (1/scale)))  with the GLOBAL absmax
// so every GPU uses exactly the scale the unsharded quantiser would, its int32 partial accumulators
// add up to the tp = 1 accumulator bit for bit (all-reduce(SUM) on int32), and
//   3. residual_acc    x += act_scale * float(acc)   is the same one-multiply epilogue as tp = 1.
// The reference has only the in-process simulation of these semantics
// (crates/bitnet-inference/src/tensor_parallel.rs:123-290: partition_tensor / all_reduce Sum, Max).
// Also here: the decode glue the TP step needs on the local heads (RoPE + KV append + attention),
// exported so that a host-side decoder can run it per rank.
#include "common.cuh"

namespace b200 {

// q = clamp(round_half_away(v * inv), -127, 127) as i8; NaN -> 0 (Rust `as i8`) -- as quant.cu
__device__ __forceinline__ int tp_quant1(float v, float inv) {
    float r = roundf(v * inv);
    if (r != r) return 0;
    r = fminf(fmaxf(r, -127.f), 127.f);
    return (int)r;
}

__device__ __forceinline__ float tp_block_max(float v, float* sh) {
    v = warp_max(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    float r = lane < nw ? sh[lane] : 0.f;
    r = warp_max(r);
    __syncthreads();
    return r;
}

// One CTA per token row. MODE 0: v = x[row, k]. MODE 2: SwiGLU glue on an interleaved (gate, up)
// row of 2K floats, v = silu(gate) * up (crates/bitnet-transformer/src/lib.rs:756-789), also
// written to h[row, k] so the quantiser reads it back instead of recomputing expf.
template <int MODE>
__global__ void __launch_bounds__(256) tp_row_absmax_kernel(const float* __restrict__ x, int K, float* __restrict__ h,
                                                           float* __restrict__ absmax) {
    __shared__ float sh[32];
    const int row = blockIdx.x;
    const float* xr = x + (size_t)row * (MODE == 2 ? 2 * (size_t)K : (size_t)K);
    float m = 0.f;
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        float v;
        if (MODE == 2) {
            const float2 gu = *reinterpret_cast<const float2*>(xr + 2 * (size_t)k);
            v = silu_mul(gu.x, gu.y);
            h[(size_t)row * K + k] = v;
        } else {
            v = xr[k];
        }
        m = fmaxf(m, fabsf(v));
    }
    m = tp_block_max(m, sh);
    if (threadIdx.x == 0) absmax[row] = m;
}

// quantize_symmetric_i8 with a given (global) absmax per row; bit-identical to the unsharded
// quantiser on the same values: scale = absmax / 127, inv = 1 / scale, all-zero row -> scale 0
__global__ void __launch_bounds__(256) tp_quant_absmax_kernel(const float* __restrict__ x, const float* __restrict__ absmax,
                                                             int K, int8_t* __restrict__ q, float* __restrict__ scale) {
    const int row = blockIdx.x;
    const float amax = absmax[row];
    const float sc = amax == 0.f ? 0.f : amax / 127.f;
    const float inv = amax == 0.f ? 0.f : 1.0f / sc;
    if (threadIdx.x == 0) scale[row] = sc;
    const float* xr = x + (size_t)row * K;
    int8_t* qr = q + (size_t)row * K;
    for (int k = threadIdx.x; k < K; k += blockDim.x) qr[k] = (int8_t)(amax == 0.f ? 0 : tp_quant1(xr[k], inv));
}

// ---- vector forms of the two row kernels above (K % 4 == 0, 16-byte aligned rows) ---------------------------------
// The scalar kernels walk a row with one dependent 4/8-byte load per thread and iteration: at batch 64 that is 64 CTAs
// of 256 threads x 56 serial iterations on the 8B FFN row (measured 31 us per call, the largest glue item of the
// tensor-parallel step). Here a CTA is 1024 threads, every load is 16 bytes and four are in flight per thread. max and
// the per-element quantiser do not depend on the order, so the results are the same bits.
template <int MODE>
__global__ void __launch_bounds__(1024) tp_row_absmax_vec_kernel(const float* __restrict__ x, int K, float* __restrict__ h,
                                                                float* __restrict__ absmax) {
    __shared__ float sh[32];
    const int row = blockIdx.x, tid = threadIdx.x, nt = blockDim.x;
    float m = 0.f;
    if (MODE == 2) {
        const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * 2 * (size_t)K);  // (g, u, g, u)
        float2* hr = reinterpret_cast<float2*>(h + (size_t)row * K);
        const int n2 = K >> 1;
#pragma unroll 4
        for (int i = tid; i < n2; i += nt) {
            const float4 gu = xr[i];
            const float h0 = silu_mul(gu.x, gu.y), h1 = silu_mul(gu.z, gu.w);
            hr[i] = make_float2(h0, h1);
            m = fmaxf(m, fmaxf(fabsf(h0), fabsf(h1)));
        }
    } else {
        const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * K);
        const int n4 = K >> 2;
#pragma unroll 4
        for (int i = tid; i < n4; i += nt) {
            const float4 v = xr[i];
            m = fmaxf(m, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
        }
    }
    m = tp_block_max(m, sh);
    if (tid == 0) absmax[row] = m;
}

__global__ void __launch_bounds__(1024) tp_quant_absmax_vec_kernel(const float* __restrict__ x, const float* __restrict__ absmax,
                                                                  int K, int8_t* __restrict__ q, float* __restrict__ scale) {
    const int row = blockIdx.x;
    const float amax = absmax[row];
    const float sc = amax == 0.f ? 0.f : amax / 127.f;
    const float inv = amax == 0.f ? 0.f : 1.0f / sc;
    if (threadIdx.x == 0) scale[row] = sc;
    const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * K);
    uint32_t* qr = reinterpret_cast<uint32_t*>(q + (size_t)row * K);
    const int n4 = K >> 2;
#pragma unroll 4
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        const float4 v = xr[i];
        uint32_t o = 0;
        if (amax != 0.f)
            o = (uint32_t)(uint8_t)tp_quant1(v.x, inv) | ((uint32_t)(uint8_t)tp_quant1(v.y, inv) << 8) |
                ((uint32_t)(uint8_t)tp_quant1(v.z, inv) << 16) | ((uint32_t)(uint8_t)tp_quant1(v.w, inv) << 24);
        qr[i] = o;
    }
}

// threads of a row CTA: enough for one 16-byte item each, a multiple of 32, at most 1024
static inline int tp_row_threads(int items) {
    int t = (items + 31) / 32 * 32;
    return t < 64 ? 64 : t > 1024 ? 1024 : t;
}

// x[m, n] = x[m, n] + act_scale[m] * float(acc[m, n])   (fmul_rn then fadd_rn, as EPI_RESIDUAL)
__global__ void __launch_bounds__(256) tp_residual_acc_kernel(const int32_t* __restrict__ acc, const float* __restrict__ act_scale,
                                                             float* __restrict__ x, int N, size_t total) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const size_t m = i / (size_t)N;
    x[i] = __fadd_rn(x[i], __fmul_rn(act_scale[m], (float)acc[i]));
}

// ---- decode glue on the local heads (same arithmetic as decode.cu's generic kernels) -------------
// RoPE (split-half pairs (i, i + D/2)) on q in place and on k into the cache; v into the cache.
// grid (NH + 2*NKV, B), block D/2. qkv row = [q (NH*D) | k (NKV*D) | v (NKV*D)].
__global__ void tp_rope_kv_kernel(float* __restrict__ qkv, const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                                  float* __restrict__ kc, float* __restrict__ vc, const int* __restrict__ pos_p, int NH,
                                  int NKV, int D, int max_ctx) {
    const int head = blockIdx.x, b = blockIdx.y, i = threadIdx.x, half = D / 2;
    const int pos = *pos_p;
    float* row = qkv + (size_t)b * (NH + 2 * NKV) * D;
    if (head < NH + NKV) {
        float* v = row + (size_t)head * D;
        const float c = cos_t[(size_t)pos * half + i], s = sin_t[(size_t)pos * half + i];
        const float a = v[i], bb = v[i + half];
        const float r0 = a * c - bb * s, r1 = a * s + bb * c;
        if (head < NH) {
            v[i] = r0;
            v[i + half] = r1;
        } else {
            float* dst = kc + (((size_t)b * NKV + (head - NH)) * max_ctx + pos) * D;
            dst[i] = r0;
            dst[i + half] = r1;
        }
    } else {
        const int g = head - NH - NKV;
        const float* v = row + (size_t)(NH + NKV + g) * D;
        float* dst = vc + (((size_t)b * NKV + g) * max_ctx + pos) * D;
        dst[i] = v[i];
        dst[i + half] = v[i + half];
    }
}

// Decode attention for one (head, sequence): scores over 0..pos, softmax, weighted V sum.
// grid (NH, B), block 128 threads, dynamic smem = max_ctx floats.
__global__ void __launch_bounds__(128) tp_attn_decode_kernel(const float* __restrict__ qkv, const float* __restrict__ kc,
                                                            const float* __restrict__ vc, float* __restrict__ att,
                                                            const int* __restrict__ pos_p, int NH, int NKV, int D,
                                                            int max_ctx) {
    extern __shared__ float sc[];
    __shared__ float red[32];
    const int head = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int pos = *pos_p, g = head / (NH / NKV);
    const float* q = qkv + (size_t)b * (NH + 2 * NKV) * D + (size_t)head * D;
    const float* kb = kc + ((size_t)b * NKV + g) * max_ctx * D;
    const float* vb = vc + ((size_t)b * NKV + g) * max_ctx * D;
    const float scale = 1.0f / sqrtf((float)D);
    float qr[8];
    const int per = D / 32;
    for (int j = 0; j < per; ++j) qr[j] = q[lane + 32 * j];
    float mx = -INFINITY;
    for (int p = warp; p <= pos; p += nwarps) {
        const float* kp = kb + (size_t)p * D;
        float d = 0.f;
        for (int j = 0; j < per; ++j) d = fmaf(qr[j], kp[lane + 32 * j], d);
        d = warp_sum(d) * scale;
        if (lane == 0) sc[p] = d;
        mx = fmaxf(mx, d);
    }
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    mx = red[0];
    for (int w = 1; w < nwarps; ++w) mx = fmaxf(mx, red[w]);
    __syncthreads();
    float sum = 0.f;
    for (int p = tid; p <= pos; p += blockDim.x) {
        const float e = expf(sc[p] - mx);
        sc[p] = e;
        sum += e;
    }
    sum = warp_sum(sum);
    if (lane == 0) red[warp] = sum;
    __syncthreads();
    sum = 0.f;
    for (int w = 0; w < nwarps; ++w) sum += red[w];
    const float inv = 1.0f / sum;
    for (int d = tid; d < D; d += blockDim.x) {
        float a = 0.f;
        for (int p = 0; p <= pos; ++p) a = fmaf(sc[p] * inv, vb[(size_t)p * D + d], a);
        att[((size_t)b * NH + head) * D + d] = a;
    }
}

// ---- RoPE + KV append + GQA decode attention in ONE kernel, one CTA per (KV head, sequence) ---------------------------
// The two kernels above cost 27 us per layer at batch 64 (in-graph, tools/tp_layer_prof.py): one CTA per QUERY head re-reads
// its group's K and V, walks the positions one dependent row load at a time, and RoPE is a launch of its own. Here the
// CTA of a KV head rotates and appends its k / v row (kept in shared memory for the new position), and each warp owns
// one (query head of the group, slice of the positions): q rotated in registers (a RoPE pair lives in one lane), one
// pass over the positions with the softmax folded in (running max / sum), four positions' K and V rows in flight per
// step, and a log-sum-exp merge of the S position slices through shared memory at the end. The K / V rows of a group are
// read from L2 once per CTA and hit in L1 for the other query heads.
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <int PER>  // head_dim / 32
__global__ void __launch_bounds__(512) tp_rope_attn_gqa_kernel(float* __restrict__ qkv, const float* __restrict__ cos_t,
                                                               const float* __restrict__ sin_t, float* __restrict__ kc,
                                                               float* __restrict__ vc, float* __restrict__ att,
                                                               const int* __restrict__ pos_p, int NH, int NKV, int max_ctx,
                                                               int S) {
    constexpr int D = 32 * PER, HALF = D / 2, CH = PER <= 4 ? 16 : 8;  // positions per shared-memory chunk
    extern __shared__ __align__(16) float sm[];
    float* knew = sm;            // [D] the rotated k row of this step
    float* vnew = sm + D;        // [D]
    float* mg_m = sm + 2 * D;    // [warps] slice maxima   (S > 1)
    const int g = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nwarps = blockDim.x >> 5, R = NH / NKV;
    float* mg_l = mg_m + nwarps;      // [warps] slice sums
    float* mg_a = mg_m + ((2 * nwarps + 3) & ~3);  // [warps][D] slice accumulators (16-byte aligned)
    const int hq = warp / S, sp = warp - hq * S, head = g * R + hq;
    const int pos = *pos_p;
    float* row = qkv + (size_t)b * (NH + 2 * NKV) * D;
    float* kb = kc + ((size_t)b * NKV + g) * max_ctx * D;
    float* vb = vc + ((size_t)b * NKV + g) * max_ctx * D;
    const float* cp = cos_t + (size_t)pos * HALF;
    const float* sp_t = sin_t + (size_t)pos * HALF;
    // k: rotate, append, keep; v: append, keep (same arithmetic as tp_rope_kv_kernel)
    for (int i = tid; i < HALF; i += blockDim.x) {
        const float* k = row + (size_t)(NH + g) * D;
        const float c = cp[i], s = sp_t[i], a = k[i], bb = k[i + HALF];
        const float r0 = a * c - bb * s, r1 = a * s + bb * c;
        kb[(size_t)pos * D + i] = r0;
        kb[(size_t)pos * D + i + HALF] = r1;
        knew[i] = r0;
        knew[i + HALF] = r1;
    }
    for (int i = tid; i < D; i += blockDim.x) {
        const float v = row[(size_t)(NH + NKV + g) * D + i];
        vb[(size_t)pos * D + i] = v;
        vnew[i] = v;
    }
    // q of this warp's head: lane holds elements lane + 32 j; the RoPE partner of element e < HALF is e + HALF = the
    // same lane, j + PER / 2
    float qr[PER];
    {
        float* q = row + (size_t)head * D;
#pragma unroll
        for (int j = 0; j < PER; ++j) qr[j] = q[lane + 32 * j];
#pragma unroll
        for (int j = 0; j < PER / 2; ++j) {
            const float c = cp[lane + 32 * j], s = sp_t[lane + 32 * j], a = qr[j], bb = qr[j + PER / 2];
            qr[j] = a * c - bb * s;
            qr[j + PER / 2] = a * s + bb * c;
        }
    }
    __syncthreads();  // every slice of every head has read its (unrotated) q row
    if (sp == 0) {    // the contract says "RoPE on q in place": only now, or a slower slice would rotate the row twice
        float* q = row + (size_t)head * D;
#pragma unroll
        for (int j = 0; j < PER; ++j) q[lane + 32 * j] = qr[j];
    }
    const float scale = 1.0f / sqrtf((float)D);
    float m = -INFINITY, l = 0.f, acc[PER];
#pragma unroll
    for (int j = 0; j < PER; ++j) acc[j] = 0.f;
    // K / V rows go through shared memory, CH positions at a time, double-buffered with cp.async: the R query heads of
    // the group (and the S slices) read every row, and R x S separate L2 reads of it were the limiter of the first
    // version (25 MB of f32 K/V per layer at batch 64 turned into 100 MB of L2 -> SM traffic).
    float* kbuf = mg_a + (S > 1 ? (size_t)nwarps * D : 0);  // [2][CH][D]
    float* vbuf = kbuf + 2 * CH * D;                         // [2][CH][D]
    const int nchunks = pos / CH + 1;
    auto load_chunk = [&](int c, int buf) {
        constexpr int D4 = D / 4;
        for (int idx = tid; idx < CH * D4; idx += blockDim.x) {
            const int p = c * CH + idx / D4, e = (idx % D4) * 4;
            float* kd = kbuf + ((size_t)buf * CH) * D + (size_t)idx * 4;
            float* vd = vbuf + ((size_t)buf * CH) * D + (size_t)idx * 4;
            if (p < pos) {
                cp_async16(kd, kb + (size_t)p * D + e);
                cp_async16(vd, vb + (size_t)p * D + e);
            } else if (p == pos) {  // this step's row: from the CTA's own copy
                *reinterpret_cast<float4*>(kd) = *reinterpret_cast<const float4*>(knew + e);
                *reinterpret_cast<float4*>(vd) = *reinterpret_cast<const float4*>(vnew + e);
            }
        }
        cp_async_commit();
    };
    load_chunk(0, 0);
    for (int c = 0; c < nchunks; ++c) {
        if (c + 1 < nchunks) {
            load_chunk(c + 1, (c + 1) & 1);
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const float* kch = kbuf + ((size_t)(c & 1) * CH) * D;
        const float* vch = vbuf + ((size_t)(c & 1) * CH) * D;
        for (int i0 = sp; i0 < CH && c * CH + i0 <= pos; i0 += 4 * S) {
            float sc[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * S;
                ok[u] = i < CH && c * CH + i <= pos;
                const float* kp = kch + (size_t)(ok[u] ? i : i0) * D;
                float d = 0.f;
#pragma unroll
                for (int j = 0; j < PER; ++j) d = fmaf(qr[j], kp[lane + 32 * j], d);
                sc[u] = d;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
                for (int u = 0; u < 4; ++u) sc[u] += __shfl_xor_sync(0xffffffffu, sc[u], o);
            }
            float mn = m;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                sc[u] = ok[u] ? sc[u] * scale : -INFINITY;
                mn = fmaxf(mn, sc[u]);
            }
            const float corr = expf(m - mn);  // first step: exp(-inf) = 0
            float w[4], ws = 0.f;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                w[u] = expf(sc[u] - mn);      // masked positions: exp(-inf) = 0
                ws += w[u];
            }
            l = l * corr + ws;
#pragma unroll
            for (int j = 0; j < PER; ++j) {
                float a = acc[j] * corr;
#pragma unroll
                for (int u = 0; u < 4; ++u) a = fmaf(w[u], vch[(size_t)(ok[u] ? i0 + u * S : i0) * D + lane + 32 * j], a);
                acc[j] = a;
            }
            m = mn;
        }
        __syncthreads();  // the buffer is refilled two chunks later
    }
    if (S > 1) {  // merge the S slices of a head (log-sum-exp); a slice that saw no position has m = -inf, l = 0
        if (lane == 0) {
            mg_m[warp] = m;
            mg_l[warp] = l;
        }
#pragma unroll
        for (int j = 0; j < PER; ++j) mg_a[(size_t)warp * D + lane + 32 * j] = acc[j];
        __syncthreads();
        if (sp != 0) return;
        float M = -INFINITY;
        for (int t = 0; t < S; ++t) M = fmaxf(M, mg_m[hq * S + t]);
        l = 0.f;
#pragma unroll
        for (int j = 0; j < PER; ++j) acc[j] = 0.f;
        for (int t = 0; t < S; ++t) {
            const float f = expf(mg_m[hq * S + t] - M);
            l = fmaf(mg_l[hq * S + t], f, l);
#pragma unroll
            for (int j = 0; j < PER; ++j) acc[j] = fmaf(mg_a[(size_t)(hq * S + t) * D + lane + 32 * j], f, acc[j]);
        }
    }
    const float inv = 1.0f / l;
#pragma unroll
    for (int j = 0; j < PER; ++j) att[((size_t)b * NH + head) * D + lane + 32 * j] = acc[j] * inv;
}

static size_t rope_attn_gqa_smem(int R, int S, int D) {
    const int warps = R * S, CH = D <= 128 ? 16 : 8;
    return ((size_t)2 * D + ((2 * warps + 3) & ~3) + (S > 1 ? (size_t)warps * D : 0) + (size_t)4 * CH * D) * 4;
}

template <int PER>
static int launch_rope_attn_gqa(const bitnet_b200_ctx* ctx, float* qkv, const float* cos_t, const float* sin_t, float* kc,
                                float* vc, const int32_t* pos_dev, int B, int NH, int NKV, int max_ctx, int S, float* att) {
    const int warps = NH / NKV * S, D = 32 * PER;
    const size_t smem = rope_attn_gqa_smem(NH / NKV, S, D);
    tp_rope_attn_gqa_kernel<PER><<<dim3(NKV, B), 32 * warps, smem, ctx->stream>>>(qkv, cos_t, sin_t, kc, vc, att, pos_dev, NH, NKV,
                                                                                 max_ctx, S);
    B200_LAUNCH_CHECK("tp_rope_attn_gqa_kernel");
    return 0;
}

}  // namespace b200

using namespace b200;

namespace b200 {
int launch_tp_xchg_sum_norm_quant(const bitnet_b200_ctx*, const bitnet_b200_tp_xchg*, const float*, float*, const float*, float, int,
                                  int, int8_t*, float*, float*);
int launch_tp_xchg_max_quant(const bitnet_b200_ctx*, const bitnet_b200_tp_xchg*, const float*, int, int, int8_t*, float*);
}  // namespace b200

extern "C" {

int bitnet_b200_row_absmax(bitnet_b200_ctx* ctx, const float* x, int M, int K, int swiglu, float* h, float* absmax) {
    if (!ctx || !x || !absmax) return fail("row_absmax: NULL argument");
    if (M <= 0 || K <= 0) return fail("row_absmax: M and K must be > 0");
    if (swiglu && !h) return fail("row_absmax: the SwiGLU mode needs the h output");
    if (swiglu && (reinterpret_cast<uintptr_t>(x) & 7)) return fail("row_absmax: x must be 8-byte aligned");
    B200_CUDA(cudaSetDevice(ctx->device));
    const bool vec = K % 4 == 0 && !(reinterpret_cast<uintptr_t>(x) & 15) && !(swiglu && (reinterpret_cast<uintptr_t>(h) & 7));
    if (vec && swiglu)
        tp_row_absmax_vec_kernel<2><<<M, tp_row_threads(K / 2), 0, ctx->stream>>>(x, K, h, absmax);
    else if (vec)
        tp_row_absmax_vec_kernel<0><<<M, tp_row_threads(K / 4), 0, ctx->stream>>>(x, K, nullptr, absmax);
    else if (swiglu)
        tp_row_absmax_kernel<2><<<M, 256, 0, ctx->stream>>>(x, K, h, absmax);
    else
        tp_row_absmax_kernel<0><<<M, 256, 0, ctx->stream>>>(x, K, nullptr, absmax);
    B200_LAUNCH_CHECK("tp_row_absmax_kernel");
    return 0;
}

int bitnet_b200_actquant_i8_absmax(bitnet_b200_ctx* ctx, const float* x, const float* absmax, int M, int K, int8_t* q,
                                   float* scale) {
    if (!ctx || !x || !absmax || !q || !scale) return fail("actquant_i8_absmax: NULL argument");
    if (M <= 0 || K <= 0) return fail("actquant_i8_absmax: M and K must be > 0");
    B200_CUDA(cudaSetDevice(ctx->device));
    if (K % 4 == 0 && !(reinterpret_cast<uintptr_t>(x) & 15) && !(reinterpret_cast<uintptr_t>(q) & 3))
        tp_quant_absmax_vec_kernel<<<M, tp_row_threads(K / 4), 0, ctx->stream>>>(x, absmax, K, q, scale);
    else
        tp_quant_absmax_kernel<<<M, 256, 0, ctx->stream>>>(x, absmax, K, q, scale);
    B200_LAUNCH_CHECK("tp_quant_absmax_kernel");
    return 0;
}

int bitnet_b200_residual_acc(bitnet_b200_ctx* ctx, const int32_t* acc, const float* act_scale, float* x, int M, int N) {
    if (!ctx || !acc || !act_scale || !x) return fail("residual_acc: NULL argument");
    if (M <= 0 || N <= 0) return fail("residual_acc: M and N must be > 0");
    B200_CUDA(cudaSetDevice(ctx->device));
    const size_t total = (size_t)M * N;
    tp_residual_acc_kernel<<<(unsigned)div_ceil(total, 256), 256, 0, ctx->stream>>>(acc, act_scale, x, N, total);
    B200_LAUNCH_CHECK("tp_residual_acc_kernel");
    return 0;
}

int bitnet_b200_tp_xchg_sum_rmsnorm_actquant(bitnet_b200_ctx* ctx, const bitnet_b200_tp_xchg* xc, const float* act_scale,
                                             float* x, const float* gamma, float eps, int M, int N, int8_t* q, float* scale,
                                             float* normed) {
    if (!ctx || !xc || !act_scale || !x || !gamma || !q || !scale) return fail("tp_xchg_sum: NULL argument");
    if (!(eps > 0.f)) return fail("tp_xchg_sum: eps must be > 0");
    B200_CUDA(cudaSetDevice(ctx->device));
    return launch_tp_xchg_sum_norm_quant(ctx, xc, act_scale, x, gamma, eps, M, N, q, scale, normed);
}

int bitnet_b200_tp_xchg_max_actquant(bitnet_b200_ctx* ctx, const bitnet_b200_tp_xchg* xc, const float* x, int M, int K,
                                     int8_t* q, float* scale) {
    if (!ctx || !xc || !x || !q || !scale) return fail("tp_xchg_max: NULL argument");
    B200_CUDA(cudaSetDevice(ctx->device));
    return launch_tp_xchg_max_quant(ctx, xc, x, M, K, q, scale);
}

int bitnet_b200_rope_attn_decode(bitnet_b200_ctx* ctx, float* qkv, const float* cos_t, const float* sin_t, float* kc,
                                 float* vc, const int32_t* pos_dev, int B, int NH, int NKV, int D, int max_ctx, float* att) {
    if (!ctx || !qkv || !cos_t || !sin_t || !kc || !vc || !pos_dev || !att) return fail("rope_attn_decode: NULL argument");
    if (B <= 0 || NH <= 0 || NKV <= 0 || NH % NKV != 0) return fail("rope_attn_decode: invalid head counts");
    if (D % 64 != 0 || D > 256 || max_ctx <= 0 || max_ctx > 12000) return fail("rope_attn_decode: unsupported head_dim / max_ctx");
    B200_CUDA(cudaSetDevice(ctx->device));
    {
        // fused kernel: position slices per head from the context bound (the position itself lives on the device)
        const int R = NH / NKV;
        int S = max_ctx <= 256 ? 1 : max_ctx <= 1024 ? 2 : max_ctx <= 4096 ? 4 : 8;
        while (S > 1 && (R * S > 16 || rope_attn_gqa_smem(R, S, D) > 48 * 1024)) S >>= 1;
        const bool off = getenv("BITNET_B200_TP_ATTN_SPLIT") != nullptr;  // diagnostics / tests: the two-kernel form
        if (!off && R * S <= 16 && rope_attn_gqa_smem(R, S, D) <= 48 * 1024) {
            if (D == 64) return launch_rope_attn_gqa<2>(ctx, qkv, cos_t, sin_t, kc, vc, pos_dev, B, NH, NKV, max_ctx, S, att);
            if (D == 128) return launch_rope_attn_gqa<4>(ctx, qkv, cos_t, sin_t, kc, vc, pos_dev, B, NH, NKV, max_ctx, S, att);
            if (D == 256) return launch_rope_attn_gqa<8>(ctx, qkv, cos_t, sin_t, kc, vc, pos_dev, B, NH, NKV, max_ctx, S, att);
        }
    }
    tp_rope_kv_kernel<<<dim3(NH + 2 * NKV, B), D / 2, 0, ctx->stream>>>(qkv, cos_t, sin_t, kc, vc, pos_dev, NH, NKV, D, max_ctx);
    B200_LAUNCH_CHECK("tp_rope_kv_kernel");
    tp_attn_decode_kernel<<<dim3(NH, B), 128, (size_t)max_ctx * 4, ctx->stream>>>(qkv, kc, vc, att, pos_dev, NH, NKV, D, max_ctx);
    B200_LAUNCH_CHECK("tp_attn_decode_kernel");
    return 0;
}

}  // extern "C"
