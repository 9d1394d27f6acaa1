// decode_fused.cuh -- glue kernels of the fused decode path (batch <= 4): 5 launches per layer,
// chained with programmatic dependent launch. Every kernel calls griddepcontrol.launch_dependents
// first and touches activations only after griddepcontrol.wait. Included by decode.cu.
#pragma once
#include "common.cuh"

namespace b200 {

__global__ void embed_pdl_kernel(const __half* __restrict__ emb, const int* __restrict__ tokens,
                                 float* __restrict__ x, int H) {
    pdl_launch_dependents();
    pdl_wait();
    const int b = blockIdx.x;
    const __half* row = emb + (size_t)tokens[b] * H;
    for (int i = threadIdx.x; i < H; i += blockDim.x) x[(size_t)b * H + i] = __half2float(row[i]);
}

// RoPE (split-half, crates/bitnet-transformer/src/lib.rs:134-184) on q and on the new k, KV-cache
// append (:1171-1202), GQA decode attention (:374-540): one CTA per (query head, sequence),
// ATTN_THREADS threads. The first head of each KV group appends k/v to the cache; every head
// takes the freshly rotated k / v of position `pos` from shared memory, so no CTA reads a cache
// line that another CTA of this launch writes.
// Latency-shaped, not bandwidth-shaped (a decode step at short context is a chain of ~150
// dependent launches): positions are spread over all 16 warps and every thread keeps four
// independent cache loads in flight, in both the score pass and the value pass.
constexpr int ATTN_THREADS = 512;
__global__ void __launch_bounds__(ATTN_THREADS) attn_fused_kernel(
    const float* __restrict__ qkv, const float* __restrict__ cos_t, const float* __restrict__ sin_t,
    float* __restrict__ kc, float* __restrict__ vc, float* __restrict__ att, const int* __restrict__ pos_p, int NH,
    int NKV, int D, int max_ctx) {
    extern __shared__ __align__(16) float sc[];  // [max_ctx] scores, then [PG][D] value partials
    __shared__ float qs[256], ks[256], vs[256];
    __shared__ float red[32];
    pdl_launch_dependents();
    pdl_wait();
    const int head = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, half = D / 2;
    const int lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int pos = *pos_p, group = NH / NKV, g = head / group;
    const float* row = qkv + (size_t)b * (NH + 2 * NKV) * D;
    float* kb = kc + ((size_t)b * NKV + g) * max_ctx * D;
    float* vb = vc + ((size_t)b * NKV + g) * max_ctx * D;
    if (tid < half) {
        const float c = cos_t[(size_t)pos * half + tid], s = sin_t[(size_t)pos * half + tid];
        const float* qh = row + (size_t)head * D;
        const float* kh = row + (size_t)(NH + g) * D;
        const float* vh = row + (size_t)(NH + NKV + g) * D;
        const float qa = qh[tid], qb = qh[tid + half], ka = kh[tid], kbv = kh[tid + half];
        qs[tid] = qa * c - qb * s;
        qs[tid + half] = qa * s + qb * c;
        const float k0 = ka * c - kbv * s, k1 = ka * s + kbv * c;
        ks[tid] = k0;
        ks[tid + half] = k1;
        const float v0 = vh[tid], v1 = vh[tid + half];
        vs[tid] = v0;
        vs[tid + half] = v1;
        if (head % group == 0) {
            kb[(size_t)pos * D + tid] = k0;
            kb[(size_t)pos * D + tid + half] = k1;
            vb[(size_t)pos * D + tid] = v0;
            vb[(size_t)pos * D + tid + half] = v1;
        }
    }
    __syncthreads();
    const float scale = 1.0f / sqrtf((float)D);
    float qr[8];
    const int per = D / 32;
    for (int j = 0; j < per; ++j) qr[j] = qs[lane + 32 * j];
    // ---- scores: warp w takes positions w, w + nwarps, ...; four positions in flight --------
    float mx = -INFINITY;
    for (int p0 = warp; p0 <= pos; p0 += 4 * nwarps) {
        float d[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = p0 + u * nwarps;
            d[u] = 0.f;
            if (p < pos) {
                const float* kp = kb + (size_t)p * D;
                for (int j = 0; j < per; ++j) d[u] = fmaf(qr[j], __ldg(kp + lane + 32 * j), d[u]);
            } else if (p == pos) {
                for (int j = 0; j < per; ++j) d[u] = fmaf(qr[j], ks[lane + 32 * j], d[u]);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int u = 0; u < 4; ++u) d[u] += __shfl_xor_sync(0xffffffffu, d[u], o);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = p0 + u * nwarps;
            if (p <= pos) {
                const float v = d[u] * scale;
                if (lane == 0) sc[p] = v;
                mx = fmaxf(mx, v);
            }
        }
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
    // ---- values: thread = (position group pg, float4 column dc); partial sums meet in smem ---
    const int ncol = D / 4, PG = blockDim.x / ncol;
    const int dc = tid % ncol, pg = tid / ncol;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (pg < PG) {
        for (int p0 = pg; p0 < pos; p0 += 4 * PG) {
            float4 v[4];
            float w[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int p = p0 + u * PG;
                v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                w[u] = 0.f;
                if (p < pos) {
                    v[u] = __ldg(reinterpret_cast<const float4*>(vb + (size_t)p * D) + dc);
                    w[u] = sc[p] * inv;
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                a.x = fmaf(w[u], v[u].x, a.x);
                a.y = fmaf(w[u], v[u].y, a.y);
                a.z = fmaf(w[u], v[u].z, a.z);
                a.w = fmaf(w[u], v[u].w, a.w);
            }
        }
    }
    const float wl = sc[pos] * inv;
    __syncthreads();  // all reads of sc[] done: reuse it for the partials
    float4* part = reinterpret_cast<float4*>(sc + ((max_ctx + 3) & ~3));
    if (pg < PG) part[pg * ncol + dc] = a;
    __syncthreads();
    if (tid < ncol) {
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int i = 0; i < PG; ++i) {
            const float4 v = part[i * ncol + tid];
            t.x += v.x;
            t.y += v.y;
            t.z += v.z;
            t.w += v.w;
        }
        const float4 vl = reinterpret_cast<const float4*>(vs)[tid];
        t.x = fmaf(wl, vl.x, t.x);
        t.y = fmaf(wl, vl.y, t.y);
        t.z = fmaf(wl, vl.z, t.z);
        t.w = fmaf(wl, vl.w, t.w);
        reinterpret_cast<float4*>(att + ((size_t)b * NH + head) * D)[tid] = t;
    }
}

// Tied-embedding logits (crates/bitnet-transformer/src/lib.rs:1599-1630) with the final RMSNorm
// fused in front (every CTA normalises the 10 KB hidden row itself) and a per-CTA argmax
// partial behind: f16 table [V, H] x f32 hidden. Thread t owns k = 8t..8t+7 of every row.
template <int MT, int R>
__global__ void __launch_bounds__(1024) logits_fused_kernel(const __half* __restrict__ emb,
                                                            const float* __restrict__ x,
                                                            const float* __restrict__ gamma, float eps,
                                                            float* __restrict__ logits,
                                                            float* __restrict__ part_val,
                                                            int* __restrict__ part_idx, int V, int H,
                                                            int M, int rows_per_cta) {
    constexpr int NV = R * MT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* red = reinterpret_cast<float*>(smem_raw);  // [2][nwarps][NV]
    __shared__ double shd[32 * MT];
    __shared__ float cand_v[NV];
    __shared__ int cand_i[NV];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int nvec = H / 8;
    const bool active = tid < nvec;
    const int row_begin = blockIdx.x * rows_per_cta;
    const int row_end = min(V, row_begin + rows_per_cta);
    // the table does not depend on the previous kernel: first rows in flight across the wait
    uint4 w[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int row = row_begin + r;
        w[r] = (active && row < row_end) ? ldg_nc_u128(emb + (size_t)row * H + (size_t)tid * 8)
                                         : make_uint4(0, 0, 0, 0);
    }
    pdl_launch_dependents();
    pdl_wait();
    float xv[MT][8];
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        double ss = 0.0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            xv[m][j] = (active && m < M) ? x[(size_t)m * H + tid * 8 + j] : 0.f;
            ss += (double)xv[m][j] * (double)xv[m][j];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        if (lane == 0) shd[warp * MT + m] = ss;
    }
    __syncthreads();
#pragma unroll
    for (int m = 0; m < MT; ++m) {
        double tot = 0.0;
        for (int wv = 0; wv < nwarps; ++wv) tot += shd[wv * MT + m];
        const float ms = (float)(tot / (double)H);
        const float inv_rms = 1.0f / sqrtf(ms + eps);
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (active) xv[m][j] = xv[m][j] * inv_rms * gamma[tid * 8 + j];
    }
    float best = -INFINITY;
    int best_i = 0x7fffffff;
    int buf = 0;
    for (int r0 = row_begin; r0 < row_end; r0 += R, buf ^= 1) {
        float v[NV];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const __half2* hp = reinterpret_cast<const __half2*>(&w[r]);
            float f[8];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 t = __half22float2(hp[j]);
                f[2 * j] = t.x;
                f[2 * j + 1] = t.y;
            }
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                float a = 0.f;
#pragma unroll
                for (int j = 0; j < 8; ++j) a = fmaf(f[j], xv[m][j], a);
                v[r * MT + m] = a;
            }
        }
        // next R rows in flight while this group is reduced
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int row = r0 + R + r;
            w[r] = (active && row < row_end) ? ldg_nc_u128(emb + (size_t)row * H + (size_t)tid * 8)
                                             : make_uint4(0, 0, 0, 0);
        }
        warp_reduce_transpose<NV, float>(v);
        constexpr int SHIFT = 5 - Log2<NV>::value;
        if ((lane & ((1 << SHIFT) - 1)) == 0) red[(buf * nwarps + warp) * NV + (lane >> SHIFT)] = v[0];
        __syncthreads();
        if (tid < NV) {
            float tot = 0.f;
            for (int wv = 0; wv < nwarps; ++wv) tot += red[(buf * nwarps + wv) * NV + tid];
            const int r = tid / MT, m = tid % MT, row = r0 + r;
            if (row < row_end && m < M) {
                logits[(size_t)m * V + row] = tot;
                if (tot > best) {  // rows come in increasing order: first maximum = lowest index
                    best = tot;
                    best_i = row;
                }
            }
        }
    }
    if (tid < NV) {
        cand_v[tid] = best;
        cand_i[tid] = best_i;
    }
    __syncthreads();
    if (tid < MT && tid < M) {
        float bv = -INFINITY;
        int bi = 0x7fffffff;
        for (int r = 0; r < R; ++r) {
            const float cv = cand_v[r * MT + tid];
            const int ci = cand_i[r * MT + tid];
            if (cv > bv || (cv == bv && ci < bi)) {
                bv = cv;
                bi = ci;
            }
        }
        part_val[(size_t)tid * gridDim.x + blockIdx.x] = bv;
        part_idx[(size_t)tid * gridDim.x + blockIdx.x] = bi;
    }
}

// greedy argmax over the per-CTA partials (lowest index on ties); feeds the token back and
// advances the position
__global__ void __launch_bounds__(256) argmax_final_kernel(const float* __restrict__ part_val,
                                                           const int* __restrict__ part_idx, int nparts,
                                                           int* __restrict__ next, int* __restrict__ tokens,
                                                           int* __restrict__ pos_p) {
    __shared__ float bv[8];
    __shared__ int bi[8];
    pdl_launch_dependents();
    pdl_wait();
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float best = -INFINITY;
    int idx = 0x7fffffff;
    for (int i = tid; i < nparts; i += blockDim.x) {
        const float v = part_val[(size_t)b * nparts + i];
        const int vi = part_idx[(size_t)b * nparts + i];
        if (v > best || (v == best && vi < idx)) {
            best = v;
            idx = vi;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ov > best || (ov == best && oi < idx)) {
            best = ov;
            idx = oi;
        }
    }
    if (lane == 0) {
        bv[warp] = best;
        bi[warp] = idx;
    }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w)
            if (bv[w] > best || (bv[w] == best && bi[w] < idx)) {
                best = bv[w];
                idx = bi[w];
            }
        if (idx == 0x7fffffff) idx = 0;
        next[b] = idx;
        tokens[b] = idx;
        if (b == 0) *pos_p = *pos_p + 1;
    }
}

}  // namespace b200
