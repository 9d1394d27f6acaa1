// gemm_f16_tcgen05.cu -- the batched / prefill regime (M >= 9) of the ternary linear for the
// BitNet32-F16 layout (32-element blocks with an inline f16 scale, crates/bitnet-models/src/quant/
// i2s.rs:66-140): tcgen05.mma kind::f16 with the fp32 accumulator in TMEM.
//
// The per-32-column scales cannot be applied after an int32 accumulation over all of K, so this path
// follows the reference's own arithmetic (quant/i2s.rs:111: w = s * LUT[code]): the decoder warps
// expand the 2-bit codes to f16 values s * LUT[code] -- exact in f16 for the LUT values {-2,-1,0,1,2}
// short of overflow -- and the int8 activations are widened to f16 (exact) by a pre-pass, so every
// product is exact in the fp32 accumulator and only the fp32 summation order differs from the
// block-ordered GEMV: y = act_scale[m] * sum_k f16(s*w)[n,k] * q[m,k], within 1e-3 relative of the
// dequantise-then-f32 reference. Same persistent warp-specialised structure as
// gemm_tcgen05.cu (see there): weights = A operand (128 rows), tokens = B operand (BN <= 256),
// k-block = 64 columns = one 128-byte swizzle row of f16; a packed stage (64 B x 128 rows) holds
// four k-blocks, one per decoder group.
#include <cuda_fp16.h>

#include <mutex>

#include "tcgen05.cuh"

namespace b200 {

constexpr int H_ROWS = 128, H_BKE = 64;   // weight rows per tile; f16 k per block (128 bytes)
constexpr int H_PKQ = 4;                  // k-blocks per packed stage
constexpr int H_A_BYTES = H_ROWS * 128, H_PK_BYTES = H_ROWS * 64;
constexpr int H_THREADS = 15 * 32, H_DEC_WARP0 = 6, H_BAR_BYTES = 1024;
constexpr size_t H_SCRATCH_BYTES = 16u << 20;
constexpr int H_MAX_TILE_COUNTERS = 8192;

struct GemmF16Params {
    const float* act_scale;   // [M]
    const float* wscale;      // [N, K/32] policy-applied block scales (f32)
    float* y;                 // [M, N]
    float* scratch;           // split-K accumulation blocks (zero between launches)
    int* counters;
    uint32_t lut;             // 4 int8 LUT entries
    int M, N, K, BN, n_rt, n_mt, S, n_pk, n_b, n_a, n_items;
};

__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// instruction descriptor, kind::f16: D = f32 (bits 4-5 = 1), A = B = f16 (bits 7-9, 10-12 = 0), both K-major,
// N >> 3 at bits 17-22, M >> 4 at bits 24-28
__device__ __forceinline__ uint32_t umma_idesc_f16(int m, int n) {
    return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

struct ItemH {
    int row0, m0, kb0, kb1, tile, split;
};
__device__ __forceinline__ ItemH item_h(const GemmF16Params& p, int it) {
    ItemH r;
    r.split = it % p.S;
    r.tile = it / p.S;
    const int rt = r.tile % p.n_rt, mt = r.tile / p.n_rt;
    const int nkb = p.K / H_BKE;
    r.row0 = rt * H_ROWS;
    r.m0 = mt * p.BN;
    r.kb0 = (int)(((long long)nkb * r.split) / p.S);
    r.kb1 = (int)(((long long)nkb * (r.split + 1)) / p.S);
    return r;
}

// int8 activations -> f16 (exact), 4 per thread
__global__ void q8_to_f16_kernel(const int8_t* __restrict__ q, __half* __restrict__ out, size_t n4) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const char4 v = reinterpret_cast<const char4*>(q)[i];
    reinterpret_cast<__half2*>(out)[2 * i] = __halves2half2(__int2half_rn(v.x), __int2half_rn(v.y));
    reinterpret_cast<__half2*>(out)[2 * i + 1] = __halves2half2(__int2half_rn(v.z), __int2half_rn(v.w));
}

__global__ void __launch_bounds__(H_THREADS, 1)
gemm_f16_tcgen05_kernel(const __grid_constant__ CUtensorMap tm_codes, const __grid_constant__ CUtensorMap tm_q,
                        const GemmF16Params p) {
    extern __shared__ unsigned char hsm_raw[];
    unsigned char* base = hsm_raw + ((1024u - (smem_u32(hsm_raw) & 1023u)) & 1023u);  // offset arithmetic keeps the shared address space (LDS / STS, not generic LD / ST)
    const uint32_t b_stage_bytes = (uint32_t)p.BN * 128u;
    unsigned char* b_base = base + (size_t)p.n_a * H_A_BYTES;
    unsigned char* pk_base = b_base + (size_t)p.n_b * b_stage_bytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(pk_base + (size_t)p.n_pk * H_PK_BYTES);
    uint64_t* pk_full = bars;            // [8]   count 1 (TMA tx)
    uint64_t* pk_empty = bars + 8;       // [8]   count 8 (4 k-blocks x 2 decoder warps)
    uint64_t* a_full = bars + 16;        // [8]   count 64 (the threads of the owning decoder group)
    uint64_t* a_empty = bars + 24;       // [8]   count 1 (tcgen05.commit)
    uint64_t* b_full = bars + 32;        // [16]  count 1 (TMA tx)
    uint64_t* b_empty = bars + 48;       // [16]  count 1 (tcgen05.commit)
    uint64_t* tmem_full = bars + 64;     // [2]   count 1
    uint64_t* tmem_empty = bars + 66;    // [2]   count 4 (epilogue warps)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 68);
    volatile int* ticket_s = reinterpret_cast<volatile int*>(tmem_slot + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid >= 64 && tid < 64 + 68) {
        const int i = tid - 64;
        uint32_t cnt = 1;
        if (i >= 8 && i < 16) cnt = 8u;
        if (i >= 16 && i < 24) cnt = 64u;
        if (i >= 66) cnt = 4u;
        mbar_init(&bars[i], cnt);
        mbar_fence_init();
    }
    if (warp == 0 && lane == 0) tma_prefetch_desc(&tm_q);
    if (warp == 14 && lane == 0) tma_prefetch_desc(&tm_codes);
    if (warp == 1) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = *tmem_slot;

    if (warp == 0) {
        // ---- TMA producer: f16 token tiles (B operand), 64 k = 128 bytes per row
        if (lane == 0) {
            Ring rb((uint32_t)p.n_b);
            for (int it = blockIdx.x; it < p.n_items; it += gridDim.x) {
                const ItemH w = item_h(p, it);
                for (int kb = w.kb0; kb < w.kb1; ++kb) {
                    const uint32_t s = rb.stage();
                    mbar_wait(&b_empty[s], rb.parity() ^ 1u);
                    mbar_expect_tx(&b_full[s], b_stage_bytes);
                    tma_load_2d(b_base + (size_t)s * b_stage_bytes, &tm_q, kb * 128, w.m0, &b_full[s]);
                    rb.next();
                }
            }
        }
    } else if (warp == 14) {
        // ---- TMA producer: packed codes, 64 B x 128 rows = four k-blocks per stage
        if (lane == 0) {
            Ring pk((uint32_t)p.n_pk);
            for (int it = blockIdx.x; it < p.n_items; it += gridDim.x) {
                const ItemH w = item_h(p, it);
                for (int kb = w.kb0; kb < w.kb1; kb += H_PKQ) {
                    const uint32_t s = pk.stage();
                    mbar_wait(&pk_empty[s], pk.parity() ^ 1u);
                    mbar_expect_tx(&pk_full[s], (uint32_t)H_PK_BYTES);
                    tma_load_2d(pk_base + (size_t)s * H_PK_BYTES, &tm_codes, kb * (H_BKE / 4), w.row0, &pk_full[s]);
                    pk.next();
                }
            }
        }
    } else if (warp == 1) {
        // ---- MMA issuer (warp-uniform loop, one elected lane issues)
        const uint32_t idesc = umma_idesc_f16(H_ROWS, p.BN);
        const uint64_t adesc0 = umma_desc_sw128(smem_u32(base));
        const uint64_t bdesc0 = umma_desc_sw128(smem_u32(b_base));
        const uint32_t b_step = b_stage_bytes >> 4;
        const uint32_t n_a = (uint32_t)p.n_a, n_b = (uint32_t)p.n_b;
        uint32_t sa = 0, pa = 0, sb = 0, pb = 0, acc = 0, acc_ph = 0;
        for (int it = blockIdx.x; it < p.n_items; it += gridDim.x) {
            const ItemH w = item_h(p, it);
            mbar_wait(&tmem_empty[acc], acc_ph ^ 1u);
            tc_fence_after();
            const uint32_t d_addr = tmem_d + acc * 256u;
            uint32_t accumulate = 0;
            for (int kb = w.kb0; kb < w.kb1; ++kb) {
                mbar_wait(&b_full[sb], pb);
                mbar_wait(&a_full[sa], pa);
                tc_fence_after();
                const uint64_t adesc = adesc0 + (uint64_t)(sa * (uint32_t)(H_A_BYTES >> 4));
                const uint64_t bdesc = bdesc0 + (uint64_t)(sb * b_step);
                if (elect_one()) {
                    umma_f16(d_addr, adesc, bdesc, idesc, accumulate);  // K = 16 f16 = 32 bytes per MMA
                    umma_f16(d_addr, adesc + 2, bdesc + 2, idesc, 1u);
                    umma_f16(d_addr, adesc + 4, bdesc + 4, idesc, 1u);
                    umma_f16(d_addr, adesc + 6, bdesc + 6, idesc, 1u);
                    umma_commit(&a_empty[sa]);
                    umma_commit(&b_empty[sb]);
                    if (kb == w.kb1 - 1) umma_commit(&tmem_full[acc]);
                }
                __syncwarp();
                accumulate = 1u;
                if (++sa == n_a) { sa = 0; pa ^= 1u; }
                if (++sb == n_b) { sb = 0; pb ^= 1u; }
            }
            acc ^= 1u;
            if (acc == 0) acc_ph ^= 1u;
        }
    } else if (warp >= H_DEC_WARP0 && warp < H_DEC_WARP0 + 8) {
        // ---- decoders: codes -> f16 s * LUT[code] into the K-major SWIZZLE_128B A tile.
        // Group g (two warps) expands k-blocks g, g+4, ...; a thread owns four (row, 32-k half) units.
        const int dw = warp - H_DEC_WARP0;
        const int grp = dw >> 1, gw = dw & 1;
        const uint32_t half = (uint32_t)(lane >> 4);
        const int nblk32 = p.K >> 5;
        const __half2 lutA = __halves2half2(__int2half_rn((int)(int8_t)(p.lut & 0xFF)), __int2half_rn((int)(int8_t)((p.lut >> 8) & 0xFF)));
        const __half2 lutB = __halves2half2(__int2half_rn((int)(int8_t)((p.lut >> 16) & 0xFF)), __int2half_rn((int)(int8_t)(p.lut >> 24)));
        const uint32_t LSEL = 0x76543210u;  // byte c = selector nibbles (2c, 2c+1): the two bytes of f16 entry c
        uint32_t rloc[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) rloc[u] = (uint32_t)(gw * 64 + u * 16 + (lane & 15));
        Ring ra((uint32_t)p.n_a), pk((uint32_t)p.n_pk);
        for (int it = blockIdx.x; it < p.n_items; it += gridDim.x) {
            const ItemH w = item_h(p, it);
            const int nk = w.kb1 - w.kb0, nk4 = (nk + 3) & ~3;  // a packed stage always gets all its arrivals
            for (int rel = 0; rel < nk4; ++rel) {
                const uint32_t ps = pk.stage();
                if ((rel & 3) == grp) {
                    const bool real = rel < nk;
                    const int kb = w.kb0 + rel;
                    float sc[4];
                    uint2 cw[4];
                    if (real) {
#pragma unroll
                        for (int u = 0; u < 4; ++u) {  // block scales first: their L2 latency overlaps the barrier wait
                            const int n = w.row0 + (int)rloc[u];
                            sc[u] = n < p.N ? __ldg(p.wscale + (size_t)n * nblk32 + 2 * kb + (int)half) : 0.f;
                        }
                        mbar_wait(&pk_full[ps], pk.parity());
                        const unsigned char* pks = pk_base + (size_t)ps * H_PK_BYTES;
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const uint32_t off = rloc[u] * 64u;  // SWIZZLE_64B: 16-byte chunk ^= (byte offset >> 7) & 3
                            cw[u] = *reinterpret_cast<const uint2*>(pks + off + ((((uint32_t)rel & 3u) ^ ((off >> 7) & 3u)) << 4) + half * 8u);
                        }
                        const uint32_t sa = ra.stage();
                        mbar_wait(&a_empty[sa], ra.parity() ^ 1u);
                        unsigned char* a_tile = base + (size_t)sa * H_A_BYTES;
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const __half2 sv = __float2half2_rn(sc[u]);
                            const __half2 ha = __hmul2(sv, lutA), hb = __hmul2(sv, lutB);
                            const uint32_t pa = *reinterpret_cast<const uint32_t*>(&ha), pb = *reinterpret_cast<const uint32_t*>(&hb);
                            const uint32_t rx = rloc[u] & 7u, row_off = rloc[u] * 128u;
                            const uint32_t wds[2] = {cw[u].x, cw[u].y};
#pragma unroll
                            for (int wi = 0; wi < 2; ++wi) {
                                const uint32_t x = wds[wi] & 0x33333333u;
                                const uint32_t y = __umulhi(wds[wi], 0x40000000u) & 0x33333333u;
                                const uint32_t s0 = prmt(LSEL, 0u, x), s1 = prmt(LSEL, 0u, y);
                                const uint32_t s2 = prmt(LSEL, 0u, x >> 16), s3 = prmt(LSEL, 0u, y >> 16);
                                const uint4 c0 = make_uint4(prmt(pa, pb, s0), prmt(pa, pb, s0 >> 16), prmt(pa, pb, s1), prmt(pa, pb, s1 >> 16));
                                const uint4 c1 = make_uint4(prmt(pa, pb, s2), prmt(pa, pb, s2 >> 16), prmt(pa, pb, s3), prmt(pa, pb, s3 >> 16));
                                const uint32_t j0 = half * 4u + (uint32_t)(2 * wi);
                                *reinterpret_cast<uint4*>(a_tile + row_off + (((j0) ^ rx) << 4)) = c0;
                                *reinterpret_cast<uint4*>(a_tile + row_off + (((j0 + 1u) ^ rx) << 4)) = c1;
                            }
                        }
                        fence_proxy_async_smem();
                        mbar_arrive(&a_full[sa]);
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&pk_empty[ps]);  // only after the codes have been consumed
                }
                if (rel < nk) ra.next();
                if ((rel & 3) == 3) pk.next();
            }
        }
    } else {
        // ---- epilogue (warps 2-5): TMEM (fp32) -> y
        const int qd = warp & 3, nl = qd * 32 + lane, etid = (warp - 2) * 32 + lane;
        uint32_t nacc = 0;
        for (int it = blockIdx.x; it < p.n_items; it += gridDim.x, ++nacc) {
            const ItemH w = item_h(p, it);
            const uint32_t acc = nacc & 1u;
            const int n = w.row0 + nl;
            const bool nv = n < p.N;
            const int mv = min(p.BN, p.M - w.m0);
            float slv[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) slv[c] = (32 * c + lane < mv) ? __ldg(p.act_scale + w.m0 + 32 * c + lane) : 0.f;
            mbar_wait(&tmem_full[acc], (nacc >> 1) & 1u);
            tc_fence_after();
            const uint32_t t_addr = tmem_d + ((uint32_t)(qd * 32) << 16) + acc * 256u;
            if (p.S == 1) {
#pragma unroll 1
                for (int c0 = 0; c0 < mv; c0 += 32) {
                    uint32_t r[32];
                    float sl = slv[0];
#pragma unroll
                    for (int c = 1; c < 8; ++c) sl = (c0 == 32 * c) ? slv[c] : sl;
                    tmem_ld_32x32(t_addr + (uint32_t)c0, r);
                    float* yp = p.y + (size_t)(w.m0 + c0) * p.N + n;
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float sj = __shfl_sync(0xffffffffu, sl, j);
                        if (nv && c0 + j < mv) yp[(size_t)j * p.N] = __fmul_rn(sj, __uint_as_float(r[j]));
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tmem_empty[acc]);
            } else {
                float* blk = p.scratch + (size_t)w.tile * ((size_t)p.BN * H_ROWS) + nl;
                for (int c0 = 0; c0 < mv; c0 += 32) {
                    uint32_t r[32];
                    tmem_ld_32x32(t_addr + (uint32_t)c0, r);
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (c0 + j < mv) atomicAdd(blk + (size_t)(c0 + j) * H_ROWS, __uint_as_float(r[j]));
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tmem_empty[acc]);
                __threadfence();
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (etid == 0) *ticket_s = atomicAdd(p.counters + w.tile, 1);
                asm volatile("bar.sync 1, 128;" ::: "memory");
                const int ticket = *ticket_s;
                if (ticket == p.S - 1) {
                    __threadfence();
                    for (int c0 = 0; c0 < mv; c0 += 32) {
                        float sl = slv[0];
#pragma unroll
                        for (int c = 1; c < 8; ++c) sl = (c0 == 32 * c) ? slv[c] : sl;
                        float v[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = (c0 + j < mv) ? __ldcg(blk + (size_t)(c0 + j) * H_ROWS) : 0.f;
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const float sj = __shfl_sync(0xffffffffu, sl, j);
                            if (c0 + j < mv) {
                                __stcg(blk + (size_t)(c0 + j) * H_ROWS, 0.f);
                                if (nv) p.y[(size_t)(w.m0 + c0 + j) * p.N + n] = __fmul_rn(sj, v[j]);
                            }
                        }
                    }
                    if (etid == 0) p.counters[w.tile] = 0;
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_d, 512);
    }
}

// ---- host side ------------------------------------------------------------------------------
bool gemm_f16_supported(const bitnet_b200_weight* w, int M) {
    static const bool off = getenv("BITNET_B200_NO_GEMM") != nullptr;
    if (off) return false;
    return w->layout == BITNET_B200_LAYOUT_BITNET32_F16 && M >= 9 && w->cols % H_BKE == 0 &&
           w->code_stride == w->cols / 4 && w->scales_f32 != nullptr && w->f16_scales_exact;
}

int launch_gemm_f16(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q, const float* act_scale,
                    int M, float* y) {
    if (!gemm_f16_supported(w, M)) return fail("linear_i8 GEMM (BitNet32-F16): shape not covered by the tcgen05 kernel");
    if (reinterpret_cast<uintptr_t>(q) & 3) return fail("linear_i8 GEMM (BitNet32-F16): q must be 4-byte aligned");
    if (!ctx->gemm_scratch || !ctx->gemm_counters) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(ctx->stream, &cs);
        if (cs != cudaStreamCaptureStatusNone)
            return fail("linear_i8 GEMM: run the M > 8 path once on this context before capturing it into a CUDA graph");
        static std::mutex mu;
        std::lock_guard<std::mutex> lk(mu);
        auto* c = const_cast<bitnet_b200_ctx*>(ctx);
        if (!c->gemm_counters) {
            B200_CUDA(cudaMalloc(&c->gemm_counters, H_MAX_TILE_COUNTERS * sizeof(int)));
            B200_CUDA(cudaMemsetAsync(c->gemm_counters, 0, H_MAX_TILE_COUNTERS * sizeof(int), ctx->stream));
        }
        if (!c->gemm_scratch) {
            B200_CUDA(cudaMalloc(&c->gemm_scratch, H_SCRATCH_BYTES));
            B200_CUDA(cudaMemsetAsync(c->gemm_scratch, 0, H_SCRATCH_BYTES, ctx->stream));
        }
    }
    GemmF16Params p{};
    p.act_scale = act_scale;
    p.wscale = w->scales_f32;
    p.y = y;
    p.scratch = reinterpret_cast<float*>(ctx->gemm_scratch);
    p.counters = ctx->gemm_counters;
    p.lut = w->lut_word;
    p.M = M;
    p.N = (int)w->rows;
    p.K = (int)w->cols;
    p.n_rt = (int)div_ceil(w->rows, H_ROWS);
    const int nkb = p.K / H_BKE;
    const int m16 = (int)(div_ceil((size_t)M, 16) * 16);
    int bestBN = 0, bestS = 1;
    double best = 1e30;
    static const int bn_c[] = {16, 32, 64, 128, 256};
    static const int s_c[] = {1, 2, 3, 4, 6, 8};
    for (int bi = 0; bi < 5; ++bi) {
        int BN = bn_c[bi] < m16 ? bn_c[bi] : m16;
        if (BN > 256) BN = 256;
        const int n_mt = (int)div_ceil((size_t)M, (size_t)BN);
        BN = (int)(div_ceil(div_ceil((size_t)M, (size_t)n_mt), 16) * 16);
        const long tiles = (long)p.n_rt * n_mt;
        if (tiles > H_MAX_TILE_COUNTERS) continue;
        for (int si = 0; si < 6; ++si) {
            const int S = s_c[si];
            if (S > 1 && (S > nkb / 4 || (size_t)tiles * BN * H_ROWS * 4 > H_SCRATCH_BYTES)) continue;
            const long items = tiles * S;
            const double rounds = (double)div_ceil((size_t)items, (size_t)ctx->num_sms);
            const double t_kb = 0.43 * BN / 256.0 > 0.30 ? 0.43 * BN / 256.0 : 0.30;
            const double t = rounds * (div_ceil((size_t)nkb, (size_t)S) * t_kb + BN * (S > 1 ? 0.04 : 0.008) + 1.0) +
                             (S > 1 ? 1.0 : 0.0);
            if (t < best) {
                best = t;
                bestBN = BN;
                bestS = S;
            }
        }
    }
    if (!bestBN) return fail("linear_i8 GEMM (BitNet32-F16): no tile shape for this problem");
    p.BN = bestBN;
    p.n_mt = (int)div_ceil((size_t)M, (size_t)p.BN);
    p.S = bestS;
    const int tiles = p.n_rt * p.n_mt;
    p.n_pk = p.BN > 128 ? 2 : 6;
    p.n_a = p.BN > 128 ? 5 : 6;
    const size_t fixed = (size_t)p.n_a * H_A_BYTES + (size_t)p.n_pk * H_PK_BYTES + 1024 + H_BAR_BYTES;
    p.n_b = (int)((227 * 1024 - fixed) / ((size_t)p.BN * 128));
    if (p.n_b > 16) p.n_b = 16;
    if (p.n_b < 2) return fail("linear_i8 GEMM (BitNet32-F16): shared memory budget");
    p.n_items = tiles * p.S;
    const size_t smem = fixed + (size_t)p.n_b * p.BN * 128;

    // int8 -> f16 activations (exact), stream-ordered scratch
    __half* qh = nullptr;
    const size_t nelem = (size_t)M * p.K;
    B200_CUDA(cudaMallocAsync(&qh, nelem * sizeof(__half), ctx->stream));
    q8_to_f16_kernel<<<(unsigned)div_ceil(nelem / 4, 256), 256, 0, ctx->stream>>>(q, qh, nelem / 4);
    B200_LAUNCH_CHECK("q8_to_f16_kernel");

    CUtensorMap tm_codes, tm_q;
    if (make_map_2d(&tm_codes, w->codes, w->code_stride, w->rows, w->code_stride, H_ROWS, 64)) return -1;
    if (make_map_2d(&tm_q, qh, (size_t)p.K * 2, (size_t)M, (size_t)p.K * 2, p.BN)) return -1;

    static bool attr_set[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 16 && !attr_set[dev]) {
        B200_CUDA(cudaFuncSetAttribute(gemm_f16_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set[dev] = true;
    }
    const int grid = p.n_items < ctx->num_sms ? p.n_items : ctx->num_sms;
    gemm_f16_tcgen05_kernel<<<grid, H_THREADS, smem, ctx->stream>>>(tm_codes, tm_q, p);
    B200_LAUNCH_CHECK("gemm_f16_tcgen05_kernel");
    B200_CUDA(cudaFreeAsync(qh, ctx->stream));
    return 0;
}

}  // namespace b200
