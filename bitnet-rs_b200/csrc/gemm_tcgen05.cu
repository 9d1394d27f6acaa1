// gemm_tcgen05.cu -- the batched / prefill regime of the ternary linear (M >= 9): a dense
// contraction on the 5th-generation tensor cores, tcgen05.mma kind::i8 with the int32 accumulator
// in tensor memory (TMEM), operands fed by TMA. QK256 layout.
//
// Replaces, for device-resident packed weights and int8 activations, what the reference does with
// an f32 candle GEMM over load-time-dequantised weights (crates/bitnet-models/src/formats/gguf/
// loader.rs:1522-1617 + crates/bitnet-transformer/src/lib.rs:585,825) and per-token-row
// gemv_qk256 loops (lib.rs:589-702); fills launch_i2s_matmul (crates/bitnet-kernels/src/cuda/
// quantized_matmul.rs:282-303) for the packed layout. Result: acc[m,n] = sum_k q[m,k]*LUT[code(n,k)]
// in int32 -- bit-identical to gemv_qk256 on the same integers -- and y = act_scale[m] * acc.
//
// Shape of the kernel (persistent, one CTA per SM, 15 warps with fixed roles):
//   * the WEIGHTS are the A operand (UMMA M = 128 weight rows per tile), the TOKENS the B operand
//     (UMMA N = BN = 16..256 tokens): one expansion of a 128 x 128 block of 2-bit codes feeds
//     128 x BN x 128 MACs, and the accumulator sits in TMEM as [weight row = lane][token = column],
//     so the epilogue's stores of y[m, n0 .. n0+31] are 128-byte coalesced per warp;
//   * warp 0  : TMA producer of the int8 token tiles  (2-D tensor map over q[M,K], box 128 B x BN rows,
//               SWIZZLE_128B; rows beyond M are zero-filled by the TMA unit -> ragged M costs nothing);
//               its own ring of up to 16 stages: this is where L2 latency is hidden;
//   * warp 14 : TMA producer of the packed 2-bit codes (box 64 B x 128 rows = 256 k of 128 rows, 8 KB,
//               SWIZZLE_64B), its own ring: in the small-M regime this is the HBM stream;
//   * warps 6-13: decoders, four groups of two warps, group g expands k-blocks g, g+4, ... so four
//               k-blocks are in flight. Thread = 4 x (weight row, 64-k half): one conflict-free 128-bit
//               shared load of packed codes, four PRMT LUT decodes (common.cuh), four 128-bit stores into
//               the canonical K-major SWIZZLE_128B operand tile each, fence.proxy.async, mbarrier arrive.
//               A packed stage is released only AFTER its codes have been consumed (an early release let
//               the next TMA box overwrite codes whose shared loads were still in flight);
//   * warp 1  : MMA issuer. The loop is warp-uniform (descriptors and barrier addresses in uniform
//               registers) and one elected lane issues four tcgen05.mma (K = 32 each) per 128-k block, then
//               tcgen05.commit -> mbarrier frees the A and B stages. Its instruction count per k-block is
//               the critical path of the CTA at small BN (measured: 0.3 us per k-block issue-bound vs
//               0.43 us MMA-bound at BN = 256 under the 1 kW power cap), so nothing else lives in it.
//               Two TMEM accumulator stages (2 x 256 columns): the epilogue of tile i overlaps tile i+1;
//   * warps 2-5: epilogue: act scales prefetched before the accumulator is ready, tcgen05.ld (32 lanes x
//               32 columns) -> act_scale * float(acc) -> y.
//   * few output tiles (small M): the host picks BN and a split-K factor S from a small cost model so
//     that all 148 SMs stream weights; the int32 partials meet in an L2-resident block through RED.ADD,
//     the last CTA of a tile (ticket counter) scales it into y and re-zeroes the block -- integer
//     addition, so the result is independent of arrival order and stays bit-exact.
#include <cstdio>
#include <mutex>

#include "tcgen05.cuh"

namespace b200 {

constexpr int G_ROWS = 128;               // weight rows per tile (UMMA M)
constexpr int G_BK = 128;                 // int8 k per block = one 128-byte swizzle row
constexpr int G_PKQ = 2;                  // k-blocks per packed-code stage (64 B of codes per row)
constexpr int G_NA_MAX = 8;               // decoded-weight (A operand) stages
constexpr int G_NB_MAX = 16;              // token (B operand) stages: as many as fit, L2 latency is hidden here
constexpr int G_A_BYTES = G_ROWS * G_BK;  // 16 KB decoded weights per stage
constexpr int G_PK_BYTES = G_ROWS * 64;   // 8 KB packed codes per stage
constexpr int G_THREADS = 15 * 32;
constexpr int G_DEC_WARP0 = 6, G_DEC_WARPS = 8, G_DEC_GROUPS = 4;  // 4 groups of 2 warps, one k-block each
constexpr int G_BAR_BYTES = 1024;
constexpr size_t G_SCRATCH_BYTES = 16u << 20;  // split-K accumulation blocks (ctx-owned, zero between launches)
constexpr int G_MAX_TILE_COUNTERS = 8192;

struct GemmParams {
    const float* act_scale;  // [M]
    float* y;                // [M, N]
    int32_t* acc_out;        // optional [M, N]
    float* h_out;            // EPI_SWIGLU: [M, N/2] = silu(row 2i) * row 2i+1 instead of y (gate/up rows interleaved)
    int32_t* scratch;        // split-K partials
    int* counters;           // split-K tickets, zero between launches
    uint32_t lut;
    int M, N, K;
    int BN;        // tokens per tile (multiple of 16, <= 256)
    int n_rt;      // weight-row tiles
    int n_mt;      // token tiles
    int S;         // split-K factor
    int n_pk;      // packed-code stages
    int n_b;       // token stages
    int n_a;       // decoded-weight stages
    int n_items;   // n_rt * n_mt * S
};

struct Item {
    int row0, m0, kb0, kb1, tile, split;
};
__device__ __forceinline__ Item item_of(const GemmParams& p, int it) {
    Item r;
    r.split = it % p.S;
    r.tile = it / p.S;
    const int rt = r.tile % p.n_rt, mt = r.tile / p.n_rt;
    const int nkb = p.K / G_BK;
    r.row0 = rt * G_ROWS;
    r.m0 = mt * p.BN;
    r.kb0 = (int)(((long long)nkb * r.split) / p.S);
    r.kb1 = (int)(((long long)nkb * (r.split + 1)) / p.S);
    return r;
}

// SWIGLU = the epilogue applies the SwiGLU glue (EPI_SWIGLU); a separate instantiation so that the plain kernel's code
// (and its scheduling, which the added path perturbed by 17 % when it shared one body) stays exactly as it was.
template <bool SWIGLU>
__global__ void __launch_bounds__(G_THREADS, 1)
gemm_i8_tcgen05_kernel(const __grid_constant__ CUtensorMap tm_codes, const __grid_constant__ CUtensorMap tm_q,
                       const GemmParams p) {
    extern __shared__ unsigned char gsm_raw[];
    // 1024-byte alignment: the swizzle pattern is a function of the absolute shared address
    unsigned char* base = gsm_raw + ((1024u - (smem_u32(gsm_raw) & 1023u)) & 1023u);  // offset arithmetic keeps the shared address space (LDS / STS, not generic LD / ST)
    const uint32_t b_stage_bytes = (uint32_t)p.BN * 128u;
    unsigned char* b_base = base + (size_t)p.n_a * G_A_BYTES;
    unsigned char* pk_base = b_base + (size_t)p.n_b * b_stage_bytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(pk_base + (size_t)p.n_pk * G_PK_BYTES);
    uint64_t* pk_full = bars;            // [8]   count 1 (TMA tx)
    uint64_t* pk_empty = bars + 8;       // [8]   count 4 (2 k-blocks x 2 decoder warps)
    uint64_t* a_full = bars + 16;        // [8]   count 2 (the two warps of the owning decoder group)
    uint64_t* a_empty = bars + 24;       // [8]   count 1 (tcgen05.commit)
    uint64_t* b_full = bars + 32;        // [16]  count 1 (TMA tx)
    uint64_t* b_empty = bars + 48;       // [16]  count 1 (tcgen05.commit)
    uint64_t* tmem_full = bars + 64;     // [2]   count 1
    uint64_t* tmem_empty = bars + 66;    // [2]   count 4 (epilogue warps)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 68);
    volatile int* ticket_s = reinterpret_cast<volatile int*>(tmem_slot + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid >= 64 && tid < 64 + 68) {  // one barrier per thread
        const int i = tid - 64;
        uint32_t cnt = 1;
        if (i >= 8 && i < 16) cnt = 4u;
        if (i >= 16 && i < 24) cnt = 64u;
        if (i >= 66) cnt = 4u;
        mbar_init(&bars[i], cnt);
        mbar_fence_init();
    }
    if (warp == 0 && lane == 0) tma_prefetch_desc(&tm_q);
    if (warp == 14 && lane == 0) tma_prefetch_desc(&tm_codes);
    if (warp == 1) tmem_alloc(tmem_slot, 512);  // two accumulator stages of 256 columns
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = *tmem_slot;

    if (warp == 0) {
        // ================= TMA producer: int8 token tiles (B operand) =================
        if (lane == 0) {
            Ring rb((uint32_t)p.n_b);
            for (int it = blockIdx.x; it < p.n_items; it += gridDim.x) {
                const Item w = item_of(p, it);
                for (int kb = w.kb0; kb < w.kb1; ++kb) {
                    const uint32_t s = rb.stage();
                    mbar_wait(&b_empty[s], rb.parity() ^ 1u);
                    mbar_expect_tx(&b_full[s], b_stage_bytes);
                    tma_load_2d(b_base + (size_t)s * b_stage_bytes, &tm_q, kb * G_BK, w.m0, &b_full[s]);
                    rb.next();
                }
            }
        }
    } else if (warp == 14) {
        // ================= TMA producer: packed 2-bit codes (the HBM stream) =================
        if (lane == 0) {
            Ring pk((uint32_t)p.n_pk);
            for (int it = blockIdx.x; it < p.n_items; it += gridDim.x) {
                const Item w = item_of(p, it);
                for (int kb = w.kb0; kb < w.kb1; kb += G_PKQ) {
                    const uint32_t s = pk.stage();
                    mbar_wait(&pk_empty[s], pk.parity() ^ 1u);
                    mbar_expect_tx(&pk_full[s], (uint32_t)G_PK_BYTES);
                    tma_load_2d(pk_base + (size_t)s * G_PK_BYTES, &tm_codes, kb * (G_BK / 4), w.row0, &pk_full[s]);  // 64 B x 128 rows
                    pk.next();
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        // The whole warp runs the (warp-uniform) loop so that descriptors and barrier addresses live in
        // uniform registers; one elected lane issues. The issuing thread's instruction count per k-block
        // is the critical path of the CTA at small BN, so nothing else happens in here.
        const uint32_t idesc = umma_idesc_i8(G_ROWS, p.BN);
        const uint64_t adesc0 = umma_desc_sw128(smem_u32(base));
        const uint64_t bdesc0 = umma_desc_sw128(smem_u32(b_base));
        const uint32_t b_step = b_stage_bytes >> 4;  // descriptor address field is in 16-byte units
        const uint32_t n_a = (uint32_t)p.n_a, n_b = (uint32_t)p.n_b;
        uint32_t sa = 0, pa = 0, sb = 0, pb = 0, acc = 0, acc_ph = 0;
        for (int it = blockIdx.x; it < p.n_items; it += gridDim.x) {
            const Item w = item_of(p, it);
            mbar_wait(&tmem_empty[acc], acc_ph ^ 1u);
            tc_fence_after();
            const uint32_t d_addr = tmem_d + acc * 256u;
            uint32_t accumulate = 0;
            for (int kb = w.kb0; kb < w.kb1; ++kb) {
                mbar_wait(&b_full[sb], pb);
                mbar_wait(&a_full[sa], pa);
                tc_fence_after();
                const uint64_t adesc = adesc0 + (uint64_t)(sa * (uint32_t)(G_A_BYTES >> 4));
                const uint64_t bdesc = bdesc0 + (uint64_t)(sb * b_step);
                if (elect_one()) {
                    umma_i8(d_addr, adesc, bdesc, idesc, accumulate);  // +32 bytes along K = +2 in the address field
                    umma_i8(d_addr, adesc + 2, bdesc + 2, idesc, 1u);
                    umma_i8(d_addr, adesc + 4, bdesc + 4, idesc, 1u);
                    umma_i8(d_addr, adesc + 6, bdesc + 6, idesc, 1u);
                    umma_commit(&a_empty[sa]);  // stages reusable once these MMAs have read them
                    umma_commit(&b_empty[sb]);
                    if (kb == w.kb1 - 1) umma_commit(&tmem_full[acc]);  // accumulator complete
                }
                __syncwarp();
                accumulate = 1u;
                if (++sa == n_a) { sa = 0; pa ^= 1u; }
                if (++sb == n_b) { sb = 0; pb ^= 1u; }
            }
            acc ^= 1u;
            if (acc == 0) acc_ph ^= 1u;
        }
    } else if (warp >= G_DEC_WARP0 && warp < G_DEC_WARP0 + 8) {
        // ================= decoders: packed codes -> int8 A tile (SWIZZLE_128B, K-major) =================
        // Four groups of two warps; group g expands the k-blocks with (index & 3) == g, so four k-blocks
        // are in flight and the load -> PRMT -> store -> proxy-fence chain of one hides behind the others.
        // A thread owns four (row, 64-k half) units: one 128-bit shared load of codes and four 128-bit
        // stores of int8 each; a quarter-warp always covers 8 consecutive rows -> conflict-free both ways.
        const int dw = warp - G_DEC_WARP0;
        const int grp = dw >> 1, gw = dw & 1;
        const uint32_t half = (uint32_t)(lane >> 4);
        uint32_t pk_off[4], a_off[4];  // a_off: byte offset of this unit's chunk 0 position (before XOR)
        uint32_t rx[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const uint32_t r = (uint32_t)(gw * 64 + u * 16 + (lane & 15));
            rx[u] = r & 7u;
            a_off[u] = r * 128u;
            pk_off[u] = r * 64u;  // packed stage: rows of 64 bytes, SWIZZLE_64B: chunk ^= (byte offset >> 7) & 3
        }
        Ring ra((uint32_t)p.n_a), pk((uint32_t)p.n_pk);
        for (int it = blockIdx.x; it < p.n_items; it += gridDim.x) {
            const Item w = item_of(p, it);
            const int nk = w.kb1 - w.kb0, nk2 = (nk + 1) & ~1;  // a packed stage always gets both its arrivals
            for (int rel = 0; rel < nk2; ++rel) {
                const uint32_t ps = pk.stage();
                if ((rel & (G_DEC_GROUPS - 1)) == grp) {
                    const bool real = rel < nk;
                    uint4 cw[4];
                    if (real) {
                        mbar_wait(&pk_full[ps], pk.parity());
                        const unsigned char* pks = pk_base + (size_t)ps * G_PK_BYTES;
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const uint32_t c = (uint32_t)((rel & 1) * 2) + half;  // 16-byte chunk of the 64-byte row
                            cw[u] = *reinterpret_cast<const uint4*>(pks + pk_off[u] + ((c ^ ((pk_off[u] >> 7) & 3u)) << 4));
                        }
                    }
                    if (real) {
                        const uint32_t sa = ra.stage();
                        mbar_wait(&a_empty[sa], ra.parity() ^ 1u);
                        unsigned char* a_tile = base + (size_t)sa * G_A_BYTES;
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const uint32_t ws[4] = {cw[u].x, cw[u].y, cw[u].z, cw[u].w};
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                *reinterpret_cast<uint4*>(a_tile + a_off[u] + (((half * 4 + (uint32_t)j) ^ rx[u]) << 4)) =
                                    decode16_fma(ws[j], p.lut);
                        }
                        fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core (async proxy)
                        mbar_arrive(&a_full[sa]);  // every writer arrives itself, after its own proxy fence
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&pk_empty[ps]);  // the codes of this stage are consumed
                }
                if (rel < nk) ra.next();
                if (rel & 1) pk.next();
            }
        }
    } else {
        // ================= epilogue (warps 2-5): TMEM -> registers -> y =================
        const int qd = warp & 3;                 // TMEM lane quarter this warp may read
        const int nl = qd * 32 + lane;           // weight row within the tile
        const int etid = (warp - 2) * 32 + lane; // 0..127
        uint32_t nacc = 0;
        for (int it = blockIdx.x; it < p.n_items; it += gridDim.x, ++nacc) {
            const Item w = item_of(p, it);
            const uint32_t acc = nacc & 1u;
            const int n = w.row0 + nl;
            const bool nv = n < p.N;
            const int mv = min(p.BN, p.M - w.m0);  // valid token columns of this tile
            // act scales of the tile's tokens: lane l holds token 32c + l of chunk c. Fetched BEFORE waiting
            // for the accumulator, so their L2 latency hides behind the main loop.
            float slv[8];
#pragma unroll
            for (int c = 0; c < 8; ++c)
                slv[c] = (32 * c + lane < mv) ? __ldg(p.act_scale + w.m0 + 32 * c + lane) : 0.f;
            mbar_wait(&tmem_full[acc], (nacc >> 1) & 1u);
            tc_fence_after();
            const uint32_t t_addr = tmem_d + ((uint32_t)(qd * 32) << 16) + acc * 256u;
            if (SWIGLU) {
                // SwiGLU epilogue (crates/bitnet-transformer/src/lib.rs:756-789): adjacent TMEM lanes hold the (gate, up)
                // rows of one FFN column; silu(gate) * up (common.cuh: silu_mul, the formula of every SwiGLU site) is
                // stored instead of y: HALF the bytes, and the following act-quant reads them once instead of a glue
                // pass reading y and writing h. The lane pair splits the tokens: the even lane (gate row) finishes the
                // even tokens, the odd lane (up row) the odd ones -- one exchange and one silu per lane per two tokens.
                // A loop of its own: sharing the c0 loop with the plain epilogue cost the plain path 17 % (scheduling).
                const size_t hN = (size_t)(p.N >> 1);
                const bool odd = lane & 1;
#pragma unroll 1
                for (int c0 = 0; c0 < mv; c0 += 32) {
                    uint32_t r[32];
                    float sl = slv[0];
#pragma unroll
                    for (int c = 1; c < 8; ++c) sl = (c0 == 32 * c) ? slv[c] : sl;
                    tmem_ld_32x32(t_addr + (uint32_t)c0, r);
                    float* hp = p.h_out + (size_t)(w.m0 + c0) * hN + (n >> 1);
#pragma unroll
                    for (int j = 0; j < 32; j += 2) {
                        const float v0 = __fmul_rn(__shfl_sync(0xffffffffu, sl, j), (float)(int)r[j]);
                        const float v1 = __fmul_rn(__shfl_sync(0xffffffffu, sl, j + 1), (float)(int)r[j + 1]);
                        const float got = __shfl_xor_sync(0xffffffffu, odd ? v0 : v1, 1);
                        const float hv = odd ? silu_mul(got, v1) : silu_mul(v0, got);
                        const int jj = j + (odd ? 1 : 0);
                        if (nv && c0 + jj < mv) hp[(size_t)jj * hN] = hv;
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tmem_empty[acc]);
            } else if (p.S == 1) {
#pragma unroll 1
                for (int c0 = 0; c0 < mv; c0 += 32) {
                    uint32_t r[32];
                    float sl = slv[0];  // scale of token c0 + lane; broadcast by shuffle below
#pragma unroll
                    for (int c = 1; c < 8; ++c) sl = (c0 == 32 * c) ? slv[c] : sl;
                    tmem_ld_32x32(t_addr + (uint32_t)c0, r);
                    float* yp = p.y + (size_t)(w.m0 + c0) * p.N + n;
                    if (c0 + 32 <= mv && !p.acc_out) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const float sj = __shfl_sync(0xffffffffu, sl, j);
                            if (nv) yp[(size_t)j * p.N] = __fmul_rn(sj, (float)(int)r[j]);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const float sj = __shfl_sync(0xffffffffu, sl, j);
                            if (nv && c0 + j < mv) {
                                yp[(size_t)j * p.N] = __fmul_rn(sj, (float)(int)r[j]);
                                if (p.acc_out) p.acc_out[(size_t)(w.m0 + c0 + j) * p.N + n] = (int)r[j];
                            }
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tmem_empty[acc]);
            } else {
                // split-K: every CTA of the tile adds its int32 partial into the tile's accumulation
                // block [token][128 rows] with fire-and-forget reductions (RED.ADD at the L2; integer,
                // hence order-independent and exact); the last arrival (ticket) turns it into y and
                // leaves the block and the ticket zeroed for the next launch
                int32_t* blk = p.scratch + (size_t)w.tile * ((size_t)p.BN * G_ROWS) + nl;
                for (int c0 = 0; c0 < mv; c0 += 32) {
                    uint32_t r[32];
                    tmem_ld_32x32(t_addr + (uint32_t)c0, r);
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (c0 + j < mv) atomicAdd(blk + (size_t)(c0 + j) * G_ROWS, (int)r[j]);
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
                        int v[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = (c0 + j < mv) ? __ldcg(blk + (size_t)(c0 + j) * G_ROWS) : 0;
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const float sj = __shfl_sync(0xffffffffu, sl, j);
                            if (c0 + j < mv) {
                                __stcg(blk + (size_t)(c0 + j) * G_ROWS, 0);
                                if (nv) {
                                    const size_t oi = (size_t)(w.m0 + c0 + j) * p.N + n;
                                    p.y[oi] = __fmul_rn(sj, (float)v[j]);
                                    if (p.acc_out) p.acc_out[oi] = v[j];
                                }
                            }
                        }
                    }
                    if (etid == 0) p.counters[w.tile] = 0;  // self-cleaning: next launch starts from zero
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");  // ticket_s is reused by the next item
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
int gemm_min_m = 9;  // test/bench knob: smallest M routed to the tcgen05 kernel

bool gemm_i8_supported(const bitnet_b200_weight* w, int M) {
    static const bool off = getenv("BITNET_B200_NO_GEMM") != nullptr;
    if (off) return false;
    return w->layout == BITNET_B200_LAYOUT_QK256 && M >= gemm_min_m && w->cols % G_BK == 0 &&
           w->code_stride == w->cols / 4 && div_ceil(w->rows, G_ROWS) * div_ceil((size_t)M, 256) <= G_MAX_TILE_COUNTERS;
}

int gemm_split_override = 0;  // test knob: force a split-K factor (0 = heuristic)

int launch_gemm_i8(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q, const float* act_scale,
                   int M, float* y, int32_t* acc_out, float* h_out) {
    if (!gemm_i8_supported(w, M)) return fail("linear_i8 GEMM: shape not covered by the tcgen05 kernel");
    if (h_out && ((w->rows & 1) || acc_out)) return fail("linear_i8 GEMM: the SwiGLU epilogue needs an even row count and no acc_out");
    if (reinterpret_cast<uintptr_t>(q) & 15) return fail("linear_i8 GEMM: q must be 16-byte aligned");
    if (!ctx->gemm_scratch || !ctx->gemm_counters) {
        // first GEMM on this context: allocate the split-K scratch (cudaMalloc is illegal under graph capture,
        // so a captured step must have run once eagerly before -- every harness here warms up first)
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(ctx->stream, &cs);
        if (cs != cudaStreamCaptureStatusNone)
            return fail("linear_i8 GEMM: run the M > 8 path once on this context before capturing it into a CUDA graph");
        static std::mutex mu;
        std::lock_guard<std::mutex> lk(mu);
        auto* c = const_cast<bitnet_b200_ctx*>(ctx);
        if (!c->gemm_counters) {
            B200_CUDA(cudaMalloc(&c->gemm_counters, G_MAX_TILE_COUNTERS * sizeof(int)));
            B200_CUDA(cudaMemsetAsync(c->gemm_counters, 0, G_MAX_TILE_COUNTERS * sizeof(int), ctx->stream));
        }
        if (!c->gemm_scratch) {
            B200_CUDA(cudaMalloc(&c->gemm_scratch, G_SCRATCH_BYTES));
            B200_CUDA(cudaMemsetAsync(c->gemm_scratch, 0, G_SCRATCH_BYTES, ctx->stream));
        }
    }
    GemmParams p{};
    p.act_scale = act_scale;
    p.y = y;
    p.acc_out = acc_out;
    p.h_out = h_out;
    p.scratch = ctx->gemm_scratch;
    p.counters = ctx->gemm_counters;
    p.lut = w->lut_word;
    p.M = M;
    p.N = (int)w->rows;
    p.K = (int)w->cols;
    p.n_rt = (int)div_ceil(w->rows, G_ROWS);
    const int nkb = p.K / G_BK;
    // Tile shape: tokens per tile BN and split-K factor S from a small cost model (us): rounds of
    // persistent CTAs x (k-blocks x max(decode, MMA) + epilogue); RED-based split-K pays per element.
    const int m16 = (int)(div_ceil((size_t)M, 16) * 16);
    int bestBN = 0, bestS = 1;
    double best = 1e30;
    static const int bn_c[] = {16, 32, 64, 128, 256};
    static const int s_c[] = {1, 2, 3, 4, 6, 8};
    for (int bi = 0; bi < 5; ++bi) {
        int BN = bn_c[bi] < m16 ? bn_c[bi] : m16;
        if (BN > 256) BN = 256;
        const int n_mt = (int)div_ceil((size_t)M, (size_t)BN);
        BN = (int)(div_ceil(div_ceil((size_t)M, (size_t)n_mt), 16) * 16);  // balance ragged token tiles
        const long tiles = (long)p.n_rt * n_mt;
        if (tiles > G_MAX_TILE_COUNTERS) continue;
        for (int si = 0; si < 6; ++si) {
            const int S = s_c[si];
            if (S > 1 && (h_out || S > nkb / 2 || (size_t)tiles * BN * G_ROWS * 4 > G_SCRATCH_BYTES)) continue;  // SwiGLU epilogue: single-pass tiles only
            const long items = tiles * S;
            const double rounds = (double)div_ceil((size_t)items, (size_t)ctx->num_sms);
            const double t_kb = 0.43 * BN / 256.0 > 0.30 ? 0.43 * BN / 256.0 : 0.30;  // measured: MMA-bound vs issue-bound
            const double t = rounds * (div_ceil((size_t)nkb, (size_t)S) * t_kb + BN * (S > 1 ? 0.04 : 0.008) + 1.0) +
                             (S > 1 ? 1.0 : 0.0);
            if (t < best) {
                best = t;
                bestBN = BN;
                bestS = S;
            }
        }
    }
    if (!bestBN) return fail("linear_i8 GEMM: no tile shape for this problem");
    if (gemm_split_override > 0 && gemm_split_override <= nkb && !h_out) bestS = gemm_split_override;
    p.BN = bestBN;
    p.n_mt = (int)div_ceil((size_t)M, (size_t)p.BN);
    p.S = bestS;
    const int tiles = p.n_rt * p.n_mt;
    // shared memory: A (decoded) ring + token ring + packed ring; the token ring hides L2 latency,
    // the packed ring HBM latency (small M), the A ring the decode chain
    p.n_pk = p.BN > 128 ? 2 : 8;
    p.n_a = p.BN > 128 ? 5 : 6;
    if (const char* e = getenv("BITNET_B200_GEMM_RINGS")) {  // tuning knob: "n_a,n_pk" (decoded-weight / packed-code stages)
        int na = 0, npk = 0;
        if (sscanf(e, "%d,%d", &na, &npk) == 2 && na >= 4 && na <= 8 && npk >= 2 && npk <= 8) p.n_a = na, p.n_pk = npk;
    }
    const size_t fixed = (size_t)p.n_a * G_A_BYTES + (size_t)p.n_pk * G_PK_BYTES + 1024 + G_BAR_BYTES;
    p.n_b = (int)((227 * 1024 - fixed) / ((size_t)p.BN * 128));
    if (p.n_b > G_NB_MAX) p.n_b = G_NB_MAX;
    if (p.n_b < 2) return fail("linear_i8 GEMM: shared memory budget");
    p.n_items = tiles * p.S;
    const size_t smem = fixed + (size_t)p.n_b * p.BN * 128;

    CUtensorMap tm_codes, tm_q;
    if (make_map_2d(&tm_codes, w->codes, w->code_stride, w->rows, w->code_stride, G_ROWS, 64)) return -1;
    if (make_map_2d(&tm_q, q, (size_t)p.K, (size_t)M, (size_t)p.K, p.BN)) return -1;

    static bool attr_set[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 16 && !attr_set[dev]) {
        B200_CUDA(cudaFuncSetAttribute(gemm_i8_tcgen05_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        B200_CUDA(cudaFuncSetAttribute(gemm_i8_tcgen05_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set[dev] = true;
    }
    const int grid = p.n_items < ctx->num_sms ? p.n_items : ctx->num_sms;
    if (h_out) gemm_i8_tcgen05_kernel<true><<<grid, G_THREADS, smem, ctx->stream>>>(tm_codes, tm_q, p);
    else gemm_i8_tcgen05_kernel<false><<<grid, G_THREADS, smem, ctx->stream>>>(tm_codes, tm_q, p);
    B200_LAUNCH_CHECK("gemm_i8_tcgen05_kernel");
    return 0;
}

// ---- the same GEMM for callers that must not allocate (no counterpart in the reference) ----------------------------
// bitnet_b200_linear_i8_static_scratch: the split-K scratch (ticket counters + int32 accumulation blocks) that
// launch_gemm_i8 otherwise takes from two cudaMalloc'ed buffers of the context lives in ONE file-scope device array of a
// fixed size. Every GPU that loads the library holds it until the process exits; every call on a GPU shares it, so calls
// that overlap on that GPU (two streams, two host threads) race on it.
constexpr size_t G_STATIC_COUNTER_BYTES = G_MAX_TILE_COUNTERS * sizeof(int);  // 32 KiB, at the start of the array
constexpr size_t G_STATIC_BLOCK_BYTES = 4u << 20;                              // 4 MiB of accumulation blocks after them
constexpr size_t G_STATIC_WORDS = (G_STATIC_COUNTER_BYTES + G_STATIC_BLOCK_BYTES) / 4;
static __device__ __align__(256) int32_t g_static_scratch[G_STATIC_WORDS];

// The name g_static_scratch is only a host-side stand-in in host code: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static int static_scratch_ptr(int device, int32_t** out) {
    static std::mutex mu;
    static int32_t* ptr[64] = {};
    if (device < 0 || device >= 64) return fail("linear_i8_static_scratch: device ordinal out of range");
    std::lock_guard<std::mutex> lk(mu);
    if (!ptr[device]) {
        int cur = -1;
        B200_CUDA(cudaGetDevice(&cur));
        if (cur != device) return fail("linear_i8_static_scratch: the context's device is not current");
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) return fail_cuda("cudaGetSymbolAddress(g_static_scratch)", e);
        ptr[device] = static_cast<int32_t*>(p);
    }
    *out = ptr[device];
    return 0;
}

// bytes of accumulation blocks the planner of launch_gemm_i8 may use for this shape: the largest over its tile shapes
static size_t static_scratch_need(const bitnet_b200_weight* w, int M) {
    const size_t n_rt = div_ceil(w->rows, G_ROWS);
    const int m16 = (int)(div_ceil((size_t)M, 16) * 16);
    static const int bn_c[] = {16, 32, 64, 128, 256};
    size_t need = 0;
    for (int bi = 0; bi < 5; ++bi) {
        int BN = bn_c[bi] < m16 ? bn_c[bi] : m16;
        if (BN > 256) BN = 256;
        const size_t n_mt = div_ceil((size_t)M, (size_t)BN);
        BN = (int)(div_ceil(div_ceil((size_t)M, n_mt), 16) * 16);
        const size_t b = n_rt * n_mt * (size_t)BN * G_ROWS * 4;
        if (b > need) need = b;
    }
    return need;
}

int launch_gemm_i8_static_scratch(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q,
                                  const float* act_scale, int M, float* y) {
    if (!gemm_i8_supported(w, M))
        return fail("linear_i8_static_scratch: QK256 weights with cols % 128 == 0 and M >= 9 only (the tcgen05 GEMM)");
    const size_t need = static_scratch_need(w, M);
    if (need > G_STATIC_BLOCK_BYTES)
        return fail("linear_i8_static_scratch: rows x M too large for the 4 MiB static scratch (needs " + std::to_string(need) +
                    " bytes: 512 bytes per 128-row tile and token); use bitnet_b200_linear_i8");
    int32_t* base = nullptr;
    if (static_scratch_ptr(ctx->device, &base)) return -1;
    // the array keeps what the last call left: the kernel re-zeroes what it used, but a call that was cut short (a
    // fault, a forced split on another shape) must not leak into this one -> clear the tickets and the blocks we may use
    B200_CUDA(cudaMemsetAsync(base, 0, G_STATIC_COUNTER_BYTES + need, ctx->stream));
    bitnet_b200_ctx view = *ctx;  // same stream / device; scratch pointers -> the static array, nothing to allocate
    view.gemm_counters = reinterpret_cast<int*>(base);
    view.gemm_scratch = base + G_STATIC_COUNTER_BYTES / 4;
    return launch_gemm_i8(&view, w, q, act_scale, M, y, nullptr, nullptr);
}

}  // namespace b200

extern "C" int bitnet_b200_linear_i8_static_scratch(bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q,
                                                    const float* act_scale, int M, float* y) {
    using namespace b200;
    if (!ctx || !w || !q || !act_scale || !y) return fail("linear_i8_static_scratch: NULL argument");
    if (M <= 0) return fail("linear_i8_static_scratch: M must be > 0");
    if (w->device != ctx->device) return fail("linear_i8_static_scratch: weight lives on another device");
    B200_CUDA(cudaSetDevice(ctx->device));
    return launch_gemm_i8_static_scratch(ctx, w, q, act_scale, M, y);
}
