// gemm_small.cu -- ternary linear for the batched-decode regime, QK256 layout, M = 1..64 tokens.
//
// Same contract as the other linear_i8 kernels (i2s_qk256.rs:346-372 per token; the reference loops gemv_qk256 over
// the tokens, quantized_linear.rs forward_qk256): int32 accumulation of LUT[code] * int8 activation over all of K,
// y = act_scale[m] * (float)acc. Integer sums are order-free, so the result is bit-identical to the GEMV and to the
// tcgen05 GEMM whatever the split below.
//
// Why a third kernel: the tcgen05 GEMM (gemm_tcgen05.cu) pays 10-25 us of fixed latency per call (TMEM allocation,
// a decode -> shared -> UMMA pipeline that needs many k blocks to fill, split-K fix-up through L2), which is the whole
// cost at M <= 64, and the tensor-pipe GEMV (gemv_mma.cu) stops at 8 tokens. Here the legacy tensor path
// (mma.sync m16n8k32 s8) is the right tool: no fixed cost, and at M <= 32 the op is bound by streaming the packed
// weights from HBM once.
//
// Shape of the kernel:
//   * a CTA owns R = 32 * RW weight rows and a contiguous K range (K split over the CTAs of a thread-block cluster
//     when there are too few row groups to fill 148 SMs); its consumer warps are a KW x RW grid: warp (kw, rw) owns
//     rows [32 rw, 32 rw + 32) (two m16 tiles) and, of every chunk of CK = 64 U KW k, the 64 U k at offset 64 U kw;
//   * the k index of an MMA is only a summation index, so lane t of a quad owns 32 CONSECUTIVE k of each 128-k
//     segment: two 32-bit words of codes per weight row (one LDS.64), each decoded with 2 LOP3 + 3 SHF + 4 PRMT into
//     the A fragments of two MMAs, and two 128-bit shared loads of int8 activations per 8 tokens, each feeding the B
//     fragments of the same two MMAs for both m16 tiles (4 IMMA per LDS.128);
//   * chunk after chunk streams into a ring of shared-memory stages with tiled TMA, TWO ops per chunk: the int8
//     activations as CK / 128 boxes of 128 bytes x 8 NT token rows (SWIZZLE_128B; rows >= M and k >= K arrive as zeros)
//     and the codes as CK / 256 boxes of 64 bytes x R rows (SWIZZLE_64B), each through a 3-D tensor map whose third
//     coordinate walks along the row (tcgen05.cuh: make_map_colblocks). Per-row 1-D bulk copies were tried first and
//     are op-bound (a chunk is ~100 copies of 128-1024 bytes at ~46 cycles of TMA service each: 3-4x slower overall),
//     and so, less, are one 2-D op per box. With the swizzle, lane t reading 16-byte chunks 2t and
//     2t + 1 of token row g is conflict-free (chunk ^ g: even rows hit even chunks, odd rows odd ones). full/empty
//     mbarriers per stage; one lane of warp 0 issues the boxes on top of its consumer share (a 17th warp would cap
//     the kernel at 96 registers); the code boxes of the first NS stages are issued BEFORE griddepcontrol.wait;
//   * int32 accumulators stay in registers over the whole K range (2 x NT x 4 per lane); the KW partial sums meet in
//     shared memory (the ring, once drained), and with a K split the CTAs of the cluster read each other's reduced
//     tiles through distributed shared memory, each finishing 1 / ksplit of the outputs. No global scratch, no atomics.
#include <algorithm>
#include <cstdio>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "common.cuh"
#include "tcgen05.cuh"

namespace b200 {

constexpr int SMALL_MAX_STAGES = 8;
constexpr int SMALL_MAX_CONS = 16;   // consumer warps
constexpr int SMALL_HDR = 128;       // full[8] + empty[8] mbarriers

struct SmallParams {
    const uint8_t* codes;    // [N, stride] device code plane
    const int8_t* q;         // [M, K]
    const float* act_scale;  // [M]
    float* y;                // [M, N]
    int32_t* acc_out;        // optional [M, N]
    uint32_t lut;
    int N, K, M, stride;
    int RW, KW, U, ksplit, NS, CK, npr;
    uint32_t stage_bytes, w_off, red_off, ring_off;
    // activation-resident mode (gemm_small_resident_kernel)
    int nseg, ngroups;
    uint32_t part_off, act_off;
};

__device__ __forceinline__ void mma_s8_small(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                             uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ int4 ld_dsmem_int4(const void* local_ptr, uint32_t rank) {
    uint32_t remote;
    int4 v;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local_ptr)), "r"(rank));
    asm volatile("ld.shared::cluster.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "r"(remote)
                 : "memory");
    return v;
}
__device__ __forceinline__ void cons_bar(int nthreads) { asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory"); }

// job = (rw, mt, nt): one m16 x n8 output fragment of the CTA tile; lane (g, t) holds rows g / g + 8, tokens 2t / 2t + 1
template <int NT>
__device__ __forceinline__ void small_epilogue(const SmallParams& p, int job, int4 v, int row0, int lane) {
    const int g = lane >> 2, t = lane & 3;
    const int rw = job / (2 * NT), rem = job - rw * 2 * NT, mt = rem / NT, nt = rem - mt * NT;
    const int rA = row0 + 32 * rw + 16 * mt + g, rB = rA + 8;
    const int m0 = nt * 8 + 2 * t, m1 = m0 + 1;
    const float s0 = m0 < p.M ? __ldg(p.act_scale + m0) : 0.f, s1 = m1 < p.M ? __ldg(p.act_scale + m1) : 0.f;
#define B200_SMALL_STORE(row, m, sc, acc)                        \
    if ((row) < p.N && (m) < p.M) {                              \
        const size_t oi = (size_t)(m)*p.N + (row);               \
        p.y[oi] = __fmul_rn((sc), (float)(acc));                 \
        if (p.acc_out) p.acc_out[oi] = (acc);                    \
    }
    B200_SMALL_STORE(rA, m0, s0, v.x)
    B200_SMALL_STORE(rA, m1, s1, v.y)
    B200_SMALL_STORE(rB, m0, s0, v.z)
    B200_SMALL_STORE(rB, m1, s1, v.w)
#undef B200_SMALL_STORE
}

template <int NT>
__global__ void __launch_bounds__(32 * SMALL_MAX_CONS)
gemm_small_kernel(const __grid_constant__ CUtensorMap tm_codes, const __grid_constant__ CUtensorMap tm_q, const SmallParams p) {
    extern __shared__ unsigned char smem_raw[];
    // 1024-byte alignment: the swizzle patterns are functions of the absolute shared address (same offset in every CTA
    // of a cluster, so the DSMEM addresses below still match)
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // offset arithmetic keeps the shared address space (LDS / STS, not generic LD / ST)
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* empty = full + SMALL_MAX_STAGES;
    int4* red = reinterpret_cast<int4*>(smem + p.red_off);
    unsigned char* ring = smem + p.ring_off;  // multiple of 1024
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ncons = p.RW * p.KW;
    const int R = 32 * p.RW;
    const int row0 = blockIdx.x * R;
    const uint32_t rank = p.ksplit > 1 ? cluster_rank() : 0u;
    const int jobs = p.RW * 2 * NT;
    constexpr uint32_t A_TILE = NT * 8 * 128;  // one 128-k segment of all token rows
    const uint32_t w_tile = (uint32_t)R * 64;    // 256 k of the CTA's rows

    if (tid == 0) {
        tma_prefetch_desc(&tm_codes);
        tma_prefetch_desc(&tm_q);
        for (int s = 0; s < p.NS; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], ncons);
        }
        mbar_fence_init();
    }
    pdl_launch_dependents();

    // ---- one lane of warp 0 doubles as the producer ---------------------------------------------------------------
    auto issue_weights = [&](int i) {
        const int s = i % p.NS;
        unsigned char* st = ring + (size_t)s * p.stage_bytes;
        mbar_expect_tx(&full[s], p.stage_bytes);
        const int kc = (int)((rank * p.npr + i) * p.CK);
        tma_load_3d(st + p.w_off, &tm_codes, 0, row0, kc / 256, &full[s]);  // CK / 256 boxes of R rows x 64 bytes
    };
    auto issue_acts = [&](int i) {
        const int s = i % p.NS;
        unsigned char* st = ring + (size_t)s * p.stage_bytes;
        const int kc = (int)((rank * p.npr + i) * p.CK);
        tma_load_3d(st, &tm_q, 0, 0, kc / 128, &full[s]);  // CK / 128 boxes of 8 NT rows x 128 bytes
    };
    if (tid == 0) {
        // the weights do not depend on the previous kernel: fill the ring with them before griddepcontrol.wait; the
        // first loads are in flight before the other warps have even passed the barrier below
        for (int i = 0; i < p.NS; ++i) issue_weights(i);
        pdl_wait();  // the activations are what the previous kernel wrote
        for (int i = 0; i < p.NS; ++i) issue_acts(i);
    }
    __syncthreads();  // barrier inits visible to every waiter
    {
        // ---- consumers ---------------------------------------------------------------------------------------
        const int g = lane >> 2, t = lane & 3;
        const int kw = warp % p.KW, rw = warp / p.KW;
        int acc[2][NT][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[mt][nt][i] = 0;
        // activations, SWIZZLE_128B: 16-byte chunk c of token row r sits at chunk c ^ (r & 7); lane t owns chunks 2t, 2t+1
        const uint32_t a_lane0 = (uint32_t)g * 128 + (uint32_t)(((2 * t) ^ g) << 4);
        const uint32_t a_lane1 = (uint32_t)g * 128 + (uint32_t)(((2 * t + 1) ^ g) << 4);
        // codes, SWIZZLE_64B: rows of 64 bytes (256 k), chunk c of row r at c ^ ((r >> 1) & 3); lane t owns the 8 bytes
        // at 8t of its segment's 32: chunk 2 (seg & 1) + (t >> 1). (r >> 1) & 3 = g >> 1 for all four rows of a lane.
        const uint32_t w_row = p.w_off + (uint32_t)(32 * rw + g) * 64 + 8 * (t & 1);
        const uint32_t w_sw0 = (uint32_t)((((t >> 1)) ^ (g >> 1)) << 4), w_sw1 = (uint32_t)(((2 + (t >> 1)) ^ (g >> 1)) << 4);
        const int k_end = p.K;
#pragma unroll 1
        for (int i = 0; i < p.npr; ++i) {
            const int s = i % p.NS;
            if (tid == 0 && i >= 1 && i - 1 + p.NS < p.npr) {  // refill the stage every warp has left
                mbar_wait(&empty[(i - 1) % p.NS], ((i - 1) / p.NS) & 1);
                issue_weights(i - 1 + p.NS);
                issue_acts(i - 1 + p.NS);
            }
            __syncwarp();
            mbar_wait(&full[s], (i / p.NS) & 1);
            const unsigned char* st = ring + (size_t)s * p.stage_bytes;
            const int kc = (int)((rank * p.npr + i) * p.CK);
#pragma unroll 1
            for (int u = 0; u < p.U; ++u) {
                const int seg = kw + p.KW * u;
                if (kc + seg * 128 >= k_end) break;  // ragged last chunk: nothing but zero-filled activations there
                const unsigned char* wp = st + w_row + (uint32_t)(seg >> 1) * w_tile + ((seg & 1) ? w_sw1 : w_sw0);
                const uint2 w0 = *reinterpret_cast<const uint2*>(wp);
                const uint2 w1 = *reinterpret_cast<const uint2*>(wp + 8 * 64);
                const uint2 w2 = *reinterpret_cast<const uint2*>(wp + 16 * 64);
                const uint2 w3 = *reinterpret_cast<const uint2*>(wp + 24 * 64);
                const unsigned char* ap = st + (uint32_t)seg * A_TILE;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const uint4 d0 = decode16(h ? w0.y : w0.x, p.lut);
                    const uint4 d1 = decode16(h ? w1.y : w1.x, p.lut);
                    const uint4 d2 = decode16(h ? w2.y : w2.x, p.lut);
                    const uint4 d3 = decode16(h ? w3.y : w3.x, p.lut);
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) {
                        const uint4 b = *reinterpret_cast<const uint4*>(ap + (h ? a_lane1 : a_lane0) + nt * 1024);
                        mma_s8_small(acc[0][nt], d0.x, d1.x, d0.y, d1.y, b.x, b.y);
                        mma_s8_small(acc[1][nt], d2.x, d3.x, d2.y, d3.y, b.x, b.y);
                        mma_s8_small(acc[0][nt], d0.z, d1.z, d0.w, d1.w, b.z, b.w);
                        mma_s8_small(acc[1][nt], d2.z, d3.z, d2.w, d3.w, b.z, b.w);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
        // ---- the KW partial sums of every fragment meet in the (drained) ring -----------------------------------
        cons_bar(32 * ncons);
        int4* part = reinterpret_cast<int4*>(ring);  // [job][kw][lane]
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
                part[((size_t)((rw * 2 + mt) * NT + nt) * p.KW + kw) * 32 + lane] =
                    make_int4(acc[mt][nt][0], acc[mt][nt][1], acc[mt][nt][2], acc[mt][nt][3]);
        cons_bar(32 * ncons);
        for (int j = warp; j < jobs; j += ncons) {
            int4 v = make_int4(0, 0, 0, 0);
            for (int k = 0; k < p.KW; ++k) {
                const int4 x = part[((size_t)j * p.KW + k) * 32 + lane];
                v.x += x.x;
                v.y += x.y;
                v.z += x.z;
                v.w += x.w;
            }
            if (p.ksplit > 1)
                red[j * 32 + lane] = v;
            else
                small_epilogue<NT>(p, j, v, row0, lane);
        }
    }
    if (p.ksplit > 1) {
        // ---- K split: every CTA of the cluster finishes the fragments j = rank (mod ksplit) -----------------------
        cluster_sync_all();
        {
            for (int j = warp; j < jobs; j += ncons) {
                if ((uint32_t)j % (uint32_t)p.ksplit != rank) continue;
                int4 v = make_int4(0, 0, 0, 0);
                for (uint32_t r = 0; r < (uint32_t)p.ksplit; ++r) {
                    const int4 x = ld_dsmem_int4(&red[j * 32 + lane], r);
                    v.x += x.x;
                    v.y += x.y;
                    v.z += x.z;
                    v.w += x.w;
                }
                small_epilogue<NT>(p, j, v, row0, lane);
            }
        }
        cluster_sync_all();  // nobody leaves while a peer may still read its shared memory
    }
}

// ---- activation-resident variant: many row groups per CTA ------------------------------------------------------------
// When all of the tokens' int8 activations fit in shared memory next to a weight ring (M x K <= ~128 KB: every M <= 8,
// M <= 16 up to K = 8192, M <= 32 up to K = 4096) and there are more row groups than SMs (the gate/up projections),
// a persistent CTA loads the activations ONCE and then only streams weights: one 3-D TMA op brings the full-K codes of a
// row group (R rows x K / 4 bytes, 20-64 KB) into a ring stage, the warps split K as above, and the KW partial sums of
// group i meet in a double-buffered shared block while the codes of groups i + 1 .. i + NS - 1 are already in flight.
// One barrier per group; no cluster, no K split across CTAs. L2 -> SM activation traffic drops from (row groups) x M x K
// to (CTAs) x M x K, and the per-CTA start-up (barrier init, first-byte latency) is paid once per SM, not per group.
template <int NT>
__global__ void __launch_bounds__(32 * SMALL_MAX_CONS)
gemm_small_resident_kernel(const __grid_constant__ CUtensorMap tm_codes, const __grid_constant__ CUtensorMap tm_q,
                           const SmallParams p) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // offset arithmetic keeps the shared address space (LDS / STS, not generic LD / ST)
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* empty = full + SMALL_MAX_STAGES;
    uint64_t* afull = empty + SMALL_MAX_STAGES;
    int4* part0 = reinterpret_cast<int4*>(smem + p.part_off);
    const unsigned char* acts = smem + p.act_off;
    unsigned char* ring = smem + p.ring_off;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ncons = p.RW * p.KW;
    const int R = 32 * p.RW;
    const int jobs = p.RW * 2 * NT;
    constexpr uint32_t A_TILE = NT * 8 * 128;
    const uint32_t w_tile = (uint32_t)R * 64;
    const int my_groups = (p.ngroups - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

    auto issue_group = [&](int i) {  // the codes of this CTA's i-th row group, all of K: one op
        const int s = i % p.NS;
        mbar_expect_tx(&full[s], p.stage_bytes);
        tma_load_3d(ring + (size_t)s * p.stage_bytes, &tm_codes, 0, ((int)blockIdx.x + i * (int)gridDim.x) * R, 0, &full[s]);
    };
    if (tid == 0) {
        tma_prefetch_desc(&tm_codes);
        tma_prefetch_desc(&tm_q);
        for (int s = 0; s < p.NS; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], ncons);
        }
        mbar_init(afull, 1);
        mbar_fence_init();
        for (int i = 0; i < p.NS && i < my_groups; ++i) issue_group(i);
        pdl_wait();  // the activations are what the previous kernel wrote
        mbar_expect_tx(afull, (uint32_t)p.nseg * A_TILE);
        tma_load_3d(smem + p.act_off, &tm_q, 0, 0, 0, afull);
    }
    pdl_launch_dependents();
    __syncthreads();

    const int g = lane >> 2, t = lane & 3;
    const int kw = warp % p.KW, rw = warp / p.KW;
    const uint32_t a_lane0 = (uint32_t)g * 128 + (uint32_t)(((2 * t) ^ g) << 4);
    const uint32_t a_lane1 = (uint32_t)g * 128 + (uint32_t)(((2 * t + 1) ^ g) << 4);
    const uint32_t w_row = (uint32_t)(32 * rw + g) * 64 + 8 * (t & 1);
    const uint32_t w_sw0 = (uint32_t)((((t >> 1)) ^ (g >> 1)) << 4), w_sw1 = (uint32_t)(((2 + (t >> 1)) ^ (g >> 1)) << 4);
    mbar_wait(afull, 0);
#pragma unroll 1
    for (int i = 0; i < my_groups; ++i) {
        const int s = i % p.NS;
        if (tid == 0 && i >= 1 && i - 1 + p.NS < my_groups) {  // refill the stage every warp has left
            mbar_wait(&empty[(i - 1) % p.NS], ((i - 1) / p.NS) & 1);
            issue_group(i - 1 + p.NS);
        }
        __syncwarp();
        int acc[2][NT][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[mt][nt][j] = 0;
        mbar_wait(&full[s], (i / p.NS) & 1);
        const unsigned char* st = ring + (size_t)s * p.stage_bytes;
        // software pipeline over the warp's segments: the code words of segment n + 1 are loaded before the MMAs of
        // segment n (3-4 warps per scheduler cannot hide an LDS -> decode -> IMMA chain on their own)
        auto w_ptr = [&](int seg) { return st + w_row + (uint32_t)(seg >> 1) * w_tile + ((seg & 1) ? w_sw1 : w_sw0); };
        uint2 w0 = make_uint2(0, 0), w1 = w0, w2 = w0, w3 = w0;
        if (kw < p.nseg) {
            const unsigned char* wp = w_ptr(kw);
            w0 = *reinterpret_cast<const uint2*>(wp);
            w1 = *reinterpret_cast<const uint2*>(wp + 8 * 64);
            w2 = *reinterpret_cast<const uint2*>(wp + 16 * 64);
            w3 = *reinterpret_cast<const uint2*>(wp + 24 * 64);
        }
#pragma unroll 1
        for (int seg = kw; seg < p.nseg; seg += p.KW) {
            const uint2 c0 = w0, c1 = w1, c2 = w2, c3 = w3;
            if (seg + p.KW < p.nseg) {
                const unsigned char* wp = w_ptr(seg + p.KW);
                w0 = *reinterpret_cast<const uint2*>(wp);
                w1 = *reinterpret_cast<const uint2*>(wp + 8 * 64);
                w2 = *reinterpret_cast<const uint2*>(wp + 16 * 64);
                w3 = *reinterpret_cast<const uint2*>(wp + 24 * 64);
            }
            const unsigned char* ap = acts + (uint32_t)seg * A_TILE;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const uint4 d0 = decode16(h ? c0.y : c0.x, p.lut);
                const uint4 d1 = decode16(h ? c1.y : c1.x, p.lut);
                const uint4 d2 = decode16(h ? c2.y : c2.x, p.lut);
                const uint4 d3 = decode16(h ? c3.y : c3.x, p.lut);
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const uint4 b = *reinterpret_cast<const uint4*>(ap + (h ? a_lane1 : a_lane0) + nt * 1024);
                    mma_s8_small(acc[0][nt], d0.x, d1.x, d0.y, d1.y, b.x, b.y);
                    mma_s8_small(acc[1][nt], d2.x, d3.x, d2.y, d3.y, b.x, b.y);
                    mma_s8_small(acc[0][nt], d0.z, d1.z, d0.w, d1.w, b.z, b.w);
                    mma_s8_small(acc[1][nt], d2.z, d3.z, d2.w, d3.w, b.z, b.w);
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
        // partial sums of group i -> block i & 1; the block is rewritten two groups later, i.e. after the next barrier,
        // which every warp only passes once it has finished reading this one
        int4* part = part0 + (size_t)(i & 1) * jobs * p.KW * 32;
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
                part[((size_t)((rw * 2 + mt) * NT + nt) * p.KW + kw) * 32 + lane] =
                    make_int4(acc[mt][nt][0], acc[mt][nt][1], acc[mt][nt][2], acc[mt][nt][3]);
        cons_bar(32 * ncons);
        const int row0 = ((int)blockIdx.x + i * (int)gridDim.x) * R;
        for (int j = warp; j < jobs; j += ncons) {
            int4 v = make_int4(0, 0, 0, 0);
            for (int k = 0; k < p.KW; ++k) {
                const int4 x = part[((size_t)j * p.KW + k) * 32 + lane];
                v.x += x.x;
                v.y += x.y;
                v.z += x.z;
                v.w += x.w;
            }
            small_epilogue<NT>(p, j, v, row0, lane);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
struct SmallPlan {
    int NT, RW, KW, U, ksplit, NS, CK, npr, grid_x, threads;
    uint32_t stage_bytes, w_off, red_off, ring_off;
    size_t smem;
    int resident, nseg, ngroups;  // activation-resident persistent variant
    uint32_t part_off, act_off;
};

static int small_nt(int M) {
    const int nt = (M + 7) / 8;
    return nt <= 1 ? 1 : nt <= 2 ? 2 : nt <= 3 ? 3 : nt <= 4 ? 4 : nt <= 6 ? 6 : 8;
}

// fill the derived fields of a (RW, KW, U, ksplit) choice; false = not realisable for this shape
static bool small_derive(const bitnet_b200_weight* w, int M, SmallPlan* pl) {
    const int K = (int)w->cols, N = (int)w->rows;
    if (pl->RW < 1 || pl->RW > 2 || pl->KW < 1 || pl->RW * pl->KW > SMALL_MAX_CONS || pl->U < 1 || pl->ksplit < 1 ||
        pl->ksplit > 8)
        return false;
    pl->NT = small_nt(M);
    pl->CK = 128 * pl->U * pl->KW;
    if (pl->CK % 256) return false;  // code boxes are 64 bytes = 256 k
    const int nchunks = (K + pl->CK - 1) / pl->CK;  // a ragged last chunk reads zero-filled activations
    if (nchunks % pl->ksplit) return false;
    pl->npr = nchunks / pl->ksplit;
    pl->w_off = (uint32_t)pl->CK * pl->NT * 8;                       // CK / 128 boxes of 8 NT rows x 128 bytes
    pl->stage_bytes = pl->w_off + (uint32_t)(pl->CK / 256) * 32 * pl->RW * 64;  // + CK / 256 boxes of R rows x 64 bytes
    const int jobs = pl->RW * 2 * pl->NT;
    pl->red_off = SMALL_HDR;
    pl->ring_off = (uint32_t)((SMALL_HDR + (pl->ksplit > 1 ? jobs * 512 : 0) + 1023) & ~1023);
    const size_t part = (size_t)jobs * pl->KW * 512;
    const size_t budget = 220 * 1024;
    if (pl->ring_off + part > budget) return false;
    int ns = (int)((budget - pl->ring_off) / pl->stage_bytes);
    if (ns > SMALL_MAX_STAGES) ns = SMALL_MAX_STAGES;
    if (ns > pl->npr) ns = pl->npr;
    if (ns < 1 || (ns < 2 && pl->npr > 1)) return false;
    pl->NS = ns;
    const size_t ringb = (size_t)ns * pl->stage_bytes;
    pl->smem = 1024 + pl->ring_off + (ringb > part ? ringb : part);  // + alignment slack
    pl->grid_x = (int)div_ceil((size_t)N, (size_t)32 * pl->RW);
    pl->threads = 32 * pl->RW * pl->KW;
    return true;
}

// the activation-resident variant of (RW, KW): false = does not fit / not worthwhile for this shape
static bool small_derive_resident(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M, SmallPlan* pl) {
    const int K = (int)w->cols, N = (int)w->rows;
    if (pl->RW < 1 || pl->RW > 2 || pl->KW < 1 || pl->RW * pl->KW > SMALL_MAX_CONS) return false;
    pl->resident = 1;
    pl->NT = small_nt(M);
    pl->ksplit = 1;
    pl->nseg = K / 128;
    if (pl->nseg > 256 || pl->nseg / 2 > 256) return false;  // TMA box extents
    pl->U = (pl->nseg + pl->KW - 1) / pl->KW;
    pl->CK = K;
    pl->npr = 1;
    pl->ngroups = (int)div_ceil((size_t)N, (size_t)32 * pl->RW);
    if (pl->ngroups <= ctx->num_sms) return false;  // one group per CTA: nothing to keep resident for
    const int jobs = pl->RW * 2 * pl->NT;
    pl->part_off = 256;
    pl->act_off = (uint32_t)((pl->part_off + 2 * jobs * pl->KW * 512 + 1023) & ~1023);
    pl->ring_off = pl->act_off + (uint32_t)K * pl->NT * 8;
    pl->stage_bytes = (uint32_t)(32 * pl->RW) * (uint32_t)(K / 4);
    pl->w_off = 0, pl->red_off = 0;
    const size_t budget = 224 * 1024;
    if (pl->ring_off + 2 * (size_t)pl->stage_bytes > budget) return false;
    int ns = (int)((budget - pl->ring_off) / pl->stage_bytes);
    if (ns > SMALL_MAX_STAGES) ns = SMALL_MAX_STAGES;
    pl->NS = ns;
    pl->smem = 1024 + pl->ring_off + (size_t)ns * pl->stage_bytes;
    pl->threads = 32 * pl->RW * pl->KW;
    pl->grid_x = pl->ngroups < ctx->num_sms ? pl->ngroups : ctx->num_sms;
    return true;
}

template <int NT>
static int small_occ_nt(int threads, size_t smem) {
    int n = 0;
    cudaFuncSetAttribute(gemm_small_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, gemm_small_kernel<NT>, threads, smem) != cudaSuccess) {
        cudaGetLastError();
        n = 1;
    }
    return n < 1 ? 1 : n;
}
static int small_occupancy(const SmallPlan& pl) {
    switch (pl.NT) {
        case 1: return small_occ_nt<1>(pl.threads, pl.smem);
        case 2: return small_occ_nt<2>(pl.threads, pl.smem);
        case 3: return small_occ_nt<3>(pl.threads, pl.smem);
        case 4: return small_occ_nt<4>(pl.threads, pl.smem);
        case 6: return small_occ_nt<6>(pl.threads, pl.smem);
        default: return small_occ_nt<8>(pl.threads, pl.smem);
    }
}

// Cost model, microseconds: a least-squares fit over ~2600 timed (shape, M, plan) points on a B200
// (tools/smallm_tune.py; shapes of the 2B-4T and 8B-class models, whole and tp = 2 shards, M = 16 / 32 / 64). It ranks
// the candidates for the autotuner below and is the fallback where nothing can be timed (first use of a shape inside a
// CUDA graph capture); on the fitted set its pick is within 1.2x of the best plan on average.
static double small_cost(const bitnet_b200_ctx* ctx, const SmallPlan& pl) {
    const int occ = small_occupancy(pl);
    const double ctas = (double)pl.grid_x * pl.ksplit;
    const double rounds = ceil(ctas / ((double)ctx->num_sms * occ));
    double conc = ceil(ctas / ctx->num_sms);
    if (conc > occ) conc = occ;
    const double per_sm = rounds * conc;  // CTAs one SM executes, worst case
    const double imma = (double)pl.npr * pl.CK / 32 * 2 * pl.RW * pl.NT;
    const double wbytes = (double)pl.npr * pl.CK / 4 * 32 * pl.RW, abytes = (double)pl.npr * pl.CK * pl.NT * 8;
    const double t_imma = imma * 2.9 / 1900 * per_sm, t_bytes = (wbytes + abytes) / 42 / 1900 * per_sm,
                 t_w = wbytes / 23 / 1900 * per_sm;
    double t_max = t_imma > t_bytes ? t_imma : t_bytes;
    if (t_w > t_max) t_max = t_w;
    const int jobs = pl.RW * 2 * pl.NT;
    return 1.864 + 1.581 * rounds + 1.352 * rounds * (pl.ksplit > 1) + 2.334 * t_imma + 0.354 * t_bytes - 0.523 * t_w +
           0.466 * rounds * pl.npr / pl.NS - 1.686 * rounds * jobs * pl.KW / 256.0 - 0.175 * t_max;
}

static bool small_shape_ok(const bitnet_b200_weight* w, int M) {
    return w->layout == BITNET_B200_LAYOUT_QK256 && M >= 1 && M <= 64 && w->cols % 256 == 0 &&
           w->code_stride == w->cols / 4 && w->rows >= 1;
}

// every realisable plan of the search space, cheapest (by the model) first
static std::vector<std::pair<double, SmallPlan>> small_candidates(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w,
                                                                  int M) {
    std::vector<std::pair<double, SmallPlan>> v;
    for (int RW = 1; RW <= 2; ++RW)
        for (int KW = 2; KW * RW <= SMALL_MAX_CONS; ++KW)
            for (int U = 1; U <= 4; U *= 2)
                for (int ks = 1; ks <= 8; ++ks) {
                    SmallPlan pl{};
                    pl.RW = RW, pl.KW = KW, pl.U = U, pl.ksplit = ks;
                    if (!small_derive(w, M, &pl)) continue;
                    v.emplace_back(small_cost(ctx, pl), pl);
                }
    std::sort(v.begin(), v.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
    return v;
}

// largest M routed here by linear_i8 (above it the tcgen05 GEMM takes over). BITNET_B200_SMALLM_MAX is read per call
// so that tests and A/B runs can switch paths inside one process.
int gemm_small_max_m() {
    const char* e = getenv("BITNET_B200_SMALLM_MAX");
    return e ? atoi(e) : 64;
}

bool gemm_small_supported(const bitnet_b200_ctx*, const bitnet_b200_weight* w, int M) {
    return M <= gemm_small_max_m() && small_shape_ok(w, M);
}

template <int NT>
static int launch_small_nt(const bitnet_b200_ctx* ctx, const SmallPlan& pl, const SmallParams& p, bool pdl) {
    CUtensorMap tm_codes, tm_q;
    const int w_blk = pl.resident ? p.K / 256 : pl.CK / 256, a_blk = pl.resident ? p.K / 128 : pl.CK / 128;
    if (make_map_colblocks(&tm_codes, p.codes, (size_t)p.stride, (size_t)p.N, (size_t)p.stride, 32 * pl.RW, 64, w_blk)) return -1;
    if (make_map_colblocks(&tm_q, p.q, (size_t)p.K, (size_t)p.M, (size_t)p.K, 8 * NT, 128, a_blk)) return -1;
    static bool attr_set[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 16 && !attr_set[dev]) {
        B200_CUDA(cudaFuncSetAttribute(gemm_small_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        B200_CUDA(cudaFuncSetAttribute(gemm_small_resident_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        attr_set[dev] = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)pl.grid_x, (unsigned)pl.ksplit);
    cfg.blockDim = dim3((unsigned)pl.threads);
    cfg.dynamicSmemBytes = pl.smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (pl.ksplit > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = 1;
        attr[na].val.clusterDim.y = (unsigned)pl.ksplit;
        attr[na].val.clusterDim.z = 1;
        ++na;
    }
    if (pdl) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    if (pl.resident)
        B200_CUDA(cudaLaunchKernelEx(&cfg, gemm_small_resident_kernel<NT>, tm_codes, tm_q, p));
    else
        B200_CUDA(cudaLaunchKernelEx(&cfg, gemm_small_kernel<NT>, tm_codes, tm_q, p));
    g_last_linear_kernel = pl.resident ? "gemm_small_resident" : "gemm_small";
    count_launch();
    return 0;
}

static int run_small_plan(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const SmallPlan& pl, const int8_t* q,
                          const float* act_scale, int M, float* y, int32_t* acc_out, bool pdl) {
    SmallParams p{};
    p.codes = w->codes;
    p.q = q;
    p.act_scale = act_scale;
    p.y = y;
    p.acc_out = acc_out;
    p.lut = w->lut_word;
    p.N = (int)w->rows;
    p.K = (int)w->cols;
    p.M = M;
    p.stride = (int)w->code_stride;
    p.RW = pl.RW, p.KW = pl.KW, p.U = pl.U, p.ksplit = pl.ksplit, p.NS = pl.NS, p.CK = pl.CK;
    p.npr = pl.npr;
    p.stage_bytes = pl.stage_bytes, p.w_off = pl.w_off, p.red_off = pl.red_off, p.ring_off = pl.ring_off;
    p.nseg = pl.nseg, p.ngroups = pl.ngroups, p.part_off = pl.part_off, p.act_off = pl.act_off;
    switch (pl.NT) {
        case 1: return launch_small_nt<1>(ctx, pl, p, pdl);
        case 2: return launch_small_nt<2>(ctx, pl, p, pdl);
        case 3: return launch_small_nt<3>(ctx, pl, p, pdl);
        case 4: return launch_small_nt<4>(ctx, pl, p, pdl);
        case 6: return launch_small_nt<6>(ctx, pl, p, pdl);
        default: return launch_small_nt<8>(ctx, pl, p, pdl);
    }
}

// ---- plan selection: timed once per (device, rows, cols, token-tile count) ------------------------------------------
// The first eager call on a shape times the model's best SMALL_TUNE_CANDS plans and the tcgen05 GEMM on the caller's own
// operands (every candidate writes the same integers, so the output is the call's result whichever ran last), weights
// evicted from L2 between runs (decode streams each layer's weights from HBM), and keeps the fastest: the routing
// between the two kernels falls out of the same measurement. Nothing can be timed inside a CUDA graph capture: a shape
// first seen there uses the model's pick and the M <= 32 routing rule, without caching it.
// BITNET_B200_SMALLM_AUTOTUNE=0 disables the timing (model only); BITNET_B200_SMALLM_VERBOSE=1 prints the choice.
constexpr int SMALL_TUNE_CANDS = 20;
struct SmallChoice {
    bool use_alt;  // the other kernel of this M range measured faster: tcgen05 GEMM (M >= 9) / GEMV (M <= 8)
    SmallPlan pl;
    float us_small, us_alt;
};
int launch_gemv_fast(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M, int pro, const int8_t* q,
                     const float* act_scale, const float* x, const float* gamma, float eps, int epi, float* y,
                     const float* residual, int32_t* acc_out, int8_t* q_out, float* scale_out, float* normed_out, bool pdl);
bool gemv_fast_supported(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M, bool f32_prologue);
bool gemm_i8_supported(const bitnet_b200_weight* w, int M);
int launch_gemm_i8(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q, const float* act_scale, int M,
                   float* y, int32_t* acc_out, float* h_out);

static bool small_autotune(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q, const float* act_scale,
                           int M, float* y, int32_t* acc_out, SmallChoice* out) {
    auto cands = small_candidates(ctx, w, M);
    if (cands.empty()) return false;
    if ((int)cands.size() > SMALL_TUNE_CANDS) cands.resize(SMALL_TUNE_CANDS);
    static const int res_grid[6][2] = {{1, 8}, {1, 12}, {1, 16}, {2, 4}, {2, 6}, {2, 8}};  // (RW, KW) of the resident variant
    for (const auto& rg : res_grid) {
        SmallPlan pl{};
        pl.RW = rg[0], pl.KW = rg[1];
        if (small_derive_resident(ctx, w, M, &pl)) cands.emplace_back(0.0, pl);
    }
    // L2 flush buffer (> L2): allocated for the duration of this one tuning pass only; without it (allocation failed) the
    // candidates are timed L2-warm, which still ranks them
    constexpr size_t FLUSH_BYTES = 256u << 20;
    void* flush = nullptr;
    if (cudaMalloc(&flush, FLUSH_BYTES) != cudaSuccess) {
        cudaGetLastError();
        flush = nullptr;
    }
    cudaEvent_t e0, e1;
    if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) {
        if (flush) cudaFree(flush);
        return false;
    }
    auto timed = [&](auto&& launch) -> float {
        float best = 1e30f;
        for (int rep = 0; rep < 5; ++rep) {
            if (flush) cudaMemsetAsync(flush, rep & 1, FLUSH_BYTES, ctx->stream);
            cudaEventRecord(e0, ctx->stream);
            if (launch() != 0) {  // e.g. a cluster shape this device cannot co-schedule: not a candidate
                cudaGetLastError();
                return 1e30f;
            }
            cudaEventRecord(e1, ctx->stream);
            if (cudaEventSynchronize(e1) != cudaSuccess) {
                cudaGetLastError();
                return 1e30f;
            }
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            if (rep > 0 && ms < best) best = ms;
        }
        return best * 1e3f;
    };
    out->us_small = 1e30f;
    for (const auto& c : cands) {
        const float us = timed([&] { return run_small_plan(ctx, w, c.second, q, act_scale, M, y, acc_out, false); });
        if (us < out->us_small) {
            out->us_small = us;
            out->pl = c.second;
        }
    }
    out->us_alt = 1e30f;
    if (M >= 9 && gemm_i8_supported(w, M))
        out->us_alt = timed([&] { return launch_gemm_i8(ctx, w, q, act_scale, M, y, acc_out, nullptr); });
    if (M <= 8 && gemv_fast_supported(ctx, w, M, false))
        out->us_alt = timed([&] {
            return launch_gemv_fast(ctx, w, M, /*PRO_Q8*/ 0, q, act_scale, nullptr, nullptr, 0.f, /*EPI_STORE*/ 0, y, nullptr, acc_out,
                                    nullptr, nullptr, nullptr, false);
        });
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (flush) {
        cudaStreamSynchronize(ctx->stream);
        cudaFree(flush);
    }
    if (out->us_small >= 1e29f && out->us_alt >= 1e29f) return false;
    out->use_alt = out->us_alt < out->us_small;
    return true;
}

// 1 = shape not covered (or the tcgen05 GEMM measured faster: the caller falls through to it), 0 = ok, -1 = error
int launch_gemm_small(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q, const float* act_scale,
                      int M, float* y, int32_t* acc_out, bool pdl) {
    if (M > gemm_small_max_m() || !small_shape_ok(w, M)) return 1;
    if (reinterpret_cast<uintptr_t>(q) & 15) return 1;
    if (const char* e = getenv("BITNET_B200_SMALLM_PLAN")) {  // tuning / test knob: "RW,KW,U,ksplit"
        SmallPlan pl{};
        if (e[0] == 'R') {  // "R,RW,KW": the activation-resident variant
            if (sscanf(e + 1, ",%d,%d", &pl.RW, &pl.KW) != 2 || !small_derive_resident(ctx, w, M, &pl))
                return fail("linear_i8: BITNET_B200_SMALLM_PLAN (resident) is not realisable for this shape");
        } else if (sscanf(e, "%d,%d,%d,%d", &pl.RW, &pl.KW, &pl.U, &pl.ksplit) != 4 || !small_derive(w, M, &pl))
            return fail("linear_i8: BITNET_B200_SMALLM_PLAN is not realisable for this shape");
        return run_small_plan(ctx, w, pl, q, act_scale, M, y, acc_out, pdl);
    }
    static std::mutex mu;
    static std::map<std::tuple<int, size_t, size_t, int>, SmallChoice> chosen;
    const auto key = std::make_tuple(ctx->device, w->rows, w->cols, small_nt(M));
    SmallChoice ch{};
    bool have = false;
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = chosen.find(key);
        if (it != chosen.end()) {
            ch = it->second;
            have = true;
        }
    }
    if (!have) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        cudaStreamIsCapturing(ctx->stream, &cs);
        const char* at = getenv("BITNET_B200_SMALLM_AUTOTUNE");
        const bool tune = cs == cudaStreamCaptureStatusNone && !(at && at[0] == '0');
        if (tune) {
            std::lock_guard<std::mutex> lk(mu);  // one tuner at a time: the timings must not overlap each other
            auto it = chosen.find(key);
            if (it != chosen.end()) {
                ch = it->second;
                have = true;
            } else if (small_autotune(ctx, w, q, act_scale, M, y, acc_out, &ch)) {
                chosen[key] = ch;
                have = true;
                if (getenv("BITNET_B200_SMALLM_VERBOSE"))
                    fprintf(stderr, "[bitnet_b200] linear_i8 %zux%zu M=%d: %s (gemm_small%s RW=%d KW=%d U=%d ksplit=%d: %.1f us, %s: %.1f us)\n",
                            w->rows, w->cols, M, ch.use_alt ? "alternative" : "gemm_small", ch.pl.resident ? " resident" : "",
                            ch.pl.RW, ch.pl.KW, ch.pl.U, ch.pl.ksplit, ch.us_small, M >= 9 ? "tcgen05 gemm" : "gemv", ch.us_alt);
                // the call's result is already in y / acc_out (every candidate computed it)
                g_last_linear_kernel = "autotune";
                return 0;
            }
        }
        if (!have) {  // model only
            auto cands = small_candidates(ctx, w, M);
            if (cands.empty()) return 1;
            ch.pl = cands.front().second;
            ch.use_alt = (M > 32 && gemm_i8_supported(w, M)) || (M <= 8 && gemv_fast_supported(ctx, w, M, false));
        }
    }
    if (ch.use_alt) return 1;
    SmallPlan pl = ch.pl;  // same NT bucket: the derived fields do not depend on M beyond NT
    return run_small_plan(ctx, w, pl, q, act_scale, M, y, acc_out, pdl);
}

}  // namespace b200
