// gemm_small.cu -- ternary linear for the batched-decode regime, QK256 layout, M = 9..64 tokens.
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
#include <map>
#include <mutex>
#include <tuple>

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
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
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
    __syncthreads();
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
        // the weights do not depend on the previous kernel: fill the ring with them before griddepcontrol.wait
        for (int i = 0; i < p.NS; ++i) issue_weights(i);
        pdl_wait();  // the activations are what the previous kernel wrote
        for (int i = 0; i < p.NS; ++i) issue_acts(i);
    }
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

// ---------------------------------------------------------------------------------------------------------------
struct SmallPlan {
    int NT, RW, KW, U, ksplit, NS, CK, npr, grid_x, threads;
    uint32_t stage_bytes, w_off, red_off, ring_off;
    size_t smem;
};

static int small_nt(int M) {
    const int nt = (M + 7) / 8;
    return nt <= 2 ? 2 : nt <= 3 ? 3 : nt <= 4 ? 4 : nt <= 6 ? 6 : 8;
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
        case 2: return small_occ_nt<2>(pl.threads, pl.smem);
        case 3: return small_occ_nt<3>(pl.threads, pl.smem);
        case 4: return small_occ_nt<4>(pl.threads, pl.smem);
        case 6: return small_occ_nt<6>(pl.threads, pl.smem);
        default: return small_occ_nt<8>(pl.threads, pl.smem);
    }
}

// Heuristic: estimated cycles per SM for a candidate, smaller = better. The terms are the three streams a CTA
// overlaps -- tensor pipe (2 clk per IMMA per SM), activations from L2, weights from HBM -- plus a fixed cost per chunk
// (barrier round trip) and per CTA; CTAs beyond one wave of the 148 SMs queue.
static double small_cost(const bitnet_b200_ctx* ctx, const SmallPlan& pl, int M, int N, int K) {
    const double ctas = (double)pl.grid_x * pl.ksplit;
    const int occ = small_occupancy(pl);
    const double slots = (double)ctx->num_sms * occ;
    const double waves = ceil(ctas / slots);
    const double imma = (double)pl.RW * pl.KW * pl.U * 4 * pl.NT * 2.0 * occ;  // co-resident CTAs share the pipe
    const double l2 = (double)M * pl.CK / 40.0 * occ;
    double chunk = imma > l2 ? imma : l2;
    const double lat = 350.0;  // un-overlapped barrier + issue latency per chunk
    if (chunk < lat) chunk = lat;
    const double fixed = 1500.0 + (pl.ksplit > 1 ? 1200.0 : 0.0) + 40.0 * pl.RW * 2 * pl.NT;
    const double hbm_floor = (double)N * K / 4 / (21.0 * ctx->num_sms);
    double body = waves * pl.npr * chunk;
    if (body < hbm_floor) body = hbm_floor;
    return body + waves * fixed;
}

static bool plan_small(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M, SmallPlan* best) {
    if (w->layout != BITNET_B200_LAYOUT_QK256) return false;
    if (M < 9 || M > 64) return false;
    if (w->cols % 256 != 0 || w->code_stride != w->cols / 4 || w->rows < 1) return false;
    if (const char* e = getenv("BITNET_B200_SMALLM_PLAN")) {  // tuning knob: "RW,KW,U,ksplit"
        SmallPlan pl{};
        if (sscanf(e, "%d,%d,%d,%d", &pl.RW, &pl.KW, &pl.U, &pl.ksplit) == 4 && small_derive(w, M, &pl)) {
            *best = pl;
            return true;
        }
        *best = SmallPlan{};
        best->NT = -1;  // forced plan not realisable: launch_gemm_small reports it instead of rerouting silently
        return false;
    }
    static std::mutex mu;
    static std::map<std::tuple<size_t, size_t, int>, SmallPlan> cache;  // (rows, cols, M) -> plan
    const auto key = std::make_tuple(w->rows, w->cols, M);
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find(key);
        if (it != cache.end()) {
            *best = it->second;
            return best->NT != 0;
        }
    }
    bool found = false;
    double bc = 0;
    for (int RW = 1; RW <= 2; ++RW)
        for (int KW = 4; KW * RW <= SMALL_MAX_CONS; ++KW)
            for (int U = 1; U <= 4; U *= 2)
                for (int ks = 1; ks <= 8; ++ks) {
                    SmallPlan pl{};
                    pl.RW = RW, pl.KW = KW, pl.U = U, pl.ksplit = ks;
                    if (!small_derive(w, M, &pl)) continue;
                    const double c = small_cost(ctx, pl, M, (int)w->rows, (int)w->cols);
                    if (!found || c < bc) {
                        found = true;
                        bc = c;
                        *best = pl;
                    }
                }
    if (!found) *best = SmallPlan{};
    std::lock_guard<std::mutex> lk(mu);
    cache[key] = *best;
    return found;
}

// largest M routed here by linear_i8 (above it the tcgen05 GEMM takes over). BITNET_B200_SMALLM_MAX is read per call
// so that tests and A/B runs can switch paths inside one process.
int gemm_small_max_m() {
    const char* e = getenv("BITNET_B200_SMALLM_MAX");
    return e ? atoi(e) : 64;
}

bool gemm_small_supported(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, int M) {
    if (M > gemm_small_max_m()) return false;
    SmallPlan pl;
    return plan_small(ctx, w, M, &pl);
}

template <int NT>
static int launch_small_nt(const bitnet_b200_ctx* ctx, const SmallPlan& pl, const SmallParams& p, bool pdl) {
    CUtensorMap tm_codes, tm_q;
    if (make_map_colblocks(&tm_codes, p.codes, (size_t)p.stride, (size_t)p.N, (size_t)p.stride, 32 * pl.RW, 64, pl.CK / 256))
        return -1;
    if (make_map_colblocks(&tm_q, p.q, (size_t)p.K, (size_t)p.M, (size_t)p.K, 8 * NT, 128, pl.CK / 128)) return -1;
    static bool attr_set[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 16 && !attr_set[dev]) {
        B200_CUDA(cudaFuncSetAttribute(gemm_small_kernel<NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
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
    B200_CUDA(cudaLaunchKernelEx(&cfg, gemm_small_kernel<NT>, tm_codes, tm_q, p));
    count_launch();
    return 0;
}

// 1 = shape not covered, 0 = ok, -1 = error (same contract as launch_gemv_mma)
int launch_gemm_small(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q, const float* act_scale,
                      int M, float* y, int32_t* acc_out, bool pdl) {
    SmallPlan pl{};
    if (M > gemm_small_max_m()) return 1;
    if (!plan_small(ctx, w, M, &pl))
        return pl.NT == -1 ? fail("linear_i8: BITNET_B200_SMALLM_PLAN is not realisable for this shape") : 1;
    if (reinterpret_cast<uintptr_t>(q) & 15) return 1;
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
    switch (pl.NT) {
        case 2: return launch_small_nt<2>(ctx, pl, p, pdl);
        case 3: return launch_small_nt<3>(ctx, pl, p, pdl);
        case 4: return launch_small_nt<4>(ctx, pl, p, pdl);
        case 6: return launch_small_nt<6>(ctx, pl, p, pdl);
        default: return launch_small_nt<8>(ctx, pl, p, pdl);
    }
}

}  // namespace b200
