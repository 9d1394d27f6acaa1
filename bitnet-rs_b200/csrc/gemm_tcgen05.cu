// gemm_tcgen05.cu -- the batched / prefill regime of the ternary linear (M > 16): a dense
// contraction on the 5th-generation tensor cores, tcgen05.mma kind::i8 with the int32 accumulator
// in tensor memory (TMEM). QK256 layout.
//
// Replaces, for device-resident packed weights and int8 activations, what the reference does with
// an f32 candle GEMM over load-time-dequantised weights (crates/bitnet-models/src/formats/gguf/
// loader.rs:1522-1617 + crates/bitnet-transformer/src/lib.rs:585,825) and per-token-row
// gemv_qk256 loops (lib.rs:589-702); fills launch_i2s_matmul (crates/bitnet-kernels/src/cuda/
// quantized_matmul.rs:282-303) for the packed layout. Result: acc[m,n] = sum_k q[m,k]*LUT[code(n,k)]
// in int32 -- bit-identical to gemv_qk256 on the same integers -- and y = act_scale[m] * acc.
//
// Shape of the kernel (one CTA = one 128 x 128 output tile, 2 CTAs resident per SM):
//   * warps 0-3 (128 threads) are the operand producers. Thread t owns weight row n0+t and
//     activation row m0+t of the tile: per 128-wide K block it reads its row's 32 bytes of packed
//     codes, pushes them through the 4-entry int8 LUT (one PRMT per four weights, common.cuh)
//     and writes 128 int8 into shared memory in the canonical K-major SWIZZLE_128B layout that
//     the UMMA descriptor names; the int8 activation row goes next to it, same layout.
//     The 2-bit weights are expanded ONCE per CTA and reused by all 128 token rows of the tile.
//   * warp 4 is the MMA issuer: one elected lane waits for a stage (mbarrier), issues four
//     tcgen05.mma (K = 32 each) accumulating into TMEM, and releases the stage with
//     tcgen05.commit -> mbarrier. 3-stage ring, 32 KB per stage.
//   * the producer warps then become the epilogue: tcgen05.ld (32 lanes x 32 columns per warp
//     and instruction) -> act_scale * float(acc) -> 128-byte row segments of y.
#include "common.cuh"

namespace b200 {

constexpr int GEMM_BM = 128, GEMM_BN = 128, GEMM_BK = 128;  // int8 elements; BK = one 128-byte swizzle row
constexpr int GEMM_STAGES = 3;
constexpr int GEMM_TILE_BYTES = GEMM_BM * GEMM_BK;          // 16 KB per operand per stage
constexpr int GEMM_THREADS = 160;                           // 4 producer/epilogue warps + 1 MMA warp
constexpr int GEMM_SMEM = GEMM_STAGES * 2 * GEMM_TILE_BYTES + 1024 /*alignment slack*/ + 256 /*barriers*/;

struct GemmParams {
    const uint8_t* codes;   // [N, K/4] device layout
    const int8_t* q;        // [M, K]
    const float* act_scale; // [M]
    float* y;               // [M, N]
    int32_t* acc_out;       // optional [M, N]
    uint32_t lut;
    int M, N, K;
};

// ---- tcgen05 / TMEM PTX ---------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// D[tmem] (+)= A[smem] * B[smem], int8 x int8 -> int32, M x N x 32
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on an mbarrier once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(void* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 32 lanes (this warp's TMEM quarter) x 32 consecutive columns of 32-bit accumulators
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor of a K-major operand tile in the canonical SWIZZLE_128B layout
// (rows of 128 bytes, 8-row atoms of 1024 bytes, 16-byte chunk index XOR (row & 7)):
//   bits  0-13 start address >> 4          bits 16-29 leading byte offset >> 4 (1: unused in this mode)
//   bits 32-45 stride byte offset >> 4 (1024 B between 8-row atoms)   bits 46-47 version = 1 (sm_100)
//   bits 61-63 layout type = 2 (SWIZZLE_128B)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor, kind::i8: D = s32 (bits 4-5 = 2), A = B = signed int8 (bits 7-9, 10-12 = 1),
// both K-major (bits 15, 16 = 0), N >> 3 at bits 17-22, M >> 4 at bits 24-28
__device__ __forceinline__ uint32_t umma_idesc_i8(int m, int n) {
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// byte offset of 16-byte chunk c (0..7) of row r inside a [rows x 128 B] SWIZZLE_128B tile
__device__ __forceinline__ uint32_t sw128_off(int r, int c) {
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
}

__global__ void __launch_bounds__(GEMM_THREADS) gemm_i8_tcgen05_kernel(const GemmParams p) {
    extern __shared__ unsigned char gsm_raw[];
    // 1024-byte alignment: the swizzle pattern is a function of the absolute shared address
    unsigned char* tiles = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(gsm_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full = reinterpret_cast<uint64_t*>(tiles + GEMM_STAGES * 2 * GEMM_TILE_BYTES);  // [STAGES]
    uint64_t* empty = full + GEMM_STAGES;                                                       // [STAGES]
    uint64_t* tmem_full = empty + GEMM_STAGES;                                                  // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n0 = blockIdx.x * GEMM_BN, m0 = blockIdx.y * GEMM_BM;
    const int nkb = p.K / GEMM_BK;

    if (tid == 0) {
        for (int s = 0; s < GEMM_STAGES; ++s) {
            mbar_init(&full[s], 128);  // every producer thread arrives
            mbar_init(&empty[s], 1);   // one tcgen05.commit
        }
        mbar_init(tmem_full, 1);
        mbar_fence_init();
    }
    if (warp == 4) tmem_alloc(tmem_slot, GEMM_BN);  // 128 columns x 128 lanes of int32
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_d = *tmem_slot;

    if (warp < 4) {
        // ================= producers: thread t fills row t of the B (weights) and A (activations) tiles
        const int n = n0 + tid, m = m0 + tid;
        const bool nv = n < p.N, mv = m < p.M;
        const uint8_t* wrow = p.codes + (size_t)(nv ? n : 0) * (size_t)(p.K >> 2);
        const int8_t* arow = p.q + (size_t)(mv ? m : 0) * (size_t)p.K;
        for (int kb = 0; kb < nkb; ++kb) {
            const int s = kb % GEMM_STAGES;
            if (kb >= GEMM_STAGES) mbar_wait(&empty[s], (uint32_t)((kb / GEMM_STAGES) - 1) & 1u);
            unsigned char* a_tile = tiles + (size_t)s * 2 * GEMM_TILE_BYTES;
            unsigned char* b_tile = a_tile + GEMM_TILE_BYTES;
            // weights: 128 codes = 32 bytes -> 128 int8
            uint4 w0 = make_uint4(0, 0, 0, 0), w1 = w0;
            if (nv) {
                w0 = ldg_nc_u128(wrow + (size_t)kb * 32);
                w1 = ldg_nc_u128(wrow + (size_t)kb * 32 + 16);
            }
            uint4 av[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                av[c] = make_uint4(0, 0, 0, 0);
                if (mv) av[c] = ldg_nc_u128(arow + (size_t)kb * GEMM_BK + c * 16);
            }
            const uint32_t ws[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                uint4 d = make_uint4(0, 0, 0, 0);
                if (nv) d = decode16(ws[c], p.lut);
                *reinterpret_cast<uint4*>(b_tile + sw128_off(tid, c)) = d;
            }
#pragma unroll
            for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4*>(a_tile + sw128_off(tid, c)) = av[c];
            fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core (async proxy)
            mbar_arrive(&full[s]);
        }
        // ================= epilogue: TMEM -> registers -> y
        mbar_wait(tmem_full, 0);
        tc_fence_after();
        const float as = mv ? p.act_scale[m] : 0.f;
        for (int c0 = 0; c0 < GEMM_BN; c0 += 32) {
            uint32_t r[32];
            tmem_ld_32x32(tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
            if (mv) {
                float* yr = p.y + (size_t)m * p.N + n0 + c0;
                int32_t* ar = p.acc_out ? p.acc_out + (size_t)m * p.N + n0 + c0 : nullptr;
                if (n0 + c0 + 32 <= p.N && ((reinterpret_cast<uintptr_t>(yr) & 15) == 0)) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        *reinterpret_cast<float4*>(yr + j) =
                            make_float4(__fmul_rn(as, (float)(int)r[j]), __fmul_rn(as, (float)(int)r[j + 1]),
                                        __fmul_rn(as, (float)(int)r[j + 2]), __fmul_rn(as, (float)(int)r[j + 3]));
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (n0 + c0 + j < p.N) yr[j] = __fmul_rn(as, (float)(int)r[j]);
                }
                if (ar) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (n0 + c0 + j < p.N) ar[j] = (int)r[j];
                }
            }
        }
        tc_fence_before();
    } else {
        // ================= MMA issuer (warp 4): one elected lane
        const uint32_t idesc = umma_idesc_i8(GEMM_BM, GEMM_BN);
        for (int kb = 0; kb < nkb; ++kb) {
            const int s = kb % GEMM_STAGES;
            mbar_wait(&full[s], (uint32_t)(kb / GEMM_STAGES) & 1u);
            tc_fence_after();
            if (lane == 0) {
                const uint32_t a_addr = smem_u32(tiles + (size_t)s * 2 * GEMM_TILE_BYTES);
                const uint64_t adesc = umma_desc_sw128(a_addr), bdesc = umma_desc_sw128(a_addr + GEMM_TILE_BYTES);
#pragma unroll
                for (int k = 0; k < GEMM_BK / 32; ++k)  // +32 bytes along K = +2 in the (>>4) address field
                    umma_i8(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (uint32_t)(kb | k));
                umma_commit(&empty[s]);                       // stage reusable once these MMAs have read it
                if (kb == nkb - 1) umma_commit(tmem_full);    // accumulator complete
            }
            __syncwarp();
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 4) {
        tc_fence_after();
        tmem_dealloc(tmem_d, GEMM_BN);
    }
}

bool gemm_i8_supported(const bitnet_b200_weight* w, int M) {
    static const bool off = getenv("BITNET_B200_NO_GEMM") != nullptr;
    if (off) return false;
    return w->layout == BITNET_B200_LAYOUT_QK256 && M > 16 && w->cols % GEMM_BK == 0 && w->code_stride == w->cols / 4;
}

int launch_gemm_i8(const bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q, const float* act_scale,
                   int M, float* y, int32_t* acc_out) {
    if (!gemm_i8_supported(w, M)) return fail("linear_i8 GEMM: shape not covered by the tcgen05 kernel");
    if (reinterpret_cast<uintptr_t>(q) & 15) return fail("linear_i8 GEMM: q must be 16-byte aligned");
    GemmParams p{};
    p.codes = w->codes;
    p.q = q;
    p.act_scale = act_scale;
    p.y = y;
    p.acc_out = acc_out;
    p.lut = w->lut_word;
    p.M = M;
    p.N = (int)w->rows;
    p.K = (int)w->cols;
    static bool attr_set[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 16 && !attr_set[dev]) {
        B200_CUDA(cudaFuncSetAttribute(gemm_i8_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM));
        attr_set[dev] = true;
    }
    dim3 grid((unsigned)div_ceil((size_t)p.N, GEMM_BN), (unsigned)div_ceil((size_t)M, GEMM_BM));
    gemm_i8_tcgen05_kernel<<<grid, GEMM_THREADS, GEMM_SMEM, ctx->stream>>>(p);
    B200_LAUNCH_CHECK("gemm_i8_tcgen05_kernel");
    return 0;
}

}  // namespace b200
