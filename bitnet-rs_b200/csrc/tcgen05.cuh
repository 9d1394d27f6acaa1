// tcgen05.cuh -- shared pieces of the tcgen05 GEMM kernels (gemm_tcgen05.cu: kind::i8 for QK256,
// gemm_f16_tcgen05.cu: kind::f16 for BitNet32-F16): TMEM / UMMA / TMA PTX wrappers, shared-memory
// descriptors, the incremental ring, and the host-side tensor-map encoder.
#pragma once
#include <cuda.h>

#include <string>

#include "common.cuh"

namespace b200 {

// ---- tcgen05 / TMEM / TMA PTX -----------------------------------------------------------------
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
// 2-D tiled TMA load: box at (c0 = innermost byte coordinate, c1 = row) -> shared, completes on mbar
__device__ __forceinline__ void tma_load_2d(void* dst_smem, const CUtensorMap* tm, int c0, int c1, void* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst_smem)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
// 3-D tiled TMA load: box at (c0 = innermost byte coordinate, c1, c2)
__device__ __forceinline__ void tma_load_3d(void* dst_smem, const CUtensorMap* tm, int c0, int c1, int c2, void* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst_smem)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// true in exactly one lane of the (converged) warp
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tm) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(tm) : "memory");
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

// running ring position -> (stage, parity)
// (kept incremental: a runtime modulo per k-block costs ~30 dependent instructions, and the single
// MMA-issuing thread is the critical path of the whole CTA)
struct Ring {
    uint32_t s, ph, stages;
    __device__ __forceinline__ Ring(uint32_t n) : s(0), ph(0), stages(n) {}
    __device__ __forceinline__ uint32_t stage() const { return s; }
    __device__ __forceinline__ uint32_t parity() const { return ph; }
    __device__ __forceinline__ void next() {
        if (++s == stages) {
            s = 0;
            ph ^= 1u;
        }
    }
};


// ---- host side ------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess)
            return nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// uint8 matrix [rows, inner] with a row pitch in bytes; box = 128 bytes x box_rows, SWIZZLE_128B;
// out-of-range rows / bytes read as zero
static inline int make_map_2d(CUtensorMap* tm, const void* ptr, size_t inner, size_t rows, size_t pitch, int box_rows,
                       int box_inner = 128) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) return fail("linear_i8 GEMM: cuTensorMapEncodeTiled is not available from this driver");
    const cuuint64_t gdim[2] = {(cuuint64_t)inner, (cuuint64_t)rows};
    const cuuint64_t gstr[1] = {(cuuint64_t)pitch};
    const cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1u, 1u};
    const CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(ptr), gdim, gstr, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, box_inner == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail("linear_i8 GEMM: cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
    return 0;
}

// The same matrix seen as [col_blocks][rows][box_inner bytes]: the third coordinate walks along the row in steps of
// box_inner bytes (its stride is SMALLER than the row pitch), so that one box {box_inner, box_rows, nblk} brings nblk
// consecutive swizzled [box_rows x box_inner] tiles -- the shared-memory image of nblk 2-D boxes -- with ONE TMA op.
static inline int make_map_colblocks(CUtensorMap* tm, const void* ptr, size_t inner, size_t rows, size_t pitch, int box_rows,
                                     int box_inner, int nblk) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) return fail("linear_i8 GEMM: cuTensorMapEncodeTiled is not available from this driver");
    const cuuint64_t gdim[3] = {(cuuint64_t)box_inner, (cuuint64_t)rows, (cuuint64_t)div_ceil(inner, (size_t)box_inner)};
    const cuuint64_t gstr[2] = {(cuuint64_t)pitch, (cuuint64_t)box_inner};
    const cuuint32_t box[3] = {(cuuint32_t)box_inner, (cuuint32_t)box_rows, (cuuint32_t)nblk};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    const CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(ptr), gdim, gstr, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, box_inner == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail("linear_i8 GEMM: cuTensorMapEncodeTiled (3-D) failed (" + std::to_string((int)r) + ")");
    return 0;
}

}  // namespace b200
