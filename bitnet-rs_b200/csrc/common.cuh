// common.cuh -- shared host/device helpers for libbitnet_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stddef.h>

#include <atomic>
#include <string>

#include "../../include/bitnet_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libbitnet_b200 targets sm_100a only"
#endif

namespace b200 {

// ---- error convention: int rc + thread-local message (ffi/cpp_bridge.cpp:19,26-28) ----
void set_error(const std::string& msg);
void clear_error();
int fail(const std::string& msg);
int fail_cuda(const char* what, cudaError_t e);
extern std::atomic<uint64_t> g_launches;
inline void count_launch(uint64_t n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }
extern thread_local const char* g_last_linear_kernel;  // diagnostics: bitnet_b200_last_linear_kernel()

#define B200_CUDA(expr)                                                  \
    do {                                                                 \
        cudaError_t _e = (expr);                                         \
        if (_e != cudaSuccess) return ::b200::fail_cuda(#expr, _e);      \
    } while (0)

#define B200_LAUNCH_CHECK(name)                                          \
    do {                                                                 \
        cudaError_t _e = cudaGetLastError();                             \
        if (_e != cudaSuccess) return ::b200::fail_cuda(name, _e);       \
        ::b200::count_launch();                                          \
    } while (0)

inline size_t div_ceil(size_t a, size_t b) { return (a + b - 1) / b; }

}  // namespace b200

// ---- opaque handles of the C ABI ------------------------------------------------------
struct bitnet_b200_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    int num_sms = 148;
    int* workspace = nullptr;  // small device scratch (tile counters etc.)
    int32_t* gemm_scratch = nullptr;  // split-K int32 partials of the tcgen05 GEMM (gemm_tcgen05.cu)
    int* gemm_counters = nullptr;     // split-K ticket counters, zero between launches
};

struct bitnet_b200_weight {
    int device = 0;
    int layout = 0;
    size_t rows = 0, cols = 0;
    size_t code_stride = 0;     // bytes per row in the code plane (multiple of 16)
    size_t nblk32 = 0;          // cols/32 (BitNet32-F16) else 0
    size_t ref_nbytes = 0;      // bytes in the reference layout
    uint8_t* codes = nullptr;   // [rows, code_stride]
    __half* scales = nullptr;   // [rows, nblk32] effective scales as f16 bits (raw) -- BitNet32 only
    float* scales_f32 = nullptr;  // [rows, nblk32] policy-applied f32 scales (GEMM epilogue)
    uint32_t lut_word = 0;      // 4 int8 LUT entries, entry c in byte c
    int8_t lut[4] = {-2, -1, 1, 2};
    int scale_policy = 0;
    // BitNet32-F16: every policy-applied block scale is exactly representable in f16 and |s| * max|LUT| stays finite
    // in f16 -- the condition under which the tcgen05 kind::f16 GEMM (which folds s into f16 weights) computes the
    // same products as the f32-scale GEMV (gemm_f16_tcgen05.cu); otherwise M >= 9 loops the GEMV.
    bool f16_scales_exact = true;
};

// ---- device helpers ---------------------------------------------------------------------
#ifdef __CUDACC__
namespace b200 {

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}

// signed x signed 4-way int8 dot product with int32 accumulate
__device__ __forceinline__ int dp4a_ss(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// signed (a) x unsigned (b)
__device__ __forceinline__ int dp4a_su(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__device__ __forceinline__ uint32_t ldg_nc_u32(const void* p) {
    uint32_t v;
    asm volatile("ld.global.nc.L1::no_allocate.b32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint2 ldg_nc_u64(const void* p) {
    uint2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.b32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ uint4 ldg_nc_u128(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ uint16_t ldg_nc_u16(const void* p) {
    uint16_t v;
    asm volatile("ld.global.nc.L1::no_allocate.b16 %0, [%1];" : "=h"(v) : "l"(p));
    return v;
}

// ---- device code-plane layout -------------------------------------------------------------
// The reference packs code k of a 16-weight word at bits [2k, 2k+1] (LSB first,
// i2s_qk256.rs:159-168). At upload every 32-bit word is bit-permuted ONCE (lossless, see
// repack_word) so that the four PRMT selectors of a word come out in natural k order:
//   k = 0..3  -> bits 4k          k = 4..7   -> bits 4(k-4)+2
//   k = 8..11 -> bits 16+4(k-8)   k = 12..15 -> bits 18+4(k-12)
// i.e. (w & 0x3333) selects k=0..3, ((w>>2) & 0x3333) k=4..7, and the same on the high half.
__host__ __device__ __forceinline__ int code_bitpos(int k) {
    return ((k & 8) ? 16 : 0) + ((k & 4) ? 2 : 0) + 4 * (k & 3);
}
__host__ __device__ __forceinline__ uint32_t repack_word(uint32_t w) {
    uint32_t r = 0;
    for (int k = 0; k < 16; ++k) r |= ((w >> (2 * k)) & 3u) << code_bitpos(k);
    return r;
}
__device__ __forceinline__ uint32_t code_of(uint32_t w_dev, int k) { return (w_dev >> code_bitpos(k)) & 3u; }

// Decode one device-layout word (sixteen 2-bit codes) through a 4-entry int8 LUT -- one PRMT
// per four weights -- and dot it with sixteen int8 activations in natural order (a.x = k0..3).
// The shifts are plain SHF: IMAD.HI (umulhi), once used here to move them to the FMA pipe, measured SLOWER on
// B200 (tools/ubench/tile_loop.cu: 7.8 vs 6.2 cycles per 512-weight fragment per SM).
__device__ __forceinline__ int dot16(uint32_t w, uint32_t lut, uint4 a, int acc) {
    const uint32_t x = w & 0x33333333u;
    const uint32_t y = (w >> 2) & 0x33333333u;
    acc = dp4a_ss(prmt(lut, 0u, x), a.x, acc);
    acc = dp4a_ss(prmt(lut, 0u, y), a.y, acc);
    acc = dp4a_ss(prmt(lut, 0u, x >> 16), a.z, acc);
    acc = dp4a_ss(prmt(lut, 0u, y >> 16), a.w, acc);
    return acc;
}
// decode a device-layout word into sixteen int8 values in natural order (4 words)
__device__ __forceinline__ uint4 decode16(uint32_t w, uint32_t lut) {
    const uint32_t x = w & 0x33333333u;
    const uint32_t y = (w >> 2) & 0x33333333u;
    return make_uint4(prmt(lut, 0u, x), prmt(lut, 0u, y), prmt(lut, 0u, x >> 16), prmt(lut, 0u, y >> 16));
}

// The same decode with the three shifts issued as IMAD.HI on the FMA pipe. Slower where the decode feeds mma.sync
// directly (above), but the tcgen05 GEMM's decoder warps, which also issue the shared-memory stores of the expanded
// tile, measured 12 % faster with it (gate/up at M = 16384: 0.56 ms vs 0.71 ms with SHF).
__device__ __forceinline__ uint4 decode16_fma(uint32_t w, uint32_t lut) {
    const uint32_t x = w & 0x33333333u;
    const uint32_t y = __umulhi(w, 0x40000000u) & 0x33333333u;
    return make_uint4(prmt(lut, 0u, x), prmt(lut, 0u, y), prmt(lut, 0u, __umulhi(x, 0x10000u)),
                      prmt(lut, 0u, __umulhi(y, 0x10000u)));
}

// SwiGLU glue silu(g) * u (crates/bitnet-transformer/src/lib.rs:756-789), ONE formula for every kernel that applies
// it (decode epilogues, the GEMM epilogue, the stand-alone passes), so all paths agree bit for bit:
//   g / (1 + exp(-g)) with the accurate expf and the approximate division (MUFU.RCP-based, <= 2 ulp).
// The IEEE division it replaces took its slow path (a ~100-instruction subroutine) whenever exp(-g) overflowed to
// inf or the quotient was denormal -- every strongly negative gate -- which made the GEMM's SwiGLU epilogue 4.6 x
// slower than its MMA loop. __fdividef(g, inf) = -0 for g < 0, exactly what g / inf gives.
__device__ __forceinline__ float silu_mul(float g, float u) { return __fmul_rn(__fdividef(g, 1.0f + expf(-g)), u); }

// ---- PTX: mbarrier, 1-D bulk copy (TMA), programmatic dependent launch ----------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(void* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(void* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(void* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// bounded wait: a lost completion traps (reported as a launch failure) instead of hanging the GPU. (A non-blocking
// mbarrier.test_wait probe in front of the try_wait loop was measured: no difference in any kernel here.)
__device__ __forceinline__ void mbar_wait(void* bar, uint32_t parity) {
    for (uint32_t it = 0; !mbar_try_wait(bar, parity); ++it)
        if (it > (1u << 28)) __trap();
}
// global -> shared bulk copy; size and both addresses multiples of 16 B
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes, void* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// PDL: let the next kernel of the stream start launching / wait for the previous one to finish
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Sum V per-thread values over the 32 lanes of a warp. On return lane l holds in v[0] the
// total of value index (l >> (5 - log2 V)).
template <int V, typename T>
__device__ __forceinline__ void warp_reduce_transpose(T (&v)[V]) {
    const int lane = threadIdx.x & 31;
    int off = 16;
#pragma unroll
    for (int cnt = V / 2; cnt >= 1; cnt >>= 1) {
        const bool hi = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < cnt; ++i) {
            T send = hi ? v[i] : v[i + cnt];
            T keep = hi ? v[i + cnt] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
        off >>= 1;
    }
#pragma unroll
    for (; off >= 1; off >>= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], off);
}

template <int V>
struct Log2 {
    static constexpr int value = 1 + Log2<V / 2>::value;
};
template <>
struct Log2<1> {
    static constexpr int value = 0;
};

}  // namespace b200
#endif
