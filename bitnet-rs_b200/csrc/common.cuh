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

// Decode one 32-bit word of sixteen 2-bit codes through a 4-entry int8 LUT (one PRMT per
// four weights) and dot it with sixteen int8 activations that were pre-permuted into
//   E0=[a0,a2,a4,a6] O0=[a1,a3,a5,a7] E1=[a8,a10,a12,a14] O1=[a9,a11,a13,a15].
__device__ __forceinline__ int dot16(uint32_t w, uint32_t lut, uint32_t E0, uint32_t O0,
                                     uint32_t E1, uint32_t O1, int acc) {
    const uint32_t x = w & 0x33333333u;
    const uint32_t y = (w >> 2) & 0x33333333u;
    acc = dp4a_ss(prmt(lut, 0u, x), E0, acc);
    acc = dp4a_ss(prmt(lut, 0u, y), O0, acc);
    acc = dp4a_ss(prmt(lut, 0u, x >> 16), E1, acc);
    acc = dp4a_ss(prmt(lut, 0u, y >> 16), O1, acc);
    return acc;
}

// permute 16 consecutive int8 activations (as 4 words) into the E/O order dot16 expects
__device__ __forceinline__ void permute_act16(uint4 a, uint32_t& E0, uint32_t& O0, uint32_t& E1,
                                              uint32_t& O1) {
    E0 = prmt(a.x, a.y, 0x6420u);
    O0 = prmt(a.x, a.y, 0x7531u);
    E1 = prmt(a.z, a.w, 0x6420u);
    O1 = prmt(a.z, a.w, 0x7531u);
}

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
