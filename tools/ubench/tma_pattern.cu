// micro-benchmark: HBM -> shared bandwidth of the access patterns the small-M ternary linears use to stream a row-major
// [N, K/4] code plane, with no compute at all (a ring of stages, each refilled as soon as it lands):
//   0  2-D TMA box  64 B x 128 rows (SWIZZLE_64B)   -- gemm_tcgen05.cu's packed-code stage
//   1  2-D TMA box 128 B x  64 rows (SWIZZLE_128B)
//   2  2-D TMA box 256 B x  32 rows (no swizzle)
//   3  1-D bulk copies of 8 KB from a tile-major copy of the plane (what a repacked layout would allow)
//   4  3-D TMA box 64 B x 32 rows x 4 column blocks  -- gemm_small.cu's code stage
// One CTA per SM, one thread issues. Every pattern moves the same bytes; the matrix is rotated through > L2.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench/tma_pattern tools/ubench/tma_pattern.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x)                                                                       \
    do {                                                                            \
        cudaError_t e = (x);                                                        \
        if (e != cudaSuccess) {                                                     \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
            exit(1);                                                                \
        }                                                                           \
    } while (0)

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c)); }
__device__ __forceinline__ void expect_tx(void* b, uint32_t n) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory");
}
__device__ __forceinline__ void wait(void* b, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok)
                     : "r"(s32(b)), "r"(parity)
                     : "memory");
}
__device__ __forceinline__ void tma2(void* dst, const CUtensorMap* tm, int c0, int c1, void* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(s32(dst)),
                 "l"(tm), "r"(s32(bar)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tma3(void* dst, const CUtensorMap* tm, int c0, int c1, int c2, void* bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(s32(dst)),
                 "l"(tm), "r"(s32(bar)), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void bulk(void* dst, const void* src, uint32_t n, void* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(dst)), "l"(src), "r"(n),
                 "r"(s32(bar))
                 : "memory");
}

constexpr int STAGE = 8192, NS = 8;

// units of 8 KB: the matrix is N rows x P bytes (P = K / 4); pattern decides how a unit maps to (row, byte) coordinates
__global__ void __launch_bounds__(32, 1) stream_kernel(const __grid_constant__ CUtensorMap tm, const uint8_t* base, int pattern, int N,
                                                       int P, int units, unsigned long long* sink) {
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    unsigned char* ring = smem + 1024;
    if (threadIdx.x != 0) return;
    for (int s = 0; s < NS; ++s) mbar_init(&bar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    auto issue = [&](int u, int s) {
        expect_tx(&bar[s], STAGE);
        unsigned char* dst = ring + (size_t)s * STAGE;
        if (pattern == 0) {  // 64 B x 128 rows: units walk along the row inside a 128-row tile
            const int per_tile = P / 64, t = u / per_tile, c = u % per_tile;
            tma2(dst, &tm, c * 64, t * 128, &bar[s]);
        } else if (pattern == 1) {  // 128 B x 64 rows
            const int per_tile = P / 128, t = u / per_tile, c = u % per_tile;
            tma2(dst, &tm, c * 128, t * 64, &bar[s]);
        } else if (pattern == 2) {  // 256 B x 32 rows
            const int per_tile = P / 256, t = u / per_tile, c = u % per_tile;
            tma2(dst, &tm, c * 256, t * 32, &bar[s]);
        } else if (pattern == 3) {
            bulk(dst, base + (size_t)u * STAGE, STAGE, &bar[s]);
        } else {  // 64 B x 32 rows x 4 column blocks
            const int per_tile = P / 256, t = u / per_tile, c = u % per_tile;
            tma3(dst, &tm, 0, t * 32, c * 4, &bar[s]);
        }
    };
    int n = 0;
    for (int u = blockIdx.x; u < units && n < NS; u += gridDim.x, ++n) issue(u, n);
    int i = 0;
    unsigned long long acc = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x, ++i) {
        const int s = i % NS;
        wait(&bar[s], (i / NS) & 1);
        acc += ring[(size_t)s * STAGE + (i & 1023)];
        const int un = u + NS * gridDim.x;
        if (un < units) issue(un, s);
    }
    if (acc == 0x123456789ull) *sink = acc;
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
    const int N = argc > 1 ? atoi(argv[1]) : 28672, K = argc > 2 ? atoi(argv[2]) : 4096;
    const int P = K / 4, copies = 12;
    const size_t bytes = (size_t)N * P;
    uint8_t* d;
    CK(cudaMalloc(&d, bytes * copies));
    CK(cudaMemset(d, 1, bytes * copies));
    unsigned long long* sink;
    CK(cudaMalloc(&sink, 8));
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
    EncodeFn enc = (EncodeFn)fp;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 1024 + NS * STAGE + 1024));
    const char* names[] = {"2-D box 64 B x 128 rows (SW64)", "2-D box 128 B x 64 rows (SW128)", "2-D box 256 B x 32 rows (none)",
                           "1-D bulk 8 KB contiguous", "3-D box 64 B x 32 rows x 4 blocks (SW64)"};
    for (int pat = 0; pat < 5; ++pat) {
        CUtensorMap tms[copies];
        for (int c = 0; c < copies; ++c) {
            uint8_t* base = d + (size_t)c * bytes;
            CUresult r = CUDA_SUCCESS;
            if (pat == 4) {
                cuuint64_t gdim[3] = {64, (cuuint64_t)N, (cuuint64_t)(P / 64)};
                cuuint64_t gstr[2] = {(cuuint64_t)P, 64};
                cuuint32_t box[3] = {64, 32, 4}, es[3] = {1, 1, 1};
                r = enc(&tms[c], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            } else {
                const int bw = pat == 0 ? 64 : pat == 1 ? 128 : 256, br = STAGE / bw;
                cuuint64_t gdim[2] = {(cuuint64_t)P, (cuuint64_t)N};
                cuuint64_t gstr[1] = {(cuuint64_t)P};
                cuuint32_t box[2] = {(cuuint32_t)bw, (cuuint32_t)br}, es[2] = {1, 1};
                r = enc(&tms[c], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        pat == 0 ? CU_TENSOR_MAP_SWIZZLE_64B : pat == 1 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            }
            if (r != CUDA_SUCCESS) {
                printf("encode failed pattern %d: %d\n", pat, (int)r);
                return 1;
            }
        }
        const int units = (int)(bytes / STAGE);
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0));
        CK(cudaEventCreate(&e1));
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {
            CK(cudaEventRecord(e0));
            for (int c = 0; c < copies; ++c)
                stream_kernel<<<sms, 32, 1024 + NS * STAGE + 1024>>>(tms[c], d + (size_t)c * bytes, pat, N, P, units, sink);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep > 0 && ms < best) best = ms;
        }
        printf("%-44s %6.1f us per %.1f MB pass  %7.1f GB/s\n", names[pat], best * 1e3 / copies, bytes / 1e6, bytes * copies / (best * 1e6));
    }
    return 0;
}
