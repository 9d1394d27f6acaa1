// micro-benchmark: latency / throughput of legacy mma.sync int8 (IMMA.16832) vs dp4a on sm_100a
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void mma(int (&c)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
template <int NACC>
__global__ void k_mma(int* out, long long* cyc, int iters, unsigned seed) {
    int c[NACC][4] = {};
    unsigned a0 = seed + threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < NACC; ++j) mma(c[j], a0, a1, a2, a3, b0, b1);
    }
    long long t1 = clock64();
    int s = 0;
    for (int j = 0; j < NACC; ++j) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int NACC>
__global__ void k_dp4a(int* out, long long* cyc, int iters, unsigned seed) {
    int c[NACC] = {};
    unsigned a = seed + threadIdx.x, b = a * 11;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < NACC; ++j) asm volatile("dp4a.s32.s32 %0, %1, %2, %0;" : "+r"(c[j]) : "r"(a), "r"(b));
    }
    long long t1 = clock64();
    int s = 0;
    for (int j = 0; j < NACC; ++j) s += c[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    int* out; long long* cyc; cudaMalloc(&out, 1 << 22); cudaMalloc(&cyc, 8);
    long long h;
    const int it = 2048;
#define RUN(name, kern, nacc, blocks, threads)                                          \
    kern<<<blocks, threads>>>(out, cyc, it, 1); cudaDeviceSynchronize();                 \
    kern<<<blocks, threads>>>(out, cyc, it, 1); cudaDeviceSynchronize();                 \
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);                                      \
    printf("%-28s warps/SM=%2d  cycles/op/warp=%.2f  (ops per SM per cycle %.3f)\n", name, threads / 32, (double)h / (it * nacc), \
           (double)(it * nacc) * (threads / 32) / h);
    RUN("imma 1 acc (latency)", k_mma<1>, 1, 148, 32)
    RUN("imma 4 acc 1 warp", k_mma<4>, 4, 148, 32)
    RUN("imma 4 acc 4 warps", k_mma<4>, 4, 148, 128)
    RUN("imma 4 acc 16 warps", k_mma<4>, 4, 148, 512)
    RUN("dp4a 1 acc (latency)", k_dp4a<1>, 1, 148, 32)
    RUN("dp4a 8 acc 4 warps", k_dp4a<8>, 8, 148, 128)
    RUN("dp4a 8 acc 16 warps", k_dp4a<8>, 8, 148, 512)
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
