// micro-benchmark: does the binary tensor path (mma.sync m16n8k256 b1 and.popc) exist on sm_100a, and how fast is it?
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void bmma(int (&c)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k256.row.col.s32.b1.b1.s32.and.popc {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
template <int NACC>
__global__ void k_bmma(int* out, long long* cyc, int iters, unsigned seed) {
    int c[NACC][4] = {};
    unsigned a0 = seed + threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < NACC; ++j) bmma(c[j], a0, a1, a2, a3, b0, b1);
    }
    long long t1 = clock64();
    int s = 0;
    for (int j = 0; j < NACC; ++j) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
// correctness probe: one warp, known pattern
__global__ void k_check(int* out) {
    int c[4] = {0, 0, 0, 0};
    const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
    // A row r: all ones in k < 8*(r+1); B col n: all ones in k < 16*(n+1)  => C[r][n] = min(8(r+1), 16(n+1))
    auto rowbits = [](int r, int word) { int lim = 8 * (r + 1); int lo = word * 32; unsigned v = 0; for (int b = 0; b < 32; ++b) if (lo + b < lim) v |= 1u << b; return v; };
    auto colbits = [](int n, int word) { int lim = 16 * (n + 1); int lo = word * 32; unsigned v = 0; for (int b = 0; b < 32; ++b) if (lo + b < lim) v |= 1u << b; return v; };
    // fragment layout (PTX ISA, m16n8k256 .b1): a0: row g, k = 32t..32t+31 (+0); a1: row g+8 same k; a2: row g, k = 128+32t..; a3: row g+8, k = 128+32t..
    // b0: col g, k = 32t..; b1: col g, k = 128 + 32t..
    unsigned a0 = rowbits(g, t), a1 = rowbits(g + 8, t), a2 = rowbits(g, 4 + t), a3 = rowbits(g + 8, 4 + t);
    unsigned b0 = colbits(g, t), b1 = colbits(g, 4 + t);
    bmma(c, a0, a1, a2, a3, b0, b1);
    out[lane * 4 + 0] = c[0]; out[lane * 4 + 1] = c[1]; out[lane * 4 + 2] = c[2]; out[lane * 4 + 3] = c[3];
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
    RUN("bmma 1 acc (latency)", k_bmma<1>, 1, 148, 32)
    RUN("bmma 4 acc 1 warp", k_bmma<4>, 4, 148, 32)
    RUN("bmma 4 acc 4 warps", k_bmma<4>, 4, 148, 128)
    RUN("bmma 4 acc 16 warps", k_bmma<4>, 4, 148, 512)
    k_check<<<1, 32>>>(out); cudaDeviceSynchronize();
    int hc[128]; cudaMemcpy(hc, out, sizeof hc, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int lane = 0; lane < 32; ++lane) {
        int g = lane >> 2, t = lane & 3;
        int exp[4] = {min(8 * (g + 1), 16 * (2 * t + 1)), min(8 * (g + 1), 16 * (2 * t + 2)), min(8 * (g + 9), 16 * (2 * t + 1)), min(8 * (g + 9), 16 * (2 * t + 2))};
        for (int j = 0; j < 4; ++j) if (hc[lane * 4 + j] != exp[j]) { if (bad < 8) printf("lane %d c[%d] = %d expected %d\n", lane, j, hc[lane * 4 + j], exp[j]); ++bad; }
    }
    printf("fragment-layout check: %s (%d mismatches)\n", bad ? "FAIL" : "ok", bad);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
