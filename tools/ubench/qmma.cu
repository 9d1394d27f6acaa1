// micro-benchmark: legacy mma.sync FP8 (e4m3 x e4m3 -> f32, m16n8k32) on sm_100a:
//  (1) is a long accumulation chain of small-integer products EXACT (no reduced-precision internal accumulate)?
//  (2) latency / throughput next to IMMA.16832
//  (3) throughput of the sign/magnitude decode  reg_j = ((w << j) & 0x88888888) | 0x40404040  feeding it
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench/qmma tools/ubench/qmma.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
__device__ __forceinline__ void qmma(float (&c)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// e4m3 encoding of a small integer n in [-16, 16]
__host__ __device__ inline uint8_t e4m3_int(int n) {
    if (n == 0) return 0;
    uint8_t s = n < 0 ? 0x80 : 0;
    int a = n < 0 ? -n : n, e = 0;
    while ((a >> (e + 1)) != 0) ++e;       // a in [2^e, 2^(e+1))
    int mant = ((a << 3) >> e) & 7;         // exact for a <= 16
    return s | (uint8_t)((e + 7) << 3) | (uint8_t)mant;
}
// exactness: each warp accumulates `chain` MMAs; A[row][k], B[k][col] small integers given per (chain step) from global
__global__ void k_exact(const uint8_t* A, const uint8_t* B, float* out, int chain) {
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    float c[4] = {0, 0, 0, 0};
    for (int s = 0; s < chain; ++s) {
        const uint8_t* a = A + (size_t)s * 16 * 32;  // [16][32]
        const uint8_t* b = B + (size_t)s * 8 * 32;   // [8 cols][32 k]
        auto ld4 = [](const uint8_t* p) { return *reinterpret_cast<const uint32_t*>(p); };
        qmma(c, ld4(a + g * 32 + 4 * t), ld4(a + (g + 8) * 32 + 4 * t), ld4(a + g * 32 + 16 + 4 * t), ld4(a + (g + 8) * 32 + 16 + 4 * t),
             ld4(b + g * 32 + 4 * t), ld4(b + g * 32 + 16 + 4 * t));
    }
    out[(g)*8 + 2 * t] = c[0]; out[g * 8 + 2 * t + 1] = c[1]; out[(g + 8) * 8 + 2 * t] = c[2]; out[(g + 8) * 8 + 2 * t + 1] = c[3];
}
template <int NACC>
__global__ void k_tput(float* out, long long* cyc, int iters, unsigned seed) {
    float c[NACC][4] = {};
    unsigned a0 = 0x40404040u | ((seed + threadIdx.x) & 0x88888888u), a1 = a0 ^ 0x80000000u, a2 = a0 ^ 0x8000u, a3 = a0 ^ 0x80u, b0 = 0x38383838u, b1 = 0x40404040u;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < NACC; ++j) qmma(c[j], a0, a1, a2, a3, b0, b1);
    }
    long long t1 = clock64();
    float s = 0;
    for (int j = 0; j < NACC; ++j) s += c[j][0] + c[j][1] + c[j][2] + c[j][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
// decode + qmma loop, 15 warps, fragments in shared memory
template <int NC, int SH>
__global__ void __launch_bounds__(512, 1) k_loop(float* out, long long* cyc, int tiles) {
    extern __shared__ __align__(128) unsigned char sm[];
    uint32_t* ring = reinterpret_cast<uint32_t*>(sm);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 40 * 1024; i += 512) ring[i] = i * 2654435761u;
    __syncthreads();
    if (warp == 15) return;
    uint2 b[NC];
    for (int c = 0; c < NC; ++c) b[c] = make_uint2(0x38404838u + (lane & 1) * 0x80, 0x40383840u);
    float acc0[4] = {0, 0, 0, 0}, acc1[4] = {0, 0, 0, 0};
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < tiles; ++i) {
        const uint32_t* base = ring + ((i & 7) * 80 + warp * NC) * 32 + lane;
        uint32_t w[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) w[c] = base[c * 32];
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const uint32_t x = w[c];
            uint32_t r0 = (x & 0x88888888u) | 0x40404040u, r1, r2, r3;
            if (SH == 0) { r1 = ((x << 1) & 0x88888888u) | 0x40404040u; r2 = ((x << 2) & 0x88888888u) | 0x40404040u; r3 = ((x << 3) & 0x88888888u) | 0x40404040u; }
            else { r1 = ((x * 2u) & 0x88888888u) | 0x40404040u; r2 = ((x * 4u) & 0x88888888u) | 0x40404040u; r3 = ((x * 8u) & 0x88888888u) | 0x40404040u; }
            if (c & 1) qmma(acc1, r0, r1, r2, r3, b[c].x, b[c].y);
            else qmma(acc0, r0, r1, r2, r3, b[c].x, b[c].y);
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * 512 + tid] = acc0[0] + acc0[1] + acc0[2] + acc0[3] + acc1[0] + acc1[1] + acc1[2] + acc1[3];
    if (tid == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    // ---- exactness
    const int chain = 240;  // K = 7680
    std::vector<uint8_t> A((size_t)chain * 16 * 32), B((size_t)chain * 8 * 32);
    std::vector<int> Ai(A.size()), Bi(B.size());
    int bad_total = 0;
    for (int trial = 0; trial < 6; ++trial) {
        srand(1234 + trial);
        for (size_t i = 0; i < A.size(); ++i) {
            int v;
            if (trial == 0) v = 4; else if (trial == 1) v = (i & 1) ? 4 : -4; else { const int tab[4] = {-4, -2, 2, 4}; v = tab[rand() & 3]; }
            Ai[i] = v; A[i] = e4m3_int(v);
        }
        for (size_t i = 0; i < B.size(); ++i) {
            int v;
            if (trial == 0) v = 8; else if (trial == 1) v = (i & 1) ? -8 : 7; else if (trial == 2) v = (rand() % 17) - 8; else if (trial == 3) v = (rand() & 1) ? 8 : -8;
            else v = (rand() % 3 == 0) ? 8 : (rand() % 17) - 8;
            Bi[i] = v; B[i] = e4m3_int(v);
        }
        uint8_t *dA, *dB; float* dO;
        cudaMalloc(&dA, A.size()); cudaMalloc(&dB, B.size()); cudaMalloc(&dO, 128 * 4);
        cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice); cudaMemcpy(dB, B.data(), B.size(), cudaMemcpyHostToDevice);
        k_exact<<<1, 32>>>(dA, dB, dO, chain); cudaDeviceSynchronize();
        float o[128]; cudaMemcpy(o, dO, sizeof(o), cudaMemcpyDeviceToHost);
        int bad = 0; long long maxabs = 0;
        for (int r = 0; r < 16; ++r) for (int cidx = 0; cidx < 8; ++cidx) {
            long long ref = 0;
            for (int s = 0; s < chain; ++s) for (int k = 0; k < 32; ++k) ref += (long long)Ai[((size_t)s * 16 + r) * 32 + k] * Bi[((size_t)s * 8 + cidx) * 32 + k];
            if ((long long)o[r * 8 + cidx] != ref || o[r * 8 + cidx] != (float)ref) { if (bad < 3) printf("  mismatch trial %d r %d c %d: got %.1f want %lld\n", trial, r, cidx, o[r * 8 + cidx], ref); ++bad; }
            if (llabs(ref) > maxabs) maxabs = llabs(ref);
        }
        printf("exactness trial %d: %d mismatches of 128 (max |sum| %lld)  %s\n", trial, bad, maxabs, cudaGetErrorString(cudaGetLastError()));
        bad_total += bad;
        cudaFree(dA); cudaFree(dB); cudaFree(dO);
    }
    printf("EXACT: %s\n", bad_total ? "NO" : "YES");
    // ---- throughput
    float* out; long long* cyc; cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&cyc, 8);
    long long h; const int it = 2048;
#define RUN(name, kern, nacc, threads)                                                   \
    kern<<<148, threads>>>(out, cyc, it, 1); cudaDeviceSynchronize();                    \
    kern<<<148, threads>>>(out, cyc, it, 1); cudaDeviceSynchronize();                    \
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);                                      \
    printf("%-28s warps/SM=%2d  cycles/op/warp=%.2f  (ops per SM per cycle %.3f)\n", name, threads / 32, (double)h / (it * nacc), (double)(it * nacc) * (threads / 32) / h);
    RUN("qmma 1 acc (latency)", k_tput<1>, 1, 32)
    RUN("qmma 4 acc 4 warps", k_tput<4>, 4, 128)
    RUN("qmma 4 acc 16 warps", k_tput<4>, 4, 512)
    const int tiles = 4096; const size_t smem = 160 * 1024;
#define RUNL(name, NC, SH)                                                                              \
    cudaFuncSetAttribute(k_loop<NC, SH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);       \
    k_loop<NC, SH><<<148, 512, smem>>>(out, cyc, tiles); cudaDeviceSynchronize();                       \
    k_loop<NC, SH><<<148, 512, smem>>>(out, cyc, tiles); cudaDeviceSynchronize();                       \
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);                                                     \
    printf("%-36s cycles/tile/warp %7.1f  cycles per chunk per SM %.2f  -> %.1f B/clk/SM  %s\n", name, (double)h / tiles, \
           (double)h / tiles / (15.0 * NC), 128.0 * 15 * NC * tiles / h, cudaGetErrorString(cudaGetLastError()));
    RUNL("sign/mag decode (shl) + qmma, 6", 6, 0)
    RUNL("sign/mag decode (imul) + qmma, 6", 6, 1)
    RUNL("sign/mag decode (shl) + qmma, 15", 15, 0)
    RUNL("sign/mag decode (imul) + qmma, 15", 15, 1)
    return 0;
}
