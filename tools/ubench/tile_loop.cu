// micro-benchmark: throughput of the decode tile body (LDS.32 fragment -> PRMT decode -> IMMA.16832) with 15 warps per SM,
// data resident in shared memory (no TMA): what the ALU / tensor pipes allow, apart from the ring and the phase structure.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench/tile_loop tools/ubench/tile_loop.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ void mma(int (&c)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t s) { uint32_t d; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s)); return d; }
// VAR 0: IMAD.HI shifts (product code)   1: SHF shifts   2: no decode (raw words into IMMA)   3: decode, no IMMA (xor)
template <int VAR>
__device__ __forceinline__ void body(uint32_t w, uint32_t lut, uint2 b, int (&c)[4]) {
    uint32_t x, y, z, v;
    if (VAR == 2) { x = w; y = w ^ lut; z = w + lut; v = w - lut; }
    else if (VAR == 4) {  // mixed: one of the three shifts on the FMA pipe (IMAD.HI), two on the ALU pipe (SHF)
        const uint32_t m0 = w & 0x33333333u;
        const uint32_t m1 = (w >> 2) & 0x33333333u;
        x = prmt(lut, 0u, m0); y = prmt(lut, 0u, m1);
        z = prmt(lut, 0u, __umulhi(m0, 0x10000u));
        v = prmt(lut, 0u, m1 >> 16);
    } else if (VAR == 5) {  // PRMT as the shifter: byte-select the upper halfword instead of shifting
        const uint32_t m0 = w & 0x33333333u;
        const uint32_t m1 = (w >> 2) & 0x33333333u;
        x = prmt(lut, 0u, m0); y = prmt(lut, 0u, m1);
        z = prmt(lut, 0u, prmt(m0, 0u, 0x4432u));
        v = prmt(lut, 0u, prmt(m1, 0u, 0x4432u));
    } else {
        const uint32_t m0 = w & 0x33333333u;
        const uint32_t m1 = (VAR == 1 ? (w >> 2) : __umulhi(w, 0x40000000u)) & 0x33333333u;
        x = prmt(lut, 0u, m0); y = prmt(lut, 0u, m1);
        z = prmt(lut, 0u, VAR == 1 ? (m0 >> 16) : __umulhi(m0, 0x10000u));
        v = prmt(lut, 0u, VAR == 1 ? (m1 >> 16) : __umulhi(m1, 0x10000u));
    }
    if (VAR == 3) { c[0] ^= x; c[1] ^= y; c[2] ^= z; c[3] ^= v + b.x + b.y; }
    else mma(c, x, y, z, v, b.x, b.y);
}
template <int VAR, int NC, bool BAR>
__global__ void __launch_bounds__(512, 1) k(int* out, long long* cyc, int tiles, uint32_t lut) {
    extern __shared__ __align__(128) unsigned char sm[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(sm);
    uint32_t* ring = reinterpret_cast<uint32_t*>(sm + 128);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 40 * 1024; i += 512) ring[i] = i * 2654435761u;
    if (tid == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1048575;" ::"r"((uint32_t)__cvta_generic_to_shared(bar + 1)));
                    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar))); 
                    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory"); }
    __syncthreads();
    if (warp == 15) return;
    uint2 b[NC];
    for (int c = 0; c < NC; ++c) b[c] = make_uint2(ring[lane + c], ring[lane + 7 * c + 1]);
    int acc0[4] = {0, 0, 0, 0}, acc1[4] = {0, 0, 0, 0};
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < tiles; ++i) {
        if (BAR) {
            uint32_t ok = 0;
            while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(0u) : "memory");
        }
        const uint32_t* base = ring + ((i & 7) * 80 + warp * NC) * 32 + lane;
        uint32_t w[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) w[c] = base[c * 32];
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            if (c & 1) body<VAR>(w[c], lut, b[c], acc1);
            else body<VAR>(w[c], lut, b[c], acc0);
        }
        if (BAR) { __syncwarp(); if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(bar + 1)) : "memory"); }
    }
    long long t1 = clock64();
    out[blockIdx.x * 512 + tid] = acc0[0] + acc0[1] + acc0[2] + acc0[3] + acc1[0] + acc1[1] + acc1[2] + acc1[3];
    if (tid == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    int* out; long long* cyc; cudaMalloc(&out, 148 * 512 * 4); cudaMalloc(&cyc, 8);
    long long h; const int tiles = 4096; const size_t smem = 128 + 160 * 1024;
#define RUN(name, VAR, NC, BAR)                                                                         \
    cudaFuncSetAttribute(k<VAR, NC, BAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);      \
    k<VAR, NC, BAR><<<148, 512, smem>>>(out, cyc, tiles, 0x0201fffe); cudaDeviceSynchronize();          \
    k<VAR, NC, BAR><<<148, 512, smem>>>(out, cyc, tiles, 0x0201fffe); cudaDeviceSynchronize();          \
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);                                                     \
    printf("%-44s cycles/tile/warp %7.1f  cycles per chunk per SM %.2f  -> %.1f B/clk/SM  %s\n", name, (double)h / tiles, \
           (double)h / tiles / (15.0 * NC), 128.0 * 15 * NC * tiles / h, cudaGetErrorString(cudaGetLastError()));
    RUN("imad.hi decode + imma, 6 chunks", 0, 6, false)
    RUN("shf decode + imma, 6 chunks", 1, 6, false)
    RUN("mixed (1 imad.hi + 2 shf) + imma, 6 chunks", 4, 6, false)
    RUN("prmt-as-shift decode + imma, 6 chunks", 5, 6, false)
    RUN("no decode, imma only, 6 chunks", 2, 6, false)
    RUN("imad.hi decode, no imma, 6 chunks", 3, 6, false)
    RUN("imad.hi decode + imma, 6 chunks, mbarrier", 0, 6, true)
    RUN("imad.hi decode + imma, 15 chunks", 0, 15, false)
    RUN("shf decode + imma, 15 chunks", 1, 15, false)
    RUN("no decode, imma only, 15 chunks", 2, 15, false)
    return 0;
}
