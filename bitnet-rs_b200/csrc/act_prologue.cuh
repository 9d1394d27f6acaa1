// act_prologue.cuh -- the activation side of the ternary linear as a GEMV PROLOGUE: every CTA
// turns the f32 token row(s) into int8 + scale in its own shared memory.
//   PRO_Q8   rows are already int8 (+ scale): copy
//   PRO_F32  per-token absmax int8, quantize_symmetric_i8(row, 8)  (cpu/quantize.rs:32-49)
//   PRO_RMS  RMSNorm first, cpu::rms_norm (cpu/layer_norm.rs:118-137: f64 mean of squares,
//            x * (1/sqrt(ms+eps)) * gamma), then the same quantiser
// Bit-exact restatements; the summation order of the f64 mean differs from the reference's
// serial loop only below f64 rounding.
#pragma once
#include "common.cuh"

namespace b200 {

enum { PRO_Q8 = 0, PRO_F32 = 1, PRO_RMS = 2 };
enum { EPI_STORE = 0, EPI_RESIDUAL = 1, EPI_SWIGLU = 2 };

constexpr int NVMAX = 8;  // float4 per thread held by the fused prologue: K <= 32 * blockDim

// q = clamp(round_half_away(v * inv), -127, 127); NaN -> 0 (Rust `as i8`)
__device__ __forceinline__ int quant_i8(float v, float inv) {
    float r = roundf(v * inv);
    if (r != r) return 0;
    r = fminf(fmaxf(r, -127.f), 127.f);
    return (int)r;
}

struct PrologueArgs {
    const int8_t* q;         // PRO_Q8: [M, K]
    const float* act_scale;  // PRO_Q8: [M]
    const float* x;          // PRO_F32 / PRO_RMS: [M, K]
    const float* gamma;      // PRO_RMS: [K]
    int8_t* q_out;           // optional, CTA 0 only: the int8 rows
    float* scale_out;        // optional, CTA 0 only
    float* normed_out;       // optional, CTA 0 only (PRO_RMS)
    float eps;
    int pro, K, M;
};

// All threads of the CTA call this. qbuf: [M][q_ld] int8 (q_ld >= K, multiple of 16),
// sc_s: [M] float, shd: [32] double, shf: [32] float scratch. Ends with a __syncthreads().
__device__ __forceinline__ void act_prologue(const PrologueArgs& a, int8_t* qbuf, int q_ld, float* sc_s,
                                             double* shd, float* shf) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    if (a.pro == PRO_Q8) {
        const int vec_per_row = a.K / 16;
        for (int i = tid; i < a.M * vec_per_row; i += blockDim.x) {
            const int m = i / vec_per_row, v = i - m * vec_per_row;
            *reinterpret_cast<uint4*>(qbuf + m * q_ld + v * 16) =
                *reinterpret_cast<const uint4*>(a.q + (size_t)m * a.K + v * 16);
        }
        if (tid < a.M) sc_s[tid] = a.act_scale[tid];
        __syncthreads();
        return;
    }
    const int ngrp = a.K / 4;  // float4 groups per row
    for (int m = 0; m < a.M; ++m) {  // one token at a time: 32 live registers, 2-3 barriers each
        float4 v[NVMAX];
#pragma unroll
        for (int j = 0; j < NVMAX; ++j) {
            const int g = tid + j * blockDim.x;
            v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (g < ngrp) v[j] = *reinterpret_cast<const float4*>(a.x + (size_t)m * a.K + 4 * g);
        }
        if (a.pro == PRO_RMS) {
            double ss = 0.0;
#pragma unroll
            for (int j = 0; j < NVMAX; ++j)
                ss += (double)v[j].x * (double)v[j].x + (double)v[j].y * (double)v[j].y +
                      (double)v[j].z * (double)v[j].z + (double)v[j].w * (double)v[j].w;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
            if (lane == 0) shd[warp] = ss;
            __syncthreads();
            double tot = 0.0;
            for (int wv = 0; wv < nwarps; ++wv) tot += shd[wv];
            const float ms = (float)(tot / (double)a.K);
            const float inv_rms = 1.0f / sqrtf(ms + a.eps);
#pragma unroll
            for (int j = 0; j < NVMAX; ++j) {
                const int g = tid + j * blockDim.x;
                if (g < ngrp) {
                    const float4 gm = *reinterpret_cast<const float4*>(a.gamma + 4 * g);
                    v[j].x = v[j].x * inv_rms * gm.x;
                    v[j].y = v[j].y * inv_rms * gm.y;
                    v[j].z = v[j].z * inv_rms * gm.z;
                    v[j].w = v[j].w * inv_rms * gm.w;
                }
            }
        }
        float amax = 0.f;
#pragma unroll
        for (int j = 0; j < NVMAX; ++j)
            amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v[j].x), fabsf(v[j].y)), fmaxf(fabsf(v[j].z), fabsf(v[j].w))));
        amax = warp_max(amax);
        if (lane == 0) shf[warp] = amax;
        __syncthreads();  // also orders the shd reads above against the next token's writes
        amax = 0.f;
        for (int wv = 0; wv < nwarps; ++wv) amax = fmaxf(amax, shf[wv]);
        const float sc = amax == 0.f ? 0.f : amax / 127.f;
        const float inv = amax == 0.f ? 0.f : 1.0f / sc;
        if (tid == 0) {
            sc_s[m] = sc;
            if (blockIdx.x == 0 && a.scale_out) a.scale_out[m] = sc;
        }
#pragma unroll
        for (int j = 0; j < NVMAX; ++j) {
            const int g = tid + j * blockDim.x;
            if (g < ngrp) {
                uint32_t wq = 0;
                if (amax != 0.f)
                    wq = (uint32_t)(uint8_t)quant_i8(v[j].x, inv) | ((uint32_t)(uint8_t)quant_i8(v[j].y, inv) << 8) |
                         ((uint32_t)(uint8_t)quant_i8(v[j].z, inv) << 16) | ((uint32_t)(uint8_t)quant_i8(v[j].w, inv) << 24);
                reinterpret_cast<uint32_t*>(qbuf + m * q_ld)[g] = wq;
                if (blockIdx.x == 0) {
                    if (a.q_out) reinterpret_cast<uint32_t*>(a.q_out + (size_t)m * a.K)[g] = wq;
                    if (a.normed_out) *reinterpret_cast<float4*>(a.normed_out + (size_t)m * a.K + 4 * g) = v[j];
                }
            }
        }
        __syncthreads();  // shf reads done before the next token overwrites it; qbuf visible
    }
}

}  // namespace b200
