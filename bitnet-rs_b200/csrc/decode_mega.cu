// decode_mega.cu -- one decode step of a 2B-4T-shaped ternary decoder as ONE persistent,
// cooperative kernel (QK256 layout, batch <= 4): one CTA per SM, a dedicated producer warp that
// streams EVERY packed weight of the step (and the f16 LM head) through a shared-memory ring with
// 1-D bulk copies (TMA, cp.async.bulk -> mbarrier) and never stops at a phase boundary, and 16
// consumer warps that run the phases
//     embed | per layer: [RMSNorm+act-quant | qkv GEMV] -> [RoPE + KV append + GQA attention]
//     -> [act-quant | o GEMV + residual] -> [RMSNorm+act-quant | gate/up GEMV | SwiGLU]
//     -> [act-quant | down GEMV + residual] | final RMSNorm + LM head + argmax
// chained by flagged (value, epoch) pairs: a phase polls exactly the outputs it consumes, there is no grid
// barrier between phases (one per step, before the argmax).
//
// Why: at batch 1 a 2B-4T projection is 0.25-1.4 us of HBM time, below any launch or barrier
// latency, so a chain of per-projection kernels is latency-bound with the memory pipe idle most
// of the time (measured: 4.3-12 us per projection launch). Here the weight stream is decoupled
// from the dependency chain: while the consumers sit in a grid barrier or in the activation
// prologue of phase p, the producer is already filling the ring with the tiles of phase p+1,
// p+2, ... (up to ~1.3 layers ahead across the 148 SMs).
//
// Hot-path arithmetic is that of gemv_mma.cu: 2-bit codes -> 4-entry int8 LUT (PRMT) -> A fragment
// of mma.sync.m16n8k32.s8, int8 activations as the B fragment, int32 accumulation (exact;
// bit-identical to gemv_qk256 on integer inputs, crates/bitnet-quantization/src/i2s_qk256.rs:
// 346-372), per-token absmax int8 quantiser (cpu/quantize.rs:32-49) and cpu::rms_norm
// (cpu/layer_norm.rs:118-137) in the prologue. Cross-warp partial sums meet through int32
// shared-memory atomics: integer addition is associative, so the result is order-independent.
#include "decode_mega.h"

namespace b200 {

constexpr int MEGA_NW = 15;                         // consumer warps
constexpr int MEGA_CONS = MEGA_NW * 32;             // consumer threads
constexpr int MEGA_THREADS = MEGA_CONS + 32;        // + the producer warp
constexpr int MEGA_SLOT_STEPS = 16;                 // 256-column steps per ring slot
constexpr int MEGA_SLOT_BYTES = 16 * MEGA_SLOT_STEPS * 64;  // 16 rows x 16 steps x 64 B
constexpr int MEGA_MAX_TILES = 16;                  // 16-row tiles per CTA per phase
constexpr int MEGA_MAXB = 4;                        // tokens per step
constexpr int MEGA_MAX_SLOTS = 12;
constexpr int MEGA_MAX_LAYERS = 64;
static_assert(MEGA_MAX_TILES * 16 <= MEGA_CONS, "epilogue: one output row per consumer thread");

// ---- small PTX helpers ------------------------------------------------------------------------
__device__ __forceinline__ void cons_sync() { asm volatile("bar.sync 1, %0;" ::"n"(MEGA_CONS) : "memory"); }
__device__ __forceinline__ float ld_cg_f32(const float* p) {
    float v;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ float4 ld_cg_f32x4(const float* p) {
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ int ld_cg_s32(const int* p) {
    int v;
    asm volatile("ld.global.cg.s32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void mma_s8_m16n8k32(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                                uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// ---- flagged dataflow between phases -------------------------------------------------------------
// Every value a phase hands to the next one travels as an 8-byte (f32 bits, epoch) pair written with ONE
// store: an aligned 8-byte store is single-copy atomic, so a reader that sees the expected epoch in the flag
// word has the value too -- no fence, no separate barrier. A consumer simply polls the pairs it needs (it
// reads the whole vector in its activation prologue anyway), which replaces "grid barrier (~2 us), then an
// L2 round trip for the activations" by a single wait of about one L2 round trip after the last producer's
// store. Epochs grow monotonically across phases and launches, so a stale pair can never match.
// Write-after-read safety without barriers: a CTA can only produce out(n+1) after it has seen ALL of out(n),
// i.e. after every CTA finished phase n including its reads of out(n-1); adjacent phase outputs live in
// different buffers (qkv, att, x, gu, x, ...), so nothing still being read is overwritten.
__device__ __forceinline__ void ll_store(uint2* buf, size_t i, float v, uint32_t ep) {
    asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(buf + i), "r"(__float_as_uint(v)), "r"(ep) : "memory");
}
__device__ __forceinline__ uint4 ll_ld2(const uint2* p) {  // two pairs
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint2 ll_ld1(const uint2* p) {
    uint2 v;
    asm volatile("ld.relaxed.gpu.global.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ll_wait1(const uint2* p, uint32_t ep) {
    uint2 v = ll_ld1(p);
    for (uint32_t it = 0; v.y != ep; ++it) {
        if (it > (1u << 24)) __trap();  // a lost producer: fail loudly instead of hanging
        v = ll_ld1(p);
    }
    return __uint_as_float(v.x);
}
// four consecutive values (pairs 4g .. 4g+3): `a`, `b` hold a first probe issued by the caller
__device__ __forceinline__ float4 ll_wait4(const uint2* p, uint4 a, uint4 b, uint32_t ep) {
    for (uint32_t it = 0; a.y != ep || a.w != ep || b.y != ep || b.w != ep; ++it) {
        if (it > (1u << 24)) __trap();
        a = ll_ld2(p);
        b = ll_ld2(p + 2);
    }
    return make_float4(__uint_as_float(a.x), __uint_as_float(a.z), __uint_as_float(b.x), __uint_as_float(b.z));
}

// L2 prefetch of [ptr, ptr + bytes) by the consumer threads, one 128-byte line each. Issued at the END of a
// phase for data the next phase reads first and that the weight stream has evicted since the last step
// (norm weights, the K/V rows of earlier positions, the RoPE row): the miss overlaps the wait for the other
// CTAs' outputs.
// Warps 1..15 only: with BITNET_B200_MEGA_BARRIERS=1 warp 0 arrives at the grid barrier next, and a prefetch in
// flight ahead of its red.release holds the release back by the very miss it is meant to hide (measured, +0.9 us).
__device__ __forceinline__ void prefetch_l2_lines(const void* ptr, size_t bytes) {
    const char* c = reinterpret_cast<const char*>(ptr);
    if (threadIdx.x < 32) return;
    for (size_t o = (size_t)(threadIdx.x - 32) * 128; o < bytes; o += (size_t)(MEGA_CONS - 32) * 128)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(c + o));
}
// q = clamp(round_half_away(v * inv), -127, 127); NaN -> 0 (Rust `as i8`). Round-half-away as
// trunc(y + copysign(0.5 - 2^-25, y)): identical to roundf(y) for every float |y| <= 200 (checked
// exhaustively on the CPU, tests: test_round_half_away_trick), 4 instructions
// instead of ~12; cvt.rzi maps NaN to 0 and saturates +-inf, the clamp does the rest.
__device__ __forceinline__ int mega_quant_i8(float v, float inv) {
    const float y = v * inv;
    const float h = __uint_as_float(0x3effffffu | (__float_as_uint(y) & 0x80000000u));
    const int r = __float2int_rz(__fadd_rn(y, h));
    return max(-127, min(127, r));
}

// diagnostics: CTA 0 / thread 0 appends a globaltimer stamp (ns) when p.prof is set
__device__ __forceinline__ void prof_stamp(const MegaParams& p, int& pi) {
    if (p.prof && blockIdx.x == 0 && threadIdx.x == 0 && pi < 16384) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        p.prof[pi] = t;
    }
    ++pi;
}

// producer-side wait: sleeps between probes (the ring is deep, the producer is never latency-critical)
__device__ __forceinline__ void mbar_wait_backoff(void* bar, uint32_t parity) {
    for (uint32_t it = 0; !mbar_try_wait(bar, parity); ++it) {
        __nanosleep(200);
        if (it > (1u << 24)) __trap();
    }
}

// Grid-wide barrier over the consumer threads of all CTAs (the kernel is launched cooperatively,
// so all CTAs are resident). Arrivals are spread over MEGA_BAR_CTRS counters on separate 256-byte
// lines: same-address atomics serialise at the L2 (~27 cycles each, i.e. ~2 us for 148 CTAs on one
// counter, which is what a single-counter barrier measured); eight lanes of warp 0 poll one counter
// each with ld.acquire and the warp sums them. Bounded spin: a lost arrival traps instead of hanging.
constexpr int MEGA_BAR_CTRS = 8, MEGA_BAR_STRIDE = 64;  // counters per parity, stride in 32-bit words
__device__ __forceinline__ void grid_sync(unsigned* ctr, unsigned& target, unsigned ncta) {
    target += ncta;
    cons_sync();
    if (threadIdx.x < 32) {
        const unsigned lane = threadIdx.x;
        if (lane == 0)
            asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr + (blockIdx.x % MEGA_BAR_CTRS) * MEGA_BAR_STRIDE)
                         : "memory");
        unsigned sum;
        uint32_t it = 0;
        do {
            unsigned v = 0;
            if (lane < MEGA_BAR_CTRS)
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr + lane * MEGA_BAR_STRIDE) : "memory");
            sum = v;
#pragma unroll
            for (int o = 4; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            sum = __shfl_sync(0xffffffffu, sum, 0);
            if (++it > (1u << 24)) __trap();
        } while (sum < target);
    }
    cons_sync();
}

// split `ns` 256-column steps into `nch` nearly equal chunks; chunk c = steps [s0, s0 + cs)
__device__ __forceinline__ void chunk_range(int ns, int nch, int c, int& s0, int& cs) {
    const int base = ns / nch, rem = ns - base * nch;
    s0 = c * base + (c < rem ? c : rem);
    cs = base + (c < rem ? 1 : 0);
}

struct MegaSmem {
    uint64_t* full;    // [n_slots]
    uint64_t* empty;   // [n_slots]
    double* shd;       // [32]
    float* shf;        // [32]
    float* sc_s;       // [MEGA_MAXB]
    int* attw;         // [8] this CTA's attention work item: on, head, seq, split, lo, hi, kv offset (2 words)
    float* bestv;      // [MEGA_NW * MEGA_MAXB]
    int* besti;        // [MEGA_NW * MEGA_MAXB]
    MegaLayer* layers; // [L]
    int* acc;          // [MEGA_MAX_TILES][16][8]
    unsigned char* act;  // int8 rows [B][K + 64]  or  f32 rows [B][H]
    float* attn;       // scores [max_ctx] | value partials [512 float4] | q,k,v rows [3][256]
    unsigned char* ring;
};

// ---- activation prologue: f32 rows -> (optional RMSNorm) -> int8 rows + scale in shared memory ---
// pro: 0 = copy the fixed int8 rows (mode 1), 1 = absmax int8 quantise, 2 = RMSNorm then quantise.
// NV = float4 groups per thread (K <= 16 * NV * MEGA_CONS... i.e. 2 -> K <= 4096, 4 -> K <= 8192).
template <int NV>
__device__ __forceinline__ void mega_prologue_rows(const MegaParams& p, const MegaSmem& sm, int pro, const uint2* xin,
                                                   uint32_t ep_in, const float* gamma, int K, int q_ld) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int8_t* qbuf = reinterpret_cast<int8_t*>(sm.act);
    const int ngrp = K >> 2;
    float4 gm[NV];  // norm weights: fetched with the first row, not after the sum-of-squares barrier
    if (pro == 2) {
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            const int g = tid + j * MEGA_CONS;
            gm[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (g < ngrp) gm[j] = __ldg(reinterpret_cast<const float4*>(gamma) + g);
        }
    }
    for (int m = 0; m < p.B; ++m) {
        float4 v[NV];
        uint4 pa[NV], pb[NV];  // first probe of every group in flight before any flag is looked at
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            const int g = tid + j * MEGA_CONS;
            if (g < ngrp) {
                pa[j] = ll_ld2(xin + (size_t)m * K + 4 * g);
                pb[j] = ll_ld2(xin + (size_t)m * K + 4 * g + 2);
            }
        }
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            const int g = tid + j * MEGA_CONS;
            v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (g < ngrp) v[j] = ll_wait4(xin + (size_t)m * K + 4 * g, pa[j], pb[j], ep_in);
        }
        if (pro == 2) {
            double ss = 0.0;
#pragma unroll
            for (int j = 0; j < NV; ++j)
                ss += (double)v[j].x * (double)v[j].x + (double)v[j].y * (double)v[j].y +
                      (double)v[j].z * (double)v[j].z + (double)v[j].w * (double)v[j].w;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
            if (lane == 0) sm.shd[warp] = ss;
            cons_sync();
            double tot = 0.0;
#pragma unroll
            for (int wv = 0; wv < MEGA_NW; ++wv) tot += sm.shd[wv];
            const float ms = (float)(tot / (double)K);
            const float inv_rms = 1.0f / sqrtf(ms + p.eps);
#pragma unroll
            for (int j = 0; j < NV; ++j) {
                const int g = tid + j * MEGA_CONS;
                if (g < ngrp) {
                    v[j].x = v[j].x * inv_rms * gm[j].x;
                    v[j].y = v[j].y * inv_rms * gm[j].y;
                    v[j].z = v[j].z * inv_rms * gm[j].z;
                    v[j].w = v[j].w * inv_rms * gm[j].w;
                }
            }
        }
        float amax = 0.f;
#pragma unroll
        for (int j = 0; j < NV; ++j)
            amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v[j].x), fabsf(v[j].y)), fmaxf(fabsf(v[j].z), fabsf(v[j].w))));
        amax = warp_max(amax);
        if (lane == 0) sm.shf[warp] = amax;
        cons_sync();
        amax = 0.f;
#pragma unroll
        for (int wv = 0; wv < MEGA_NW; ++wv) amax = fmaxf(amax, sm.shf[wv]);
        const float sc = amax == 0.f ? 0.f : amax / 127.f;
        const float inv = amax == 0.f ? 0.f : 1.0f / sc;
        if (tid == 0) sm.sc_s[m] = sc;
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            const int g = tid + j * MEGA_CONS;
            if (g < ngrp) {
                uint32_t wq = 0;
                if (amax != 0.f)
                    wq = (uint32_t)(uint8_t)mega_quant_i8(v[j].x, inv) | ((uint32_t)(uint8_t)mega_quant_i8(v[j].y, inv) << 8) |
                         ((uint32_t)(uint8_t)mega_quant_i8(v[j].z, inv) << 16) |
                         ((uint32_t)(uint8_t)mega_quant_i8(v[j].w, inv) << 24);
                reinterpret_cast<uint32_t*>(qbuf + m * q_ld)[g] = wq;
            }
        }
        cons_sync();  // shd/shf reusable; qbuf + sc_s visible
    }
}

__device__ __forceinline__ void mega_prologue(const MegaParams& p, const MegaSmem& sm, int pro, const uint2* xin,
                                           uint32_t ep_in, const float* gamma, int K, int q_ld) {
    const int tid = threadIdx.x;
    if (pro == 0) {
        int8_t* qbuf = reinterpret_cast<int8_t*>(sm.act);
        const int vpr = K >> 4;
        for (int m = 0; m < p.B; ++m)
            for (int v = tid; v < vpr; v += MEGA_CONS)
                *reinterpret_cast<uint4*>(qbuf + m * q_ld + v * 16) =
                    *reinterpret_cast<const uint4*>(p.q_fixed + (size_t)m * K + v * 16);
        if (tid < p.B) sm.sc_s[tid] = p.s_fixed[tid];
        cons_sync();
        return;
    }
    if (K <= 8 * MEGA_CONS) mega_prologue_rows<2>(p, sm, pro, xin, ep_in, gamma, K, q_ld);
    else mega_prologue_rows<4>(p, sm, pro, xin, ep_in, gamma, K, q_ld);  // K <= 8192 (mega_plan)
}

// ---- one ternary GEMV phase (consumer side) ------------------------------------------------------
// Ring position of a consumer warp / of the producer: slot index and phase parity, advanced chunk
// by chunk (all parties walk the same chunk sequence).
struct RingPos {
    uint32_t slot, par;
};
__device__ __forceinline__ void ring_advance(RingPos& r, uint32_t n_slots) {
    if (++r.slot == n_slots) {
        r.slot = 0;
        r.par ^= 1u;
    }
}

// One task = 16 weight rows x 256 columns (1 KB of codes): 8 IMMAs over PRMT-decoded fragments.
__device__ __forceinline__ void mega_task(const unsigned char* wr, int rowb, const int8_t* qk, bool tok, uint32_t lut,
                                          int (&c)[4]) {
    const uint4 wg = *reinterpret_cast<const uint4*>(wr);
    const uint4 wh = *reinterpret_cast<const uint4*>(wr + 8 * rowb);
    uint4 b[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        b[j] = make_uint4(0, 0, 0, 0);
        if (tok) b[j] = *reinterpret_cast<const uint4*>(qk + 16 * j);
    }
    int c0[4] = {0, 0, 0, 0}, c1[4] = {0, 0, 0, 0};  // two independent accumulation chains
    const uint32_t wgs[4] = {wg.x, wg.y, wg.z, wg.w};
    const uint32_t whs[4] = {wh.x, wh.y, wh.z, wh.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint4 dg = decode16(wgs[j], lut);
        const uint4 dh = decode16(whs[j], lut);
        mma_s8_m16n8k32(c0, dg.x, dh.x, dg.y, dh.y, b[j].x, b[j].y);
        mma_s8_m16n8k32(c1, dg.z, dh.z, dg.w, dh.w, b[j].z, b[j].w);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) c[i] = c0[i] + c1[i];
}

// epi: 0 store, 1 residual add (out may alias resid), 2 SwiGLU on interleaved (gate, up) rows
__device__ __forceinline__ void mega_gemv_phase(const MegaParams& p, const MegaSmem& sm, int N, int K, int pro,
                                             const uint2* xin, const float* gamma, int epi, uint2* out,
                                             const uint2* resid, uint32_t ep_out, RingPos& rp, int& tcount, int& pi) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    prof_stamp(p, pi);  // phase start
    const int G = gridDim.x, cta = blockIdx.x;
    const int NT = (N + 15) >> 4, ns = K >> 8, nch = ns > MEGA_SLOT_STEPS ? 2 : 1;
    const int split = nch == 2 ? (ns + 1) >> 1 : ns;  // chunk 0 = steps [0, split), chunk 1 = [split, ns)
    const int q_ld = K + 64;
    int first = cta - tcount;  // tcount < G
    if (first < 0) first += G;
    int nloc = 0;
    for (int tt = first; tt < NT; tt += G) ++nloc;
    const uint32_t NS = (uint32_t)p.n_slots;
    for (int i = tid; i < nloc * 128; i += MEGA_CONS) sm.acc[i] = 0;
    mega_prologue(p, sm, pro, xin, ep_out - 1u, gamma, K, q_ld);  // waits for the previous phase's output; ends with a consumer barrier
    prof_stamp(p, pi);  // prologue done
    // residual rows of this CTA: fetched now, consumed in the epilogue (nloc * 16 <= MEGA_CONS: one row per thread).
    // The old x[row] was validated by this CTA in an earlier prologue and has no writer before this phase's epilogue.
    float rv[MEGA_MAXB];
    if (epi == 1) {
        const int row = (first + (tid >> 4) * G) * 16 + (tid & 15);
#pragma unroll
        for (int m = 0; m < MEGA_MAXB; ++m) {
            rv[m] = 0.f;
            if (m < p.B && tid < nloc * 16 && row < N) rv[m] = ld_cg_f32(reinterpret_cast<const float*>(resid + (size_t)m * N + row));
        }
    }

    // tasks (local tile i, step s) are dealt round-robin to the 16 warps: q = i * ns + s = warp, warp + 16, ...
    const bool tok = g < p.B;
    const int8_t* qrow = reinterpret_cast<const int8_t*>(sm.act) + g * q_ld + 64 * t;
    const uint32_t lut = p.lut;
    const int nchunks = nloc * nch;
    RingPos r = rp;
    int ci = 0, i = 0, s = warp;
    bool waited = false;
    while (s >= ns) {
        s -= ns;
        ++i;
    }
    while (i < nloc) {
        const int ch = s >= split ? 1 : 0;
        const int ct = i * nch + ch;
        while (ci < ct) {  // done with the chunks before this task's: hand their slots back
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.empty[r.slot]);
            ring_advance(r, NS);
            ++ci;
            waited = false;
        }
        if (!waited) {
            mbar_wait(&sm.full[r.slot], r.par);
            waited = true;
        }
        const int s0 = ch ? split : 0, cs = ch ? ns - split : split;
        const int rowb = cs * 64;
        const unsigned char* wr = sm.ring + (size_t)r.slot * MEGA_SLOT_BYTES + (size_t)g * rowb + (s - s0) * 64 + 16 * t;
        int c[4];
        mega_task(wr, rowb, qrow + s * 256, tok, lut, c);
        // lane (g, t): c[0] (row g, token 2t), c[1] (row g, 2t+1), c[2] (row g+8, 2t), c[3] (row g+8, 2t+1)
        int* a = sm.acc + i * 128;
        if (2 * t < p.B) {
            atomicAdd(a + g * 8 + 2 * t, c[0]);
            atomicAdd(a + (g + 8) * 8 + 2 * t, c[2]);
        }
        if (2 * t + 1 < p.B) {
            atomicAdd(a + g * 8 + 2 * t + 1, c[1]);
            atomicAdd(a + (g + 8) * 8 + 2 * t + 1, c[3]);
        }
        s += MEGA_NW;
        while (s >= ns) {
            s -= ns;
            ++i;
        }
    }
    while (ci < nchunks) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[r.slot]);
        ring_advance(r, NS);
        ++ci;
    }
    rp = r;
    cons_sync();
    prof_stamp(p, pi);  // tiles done
    // ---- epilogue ---------------------------------------------------------------------------
    if (epi == 2) {
        const int half = N >> 1;
        for (int m = 0; m < p.B; ++m)
        for (int rr = tid; rr < nloc * 8; rr += MEGA_CONS) {
            const int ii = rr >> 3, j = rr & 7;
            const int row = (first + ii * G) * 16 + 2 * j;
            if (row + 1 < N) {
                const float as = sm.sc_s[m];
                const float v0 = __fmul_rn(as, (float)sm.acc[ii * 128 + (2 * j) * 8 + m]);
                const float v1 = __fmul_rn(as, (float)sm.acc[ii * 128 + (2 * j + 1) * 8 + m]);
                ll_store(out, (size_t)m * half + (row >> 1), __fmul_rn(v0 / (1.0f + expf(-v0)), v1), ep_out);
            }
        }
    } else {
#pragma unroll
        for (int m = 0; m < MEGA_MAXB; ++m) {
            if (m >= p.B) break;
            const int rr = tid;  // nloc <= MEGA_MAX_TILES = 16 tiles -> at most 256 rows per CTA
            if (rr < nloc * 16) {
                const int ii = rr >> 4, rw = rr & 15;
                const int row = (first + ii * G) * 16 + rw;
                if (row < N) {
                    const float v = __fmul_rn(sm.sc_s[m], (float)sm.acc[ii * 128 + rw * 8 + m]);
                    ll_store(out, (size_t)m * N + row, epi == 1 ? __fadd_rn(rv[m], v) : v, ep_out);
                }
            }
        }
    }
    tcount += NT;
    while (tcount >= G) tcount -= G;
    prof_stamp(p, pi);  // epilogue done
    cons_sync();  // no grid barrier follows: the next phase zeroes sm.acc, which slower warps may still be reading here
}

// ---- RoPE + KV append + GQA decode attention for one (head, sequence); consumer threads only ---
// (crates/bitnet-transformer/src/lib.rs:134-184, 374-540, 1171-1202; same arithmetic as
// attn_fused_kernel in decode_fused.cuh)
// Long contexts are split over S CTAs per (head, sequence) -- 128 of the 148 CTAs would otherwise idle through
// this phase while 20 walk the whole cache: split s owns positions [lo, hi) (the last split owns the current
// token), produces an un-normalised partial (max, sum, sum of e * v) as flagged pairs, and split 0 merges the
// S partials (one warp per partial polls it into shared memory, warp 0 combines) into the attention output.
// S = 1 (short contexts) is exactly the single-CTA arithmetic.
constexpr int ATT_SPLIT_MAX = 8;    // partial slots per (head, sequence)
constexpr int ATT_SPLIT_FROM = 48;   // contexts up to this many positions stay on one CTA (the merge hop costs about one pass)
constexpr int ATT_SPLIT_CHUNK = 24;  // target positions per split beyond that; measured 48/24 > 64/32 > 32/16 > 120/60
__device__ __forceinline__ int att_splits(const MegaParams& p, int pos, int G) {
    const int ntask = p.NH * p.B;
    int smax = G / ntask;
    if (smax > ATT_SPLIT_MAX) smax = ATT_SPLIT_MAX;
    if (smax > MEGA_NW) smax = MEGA_NW;
    if (!p.att_split || p.D > 128 || smax < 2 || pos + 1 <= ATT_SPLIT_FROM) return 1;
    const int want = (pos + 1 + ATT_SPLIT_CHUNK - 1) / ATT_SPLIT_CHUNK;
    return want < smax ? want : smax;
}
__device__ __forceinline__ void att_range(int pos, int S, int split, int& lo, int& hi) {
    const int chunk = (pos + 1 + S - 1) / S;
    lo = split * chunk;
    hi = lo + chunk < pos + 1 ? lo + chunk : pos + 1;
    if (lo > hi) lo = hi;
}

__device__ __forceinline__ void mega_attention(const MegaParams& p, const MegaSmem& sm, int head, int b, int pos,
                                               float* kc, float* vc, uint32_t ep_out, int split, int S, int lo, int hi) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int D = p.D, NH = p.NH, NKV = p.NKV, half = D / 2, max_ctx = p.max_ctx;
    float* sc = sm.attn;
    float4* part = reinterpret_cast<float4*>(sm.attn + ((max_ctx + 3) & ~3));
    float* qs = reinterpret_cast<float*>(part + MEGA_CONS);
    float* ks = qs + 256;
    float* vs = ks + 256;
    const int group = NH / NKV, g = head / group;
    const uint2* row = p.qkv + (size_t)b * (NH + 2 * NKV) * D;
    const uint32_t ep_in = ep_out - 1u;
    float* kb = kc + ((size_t)b * NKV + g) * max_ctx * D;
    float* vb = vc + ((size_t)b * NKV + g) * max_ctx * D;
    const bool owner = hi == pos + 1;  // this split holds the current token
    if (tid < half) {
        const float c = __ldg(p.rope_cos + (size_t)pos * half + tid), s = __ldg(p.rope_sin + (size_t)pos * half + tid);
        const uint2* qh = row + (size_t)head * D;
        const uint2* kh = row + (size_t)(NH + g) * D;
        const uint2* vh = row + (size_t)(NH + NKV + g) * D;
        // six probes in flight, then wait for the qkv phase's epoch on each
        const uint2 p0 = ll_ld1(qh + tid), p1 = ll_ld1(qh + tid + half), p2 = ll_ld1(kh + tid), p3 = ll_ld1(kh + tid + half);
        const uint2 p4 = ll_ld1(vh + tid), p5 = ll_ld1(vh + tid + half);
        const float qa = p0.y == ep_in ? __uint_as_float(p0.x) : ll_wait1(qh + tid, ep_in);
        const float qb = p1.y == ep_in ? __uint_as_float(p1.x) : ll_wait1(qh + tid + half, ep_in);
        const float ka = p2.y == ep_in ? __uint_as_float(p2.x) : ll_wait1(kh + tid, ep_in);
        const float kbv = p3.y == ep_in ? __uint_as_float(p3.x) : ll_wait1(kh + tid + half, ep_in);
        qs[tid] = qa * c - qb * s;
        qs[tid + half] = qa * s + qb * c;
        const float k0 = ka * c - kbv * s, k1 = ka * s + kbv * c;
        ks[tid] = k0;
        ks[tid + half] = k1;
        const float v0 = p4.y == ep_in ? __uint_as_float(p4.x) : ll_wait1(vh + tid, ep_in);
        const float v1 = p5.y == ep_in ? __uint_as_float(p5.x) : ll_wait1(vh + tid + half, ep_in);
        vs[tid] = v0;
        vs[tid + half] = v1;
        if (head % group == 0 && owner) {
            kb[(size_t)pos * D + tid] = k0;
            kb[(size_t)pos * D + tid + half] = k1;
            vb[(size_t)pos * D + tid] = v0;
            vb[(size_t)pos * D + tid + half] = v1;
        }
    }
    cons_sync();
    const float scale = 1.0f / sqrtf((float)D);
    float qr[8];  // D <= 256; fixed trip counts keep q in registers and put every K load of a pass in flight at once
    const int per = D / 32;
#pragma unroll
    for (int j = 0; j < 8; ++j) qr[j] = j < per ? qs[lane + 32 * j] : 0.f;
    float mx = -INFINITY;
    constexpr int AU = 4;  // positions per warp (scores) / per thread group (values) in flight per pass
    for (int p0 = lo + warp; p0 < hi; p0 += AU * MEGA_NW) {
        float d[AU];
#pragma unroll
        for (int u = 0; u < AU; ++u) {
            const int pp = p0 + u * MEGA_NW;
            d[u] = 0.f;
            if (pp < hi && pp < pos) {
                const float* kp = kb + (size_t)pp * D;
                float kv[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) kv[j] = j < per ? __ldg(kp + lane + 32 * j) : 0.f;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (j < per) d[u] = fmaf(qr[j], kv[j], d[u]);
            } else if (pp == pos) {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (j < per) d[u] = fmaf(qr[j], ks[lane + 32 * j], d[u]);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int u = 0; u < AU; ++u) d[u] += __shfl_xor_sync(0xffffffffu, d[u], o);
#pragma unroll
        for (int u = 0; u < AU; ++u) {
            const int pp = p0 + u * MEGA_NW;
            if (pp < hi) {
                const float v = d[u] * scale;
                if (lane == 0) sc[pp] = v;
                mx = fmaxf(mx, v);
            }
        }
    }
    if (lane == 0) sm.shf[warp] = mx;
    cons_sync();
    mx = sm.shf[0];
#pragma unroll
    for (int w = 1; w < MEGA_NW; ++w) mx = fmaxf(mx, sm.shf[w]);
    cons_sync();
    float sum = 0.f;
    for (int pp = lo + tid; pp < hi; pp += MEGA_CONS) {
        const float e = expf(sc[pp] - mx);
        sc[pp] = e;
        sum += e;
    }
    sum = warp_sum(sum);
    if (lane == 0) sm.shf[warp] = sum;
    cons_sync();
    sum = 0.f;
#pragma unroll
    for (int w = 0; w < MEGA_NW; ++w) sum += sm.shf[w];
    const float inv = S == 1 ? 1.0f / sum : 1.0f;  // partials stay un-normalised
    const int ncol = D / 4, PG = MEGA_CONS / ncol;
    const int dc = tid % ncol, pg = tid / ncol;
    const int vhi = hi < pos ? hi : pos;  // cached rows of this split
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (pg < PG) {
        for (int p0 = lo + pg; p0 < vhi; p0 += AU * PG) {
            float4 v[AU];
            float w[AU];
#pragma unroll
            for (int u = 0; u < AU; ++u) {
                const int pp = p0 + u * PG;
                v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                w[u] = 0.f;
                if (pp < vhi) {
                    v[u] = __ldg(reinterpret_cast<const float4*>(vb + (size_t)pp * D) + dc);
                    w[u] = sc[pp] * inv;
                }
            }
#pragma unroll
            for (int u = 0; u < AU; ++u) {
                a.x = fmaf(w[u], v[u].x, a.x);
                a.y = fmaf(w[u], v[u].y, a.y);
                a.z = fmaf(w[u], v[u].z, a.z);
                a.w = fmaf(w[u], v[u].w, a.w);
            }
        }
    }
    const float wl = owner ? sc[pos] * inv : 0.f;
    if (pg < PG) part[pg * ncol + dc] = a;
    cons_sync();
    float4 tt = make_float4(0.f, 0.f, 0.f, 0.f);
    if (tid < ncol) {
        for (int i = 0; i < PG; ++i) {
            const float4 v = part[i * ncol + tid];
            tt.x += v.x;
            tt.y += v.y;
            tt.z += v.z;
            tt.w += v.w;
        }
        if (owner) {
            const float4 vl = reinterpret_cast<const float4*>(vs)[tid];
            tt.x = fmaf(wl, vl.x, tt.x);
            tt.y = fmaf(wl, vl.y, tt.y);
            tt.z = fmaf(wl, vl.z, tt.z);
            tt.w = fmaf(wl, vl.w, tt.w);
        }
    }
    uint2* ao = p.att + ((size_t)b * NH + head) * D + 4 * tid;
    if (S == 1) {
        if (tid < ncol) {
            ll_store(ao, 0, tt.x, ep_out);
            ll_store(ao, 1, tt.y, ep_out);
            ll_store(ao, 2, tt.z, ep_out);
            ll_store(ao, 3, tt.w, ep_out);
        }
        cons_sync();  // the attention scratch may be reused by the next (head, sequence) of this CTA
        return;
    }
    // ---- split: publish the partial (split > 0) or merge all of them (split 0) --------------------
    uint2* pbase = p.attp + (((size_t)b * NH + head) * ATT_SPLIT_MAX) * (size_t)(D + 2);
    if (split != 0) {
        uint2* po = pbase + (size_t)split * (D + 2);
        if (tid < ncol) {
            ll_store(po, 4 * tid + 0, tt.x, ep_out);
            ll_store(po, 4 * tid + 1, tt.y, ep_out);
            ll_store(po, 4 * tid + 2, tt.z, ep_out);
            ll_store(po, 4 * tid + 3, tt.w, ep_out);
        }
        if (tid == 0) {
            ll_store(po, D, mx, ep_out);
            ll_store(po, D + 1, sum, ep_out);
        }
        cons_sync();
        return;
    }
    cons_sync();  // part[] (value partials) consumed by warp 0 above: reuse it for the S split partials
    if (warp == 0) {
        if (lane < ncol) part[lane] = tt;
        if (lane == 0) {
            sm.shf[0] = mx;
            sm.shf[16] = sum;
        }
    } else if (warp < S) {  // S <= MEGA_NW: one warp per foreign partial, all polls in flight together
        const uint2* pi = pbase + (size_t)warp * (D + 2);
        if (lane < ncol) {
            const uint4 pa = ll_ld2(pi + 4 * lane), pb = ll_ld2(pi + 4 * lane + 2);
            part[warp * ncol + lane] = ll_wait4(pi + 4 * lane, pa, pb, ep_out);
        }
        if (lane == 0) {
            sm.shf[warp] = ll_wait1(pi + D, ep_out);
            sm.shf[16 + warp] = ll_wait1(pi + D + 1, ep_out);
        }
    }
    cons_sync();
    if (tid < ncol) {
        float M = sm.shf[0];
        for (int s2 = 1; s2 < S; ++s2) M = fmaxf(M, sm.shf[s2]);
        float Lsum = 0.f;
        float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int s2 = 0; s2 < S; ++s2) {
            const float cf = expf(sm.shf[s2] - M);
            const float4 v = part[s2 * ncol + tid];
            Lsum = fmaf(sm.shf[16 + s2], cf, Lsum);
            o.x = fmaf(v.x, cf, o.x);
            o.y = fmaf(v.y, cf, o.y);
            o.z = fmaf(v.z, cf, o.z);
            o.w = fmaf(v.w, cf, o.w);
        }
        const float il = 1.0f / Lsum;
        ll_store(ao, 0, o.x * il, ep_out);
        ll_store(ao, 1, o.y * il, ep_out);
        ll_store(ao, 2, o.z * il, ep_out);
        ll_store(ao, 3, o.w * il, ep_out);
    }
    cons_sync();
}

// rows [r0, r1) of the LM head against NB token rows in shared memory; row j of the CTA goes to warp j % 16
template <int NB>
__device__ __forceinline__ void mega_logits_rows(const MegaParams& p, const MegaSmem& sm, const float* xn, int r0, int r1,
                                                 int RL, RingPos& rp) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int H = p.H, V = p.V, nvec = H >> 8, nrows = r1 - r0;
    const uint32_t NS = (uint32_t)p.n_slots;
    int nchunks = 0;
    for (int rr = 0; rr < nrows; rr += RL) ++nchunks;
    float best[NB];
    int besti[NB];
#pragma unroll
    for (int m = 0; m < NB; ++m) {
        best[m] = -INFINITY;
        besti[m] = 0x7fffffff;
    }
    RingPos r = rp;
    int ci = 0, cbase = 0;  // current chunk and its first row
    bool waited = false;
    for (int j = warp; j < nrows; j += MEGA_NW) {
        while (j >= cbase + RL) {  // done with the chunks before this row's
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.empty[r.slot]);
            ring_advance(r, NS);
            ++ci;
            cbase += RL;
            waited = false;
        }
        if (!waited) {
            mbar_wait(&sm.full[r.slot], r.par);
            waited = true;
        }
        const uint4* wrow = reinterpret_cast<const uint4*>(sm.ring + (size_t)r.slot * MEGA_SLOT_BYTES + (size_t)(j - cbase) * H * 2);
        float acc[NB];
#pragma unroll
        for (int m = 0; m < NB; ++m) acc[m] = 0.f;
#pragma unroll 2
        for (int i = 0; i < nvec; ++i) {
            const uint4 wv = wrow[lane + 32 * i];
            const __half2* hp = reinterpret_cast<const __half2*>(&wv);
            float f[8];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float2 tt = __half22float2(hp[q]);
                f[2 * q] = tt.x;
                f[2 * q + 1] = tt.y;
            }
#pragma unroll
            for (int m = 0; m < NB; ++m) {
                const float4 xa = *reinterpret_cast<const float4*>(xn + m * H + 8 * (lane + 32 * i));
                const float4 xb = *reinterpret_cast<const float4*>(xn + m * H + 8 * (lane + 32 * i) + 4);
                float a = acc[m];
                a = fmaf(f[0], xa.x, a);
                a = fmaf(f[1], xa.y, a);
                a = fmaf(f[2], xa.z, a);
                a = fmaf(f[3], xa.w, a);
                a = fmaf(f[4], xb.x, a);
                a = fmaf(f[5], xb.y, a);
                a = fmaf(f[6], xb.z, a);
                a = fmaf(f[7], xb.w, a);
                acc[m] = a;
            }
        }
        const int row = r0 + j;
#pragma unroll
        for (int m = 0; m < NB; ++m) {
            const float tot = warp_sum(acc[m]);
            if (lane == 0) p.logits[(size_t)m * V + row] = tot;
            if (tot > best[m]) {  // a warp sees its rows in increasing order
                best[m] = tot;
                besti[m] = row;
            }
        }
    }
    while (ci < nchunks) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[r.slot]);
        ring_advance(r, NS);
        ++ci;
    }
    rp = r;
    if (lane == 0) {
#pragma unroll
        for (int m = 0; m < NB; ++m) {
            sm.bestv[warp * MEGA_MAXB + m] = best[m];
            sm.besti[warp * MEGA_MAXB + m] = besti[m];
        }
    }
}

// ---- LM head: final RMSNorm, tied-embedding logits (f16 table through the ring), argmax partials
__device__ __forceinline__ void mega_logits_phase(const MegaParams& p, const MegaSmem& sm, RingPos& rp, uint32_t ep_in) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int H = p.H, V = p.V, G = gridDim.x, cta = blockIdx.x;
    float* xn = reinterpret_cast<float*>(sm.act);  // [B][H]
    // final RMSNorm (cpu::rms_norm) of every token row into shared memory
    for (int m = 0; m < p.B; ++m) {
        double ss = 0.0;
        for (int k = tid; k < H; k += MEGA_CONS) {
            const float v = ll_wait1(p.x + (size_t)m * H + k, ep_in);
            xn[m * H + k] = v;
            ss += (double)v * (double)v;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        if (lane == 0) sm.shd[warp] = ss;
        cons_sync();
        double tot = 0.0;
#pragma unroll
        for (int wv = 0; wv < MEGA_NW; ++wv) tot += sm.shd[wv];
        const float ms = (float)(tot / (double)H);
        const float inv_rms = 1.0f / sqrtf(ms + p.eps);
        for (int k = tid; k < H; k += MEGA_CONS) xn[m * H + k] = xn[m * H + k] * inv_rms * __ldg(p.final_norm + k);
        cons_sync();
    }
    const int rpc = (V + G - 1) / G;
    const int r0 = min(V, cta * rpc), r1 = min(V, r0 + rpc);
    const int RL = MEGA_SLOT_BYTES / (H * 2);
    switch (p.B) {
        case 1: mega_logits_rows<1>(p, sm, xn, r0, r1, RL, rp); break;
        case 2: mega_logits_rows<2>(p, sm, xn, r0, r1, RL, rp); break;
        case 3: mega_logits_rows<3>(p, sm, xn, r0, r1, RL, rp); break;
        default: mega_logits_rows<4>(p, sm, xn, r0, r1, RL, rp); break;
    }
    cons_sync();
    if (tid < p.B) {
        float bv = -INFINITY;
        int bi = 0x7fffffff;
        for (int w = 0; w < MEGA_NW; ++w) {
            const float cv = sm.bestv[w * MEGA_MAXB + tid];
            const int ci = sm.besti[w * MEGA_MAXB + tid];
            if (cv > bv || (cv == bv && ci < bi)) {
                bv = cv;
                bi = ci;
            }
        }
        p.part_val[(size_t)tid * G + cta] = bv;
        p.part_idx[(size_t)tid * G + cta] = bi;
    }
}

// ---- the producer warp: every weight byte of the step, in consumption order ----------------------
__device__ __forceinline__ void mega_producer(const MegaParams& p, const MegaSmem& sm) {
    const int lane = threadIdx.x & 31;
    const int G = gridDim.x, cta = blockIdx.x;
    uint32_t n = 0;
    int tcount = 0;
    const uint32_t NS = (uint32_t)p.n_slots;
    for (int l = 0; l < p.L; ++l) {
        const MegaLayer& L = sm.layers[l];
        for (int op = 0; op < 4; ++op) {
            const uint8_t* w = op == 0 ? L.wqkv : (op == 1 ? L.wo : (op == 2 ? L.wgu : L.wdown));
            const int N = op == 0 ? (p.NH + 2 * p.NKV) * p.D : (op == 1 ? p.H : (op == 2 ? 2 * p.F : p.H));
            const int K = op == 1 ? p.NH * p.D : (op == 3 ? p.F : p.H);
            const int NT = (N + 15) >> 4, ns = K >> 8, nch = (ns + MEGA_SLOT_STEPS - 1) / MEGA_SLOT_STEPS;
            const size_t stride = (size_t)K >> 2;
            for (int tile = (cta - tcount + G) % G; tile < NT; tile += G) {
                const int nr = min(16, N - tile * 16);
                for (int ch = 0; ch < nch; ++ch, ++n) {
                    int s0, cs;
                    chunk_range(ns, nch, ch, s0, cs);
                    const uint32_t slot = n % NS;
                    unsigned char* dst = sm.ring + (size_t)slot * MEGA_SLOT_BYTES;
                    if (lane == 0) {
                        if (n >= NS) mbar_wait_backoff(&sm.empty[slot], ((n / NS) - 1u) & 1u);
                        mbar_expect_tx(&sm.full[slot], (uint32_t)(nr * cs * 64));
                    }
                    __syncwarp();
                    if (nch == 1) {
                        if (lane == 0) bulk_g2s(dst, w + (size_t)tile * 16 * stride, (uint32_t)(nr * cs * 64), &sm.full[slot]);
                    } else if (lane < nr) {
                        bulk_g2s(dst + (size_t)lane * cs * 64, w + ((size_t)tile * 16 + lane) * stride + (size_t)s0 * 64,
                                 (uint32_t)(cs * 64), &sm.full[slot]);
                    }
                }
            }
            tcount = (tcount + NT) % G;
        }
    }
    if (p.mode == 0) {
        const int rpc = (p.V + G - 1) / G;
        const int r0 = min(p.V, cta * rpc), r1 = min(p.V, r0 + rpc);
        const int RL = MEGA_SLOT_BYTES / (p.H * 2);
        for (int rr = r0; rr < r1; rr += RL, ++n) {
            const int nr = min(RL, r1 - rr);
            const uint32_t slot = n % NS;
            if (lane == 0) {
                if (n >= NS) mbar_wait_backoff(&sm.empty[slot], ((n / NS) - 1u) & 1u);
                const uint32_t bytes = (uint32_t)nr * (uint32_t)p.H * 2u;
                mbar_expect_tx(&sm.full[slot], bytes);
                bulk_g2s(sm.ring + (size_t)slot * MEGA_SLOT_BYTES, p.emb + (size_t)rr * p.H, bytes, &sm.full[slot]);
            }
        }
    }
}

__global__ void __launch_bounds__(MEGA_THREADS, 1) decode_mega_kernel(const __grid_constant__ MegaParams p) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5;
    const int G = gridDim.x, cta = blockIdx.x;
    // ---- shared memory carve-up (must match mega_smem_bytes) -----------------------------------
    MegaSmem sm;
    size_t off = 0;
    sm.full = reinterpret_cast<uint64_t*>(smem_raw + off);
    off += MEGA_MAX_SLOTS * 8;
    sm.empty = reinterpret_cast<uint64_t*>(smem_raw + off);
    off += MEGA_MAX_SLOTS * 8;
    sm.shd = reinterpret_cast<double*>(smem_raw + off);
    off += 32 * 8;
    sm.shf = reinterpret_cast<float*>(smem_raw + off);
    off += 32 * 4;
    sm.sc_s = reinterpret_cast<float*>(smem_raw + off);
    off += 32;
    sm.attw = reinterpret_cast<int*>(smem_raw + off);
    off += 32;
    sm.bestv = reinterpret_cast<float*>(smem_raw + off);
    off += MEGA_NW * MEGA_MAXB * 4;
    sm.besti = reinterpret_cast<int*>(smem_raw + off);
    off += MEGA_NW * MEGA_MAXB * 4;
    sm.layers = reinterpret_cast<MegaLayer*>(smem_raw + off);
    off += (size_t)MEGA_MAX_LAYERS * sizeof(MegaLayer);
    sm.acc = reinterpret_cast<int*>(smem_raw + off);
    off += MEGA_MAX_TILES * 128 * 4;
    off = (off + 127) & ~(size_t)127;
    sm.act = smem_raw + off;
    off += (size_t)p.act_bytes;
    off = (off + 127) & ~(size_t)127;
    sm.attn = reinterpret_cast<float*>(smem_raw + off);
    off += (size_t)p.attn_floats * 4;
    off = (off + 127) & ~(size_t)127;
    sm.ring = smem_raw + off;

    if (tid == 0) {
        for (int s = 0; s < p.n_slots; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], MEGA_NW);
        }
        mbar_fence_init();
        if (cta == 0)  // the counters of the NEXT launch (their last user has finished)
            for (int j = 0; j < MEGA_BAR_CTRS; ++j) p.bar[((p.parity ^ 1) * MEGA_BAR_CTRS + j) * MEGA_BAR_STRIDE] = 0u;
    }
    for (int i = tid; i < p.L * (int)(sizeof(MegaLayer) / 8); i += MEGA_THREADS)
        reinterpret_cast<uint64_t*>(sm.layers)[i] = reinterpret_cast<const uint64_t*>(p.layers)[i];
    __syncthreads();  // the only CTA-wide barrier: from here on the producer warp runs free

    if (warp == MEGA_NW) {
        mega_producer(p, sm);
        return;
    }

    // ---- consumers ------------------------------------------------------------------------------
    unsigned* ctr = p.bar + p.parity * (MEGA_BAR_CTRS * MEGA_BAR_STRIDE);
    unsigned target = 0;
    RingPos rp{0u, 0u};
    int tcount = 0, pi = 0;
    const int H = p.H, F = p.F, D = p.D, NH = p.NH, NKV = p.NKV, B = p.B;
    const int QD = NH * D, QKV = (NH + 2 * NKV) * D;
    const bool full = p.mode == 0;  // mode 1 = roofline leg: the 4L ternary GEMVs only, fixed int8 inputs, no barriers
    int pos = 0;
    if (full) {
        pos = ld_cg_s32(p.pos);
        // embed: x[b] = f32(emb[token[b]]), spread over the grid
        for (int i = cta * MEGA_CONS + tid; i < B * H; i += G * MEGA_CONS) {
            const int b = i / H, k = i - b * H;
            ll_store(p.x, i, __half2float(p.emb[(size_t)ld_cg_s32(p.tokens + b) * H + k]), p.epoch);
        }
        if (p.prefetch) prefetch_l2_lines(sm.layers[0].attn_norm, (size_t)H * 4);
        if (p.barriers) grid_sync(ctr, target, G);
    }
    // attention work item of this CTA, fixed for the whole step: computed once (divisions) and parked in shared
    // memory so that it does not occupy registers across the tile loops
    const int attS = full ? att_splits(p, pos, G) : 1;
    if (full && tid == 0) {
        const bool on = cta < NH * B * attS;
        int lo = 0, hi = 0, head = 0, bb = 0, sp = 0;
        unsigned long long kv = 0;
        if (on) {
            const int task = cta % (NH * B);
            sp = cta / (NH * B);
            head = task % NH;
            bb = task / NH;
            att_range(pos, attS, sp, lo, hi);
            kv = ((unsigned long long)bb * NKV + head / (NH / NKV)) * p.max_ctx * D + (unsigned long long)lo * D;
        }
        sm.attw[0] = on ? 1 : 0;
        sm.attw[1] = head;
        sm.attw[2] = bb;
        sm.attw[3] = sp;
        sm.attw[4] = lo;
        sm.attw[5] = hi;
        sm.attw[6] = (int)(unsigned)(kv & 0xffffffffull);
        sm.attw[7] = (int)(unsigned)(kv >> 32);
    }
    if (full) cons_sync();
    // ONE call site per phase kind (the bodies are inlined once): op 0 qkv, 1 attention, 2 o, 3 gate/up, 4 down
    for (int l = 0; l < p.L; ++l) {
        const MegaLayer& L = sm.layers[l];
        for (int op = 0; op < 5; ++op) {
            const uint32_t ep = p.epoch + 1u + (uint32_t)(5 * l + op);  // epoch of this phase's output; its input carries ep - 1
            if (op == 1) {
                if (!full) continue;
                const size_t kvoff = (size_t)l * B * NKV * p.max_ctx * D;
                prof_stamp(p, pi);  // attention start
                // work item = (split, head, sequence), one per CTA (mega_plan: NH * B <= SMs); split 0 merges
                if (sm.attw[0])
                    mega_attention(p, sm, sm.attw[1], sm.attw[2], pos, p.kc + kvoff, p.vc + kvoff, ep, sm.attw[3], attS, sm.attw[4], sm.attw[5]);
                prof_stamp(p, pi);  // attention done
                if (p.barriers) grid_sync(ctr, target, G);
                continue;
            }
            const int N = op == 0 ? QKV : (op == 3 ? 2 * F : H);
            const int K = op == 2 ? QD : (op == 4 ? F : H);
            const int pro = !full ? 0 : ((op == 0 || op == 3) ? 2 : 1);
            const uint2* xin = op == 2 ? p.att : (op == 4 ? p.gu : p.x);
            const float* gamma = op == 0 ? L.attn_norm : L.ffn_norm;
            const int epi = op == 3 ? 2 : ((full && (op == 2 || op == 4)) ? 1 : 0);
            uint2* out = op == 0 ? p.qkv : (op == 3 ? p.gu : p.x);
            mega_gemv_phase(p, sm, N, K, pro, xin, gamma, epi, out, p.x, ep, rp, tcount, pi);
            if (full) {
                // what the next phase touches first and the weight stream has pushed out of L2 since the last step:
                // the miss overlaps the wait for the other CTAs' outputs
                if (!p.prefetch) {
                } else if (op == 0) {
                    const int att_lo = sm.attw[4], att_hi = sm.attw[5];
                    if (sm.attw[0] && att_hi > att_lo) {  // the cached rows this CTA's attention item reads
                        const size_t kv = (size_t)l * B * NKV * p.max_ctx * D +
                                          (((size_t)(unsigned)sm.attw[7] << 32) | (size_t)(unsigned)sm.attw[6]);
                        const size_t nb = (size_t)((att_hi < pos ? att_hi : pos) - att_lo) * D * 4;
                        prefetch_l2_lines(p.kc + kv, nb);
                        prefetch_l2_lines(p.vc + kv, nb);
                    }
                    if (cta < NH * B && l == 0) {
                        prefetch_l2_lines(p.rope_cos + (size_t)pos * (D / 2), (size_t)D * 2);
                        prefetch_l2_lines(p.rope_sin + (size_t)pos * (D / 2), (size_t)D * 2);
                    }
                } else if (op == 2) {
                    prefetch_l2_lines(L.ffn_norm, (size_t)H * 4);
                } else if (op == 4) {
                    prefetch_l2_lines(l + 1 < p.L ? sm.layers[l + 1].attn_norm : p.final_norm, (size_t)H * 4);
                }
                if (p.barriers) grid_sync(ctr, target, G);
            }
        }
    }
    if (!full) return;
    prof_stamp(p, pi);
    mega_logits_phase(p, sm, rp, p.epoch + (uint32_t)(5 * p.L));
    prof_stamp(p, pi);
    grid_sync(ctr, target, G);
    prof_stamp(p, pi);
    // greedy argmax over the per-CTA partials (lowest index on ties); feed the token back, advance
    if (cta == 0 && warp < B) {
        const int b = warp, lane = tid & 31;
        float best = -INFINITY;
        int idx = 0x7fffffff;
        for (int i = lane; i < G; i += 32) {
            const float v = ld_cg_f32(p.part_val + (size_t)b * G + i);
            const int vi = ld_cg_s32(p.part_idx + (size_t)b * G + i);
            if (v > best || (v == best && vi < idx)) {
                best = v;
                idx = vi;
            }
        }
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
            if (ov > best || (ov == best && oi < idx)) {
                best = ov;
                idx = oi;
            }
        }
        if (lane == 0) {
            if (idx == 0x7fffffff) idx = 0;
            p.next[b] = idx;
            p.tokens[b] = idx;
            if (b == 0) *p.pos = pos + 1;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static size_t mega_fixed_smem(int act_bytes, int attn_floats) {
    size_t off = MEGA_MAX_SLOTS * 16 + 32 * 8 + 32 * 4 + 32 + 32 + 2 * MEGA_NW * MEGA_MAXB * 4 +
                 (size_t)MEGA_MAX_LAYERS * sizeof(MegaLayer) + MEGA_MAX_TILES * 128 * 4;
    off = (off + 127) & ~(size_t)127;
    off += (size_t)act_bytes;
    off = (off + 127) & ~(size_t)127;
    off += (size_t)attn_floats * 4;
    off = (off + 127) & ~(size_t)127;
    return off;
}

// Fills the shape-derived fields of `p` (n_slots, act_bytes, attn_floats) and returns the dynamic
// shared memory the launch needs, or 0 when the shape is not covered by this kernel.
size_t mega_plan(MegaParams* p, int num_sms) {
    const int QD = p->NH * p->D;
    if (p->B < 1 || p->B > MEGA_MAXB || p->L > MEGA_MAX_LAYERS) return 0;
    if (p->NH * p->B > num_sms) return 0;                        // one attention work item per CTA
    if (p->H % 256 || p->F % 256 || QD % 256) return 0;          // whole 256-column steps, no tails
    if (p->H > 16 * MEGA_CONS || p->F > 16 * MEGA_CONS || QD > 16 * MEGA_CONS) return 0;  // one prologue pass per row (NV = 4)
    if (p->D % 32 || p->D > 256 || MEGA_CONS % (p->D / 4)) return 0;
    if (p->H * 2 > MEGA_SLOT_BYTES) return 0;                    // one LM-head row per slot at least
    if (((p->NH + 2 * p->NKV) * p->D) % 2 || (2 * p->F) % 16) return 0;
    const int kmax = p->H > p->F ? (p->H > QD ? p->H : QD) : (p->F > QD ? p->F : QD);
    const int tiles_max = ((2 * p->F + 15) / 16 + num_sms - 1) / num_sms + 1;
    if (tiles_max > MEGA_MAX_TILES) return 0;
    int act = p->B * (kmax + 64);
    if (p->B * p->H * 4 > act) act = p->B * p->H * 4;
    p->act_bytes = (act + 15) & ~15;
    p->attn_floats = ((p->max_ctx + 3) & ~3) + MEGA_CONS * 4 + 3 * 256 + 16;
    const size_t fixed = mega_fixed_smem(p->act_bytes, p->attn_floats);
    const size_t cap = 227 * 1024;
    if (fixed + 3 * (size_t)MEGA_SLOT_BYTES > cap) return 0;
    int slots = (int)((cap - fixed) / MEGA_SLOT_BYTES);
    if (slots > MEGA_MAX_SLOTS) slots = MEGA_MAX_SLOTS;
    p->n_slots = slots;
    return fixed + (size_t)slots * MEGA_SLOT_BYTES;
}

int launch_decode_mega(const MegaParams& p, size_t smem, int num_sms, cudaStream_t st) {
    static bool attr_set[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 16 && !attr_set[dev]) {
        B200_CUDA(cudaFuncSetAttribute(decode_mega_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set[dev] = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)num_sms);
    cfg.blockDim = dim3(MEGA_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;  // all CTAs resident: the grid barrier cannot deadlock
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    B200_CUDA(cudaLaunchKernelEx(&cfg, decode_mega_kernel, p));
    count_launch();
    return 0;
}

}  // namespace b200
