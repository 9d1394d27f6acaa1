// decode.cu -- a BitNet-b1.58-2B-4T-shaped decoder step on device, as ONE CUDA graph.
//
// The ternary linears and the act-quant that feeds them are the hot path (gemv.cu, quant.cu);
// this file adds the glue the reference runs through candle on the host
// (crates/bitnet-transformer/src/lib.rs): embedding gather (:1390-1426), split-half RoPE
// (:134-184), KV-cache append (:1171-1202), GQA decode attention (:374-540), SwiGLU glue
// (:756-789, fused into the act-quant), tied-embedding logits (:1599-1630) and greedy argmax.
// Norms are true RMSNorm (cpu::rms_norm), as SURVEY.md §3.1 prescribes for this build.
//
// The whole step (about 10 launches per layer) is captured once into a CUDA graph; the
// position lives in device memory so the same graph replays for every token, and the greedy
// token is fed back on device so a decode loop needs no host round trip.
#include <cstdio>
#include <cstring>
#include <memory>
#include <vector>

#include "common.cuh"
#include "decode_fused.cuh"
#include "decode_mega.h"

namespace b200 {

// gemv_stream.cu: 1 = shape unsupported (use the generic kernels), 0 = ok, -1 = error
int launch_gemv_fast(const bitnet_b200_ctx*, const bitnet_b200_weight*, int M, int pro, const int8_t* q,
                       const float* act_scale, const float* x, const float* gamma, float eps, int epi, float* y,
                       const float* residual, int32_t* acc_out, int8_t* q_out, float* scale_out, float* normed_out,
                       bool pdl);
bool gemv_fast_supported(const bitnet_b200_ctx*, const bitnet_b200_weight*, int M, bool f32_prologue);
enum { PRO_Q8 = 0, PRO_F32 = 1, PRO_RMS = 2 };
enum { EPI_STORE = 0, EPI_RESIDUAL = 1, EPI_SWIGLU = 2 };

int launch_gemv_i8(const bitnet_b200_ctx*, const bitnet_b200_weight*, const int8_t*, const float*,
                   int, float*, int32_t*, const float* residual);
int launch_actquant(const bitnet_b200_ctx*, const float*, const float*, float, int, int, int8_t*,
                    float*, float*, int mode);

// ---------------------------------------------------------------------------------------
__global__ void embed_kernel(const __half* __restrict__ emb, const int* __restrict__ tokens,
                             float* __restrict__ x, int H) {
    const int b = blockIdx.x;
    const __half* row = emb + (size_t)tokens[b] * H;
    for (int i = threadIdx.x; i < H; i += blockDim.x) x[(size_t)b * H + i] = __half2float(row[i]);
}

// RoPE (split-half pairs (i, i + D/2)) on q in place and on k into the cache; v into the cache.
// grid (NH + 2*NKV, B), block D/2. qkv row = [q (NH*D) | k (NKV*D) | v (NKV*D)].
__global__ void rope_kv_kernel(float* __restrict__ qkv, const float* __restrict__ cos_t,
                               const float* __restrict__ sin_t, float* __restrict__ kc,
                               float* __restrict__ vc, const int* __restrict__ pos_p, int NH,
                               int NKV, int D, int max_ctx) {
    const int head = blockIdx.x, b = blockIdx.y, i = threadIdx.x, half = D / 2;
    const int pos = *pos_p;
    float* row = qkv + (size_t)b * (NH + 2 * NKV) * D;
    if (head < NH + NKV) {
        float* v = row + (size_t)head * D;
        const float c = cos_t[(size_t)pos * half + i], s = sin_t[(size_t)pos * half + i];
        const float a = v[i], bb = v[i + half];
        const float r0 = a * c - bb * s, r1 = a * s + bb * c;
        if (head < NH) {
            v[i] = r0;
            v[i + half] = r1;
        } else {
            float* dst = kc + (((size_t)b * NKV + (head - NH)) * max_ctx + pos) * D;
            dst[i] = r0;
            dst[i + half] = r1;
        }
    } else {
        const int g = head - NH - NKV;
        const float* v = row + (size_t)(NH + NKV + g) * D;
        float* dst = vc + (((size_t)b * NKV + g) * max_ctx + pos) * D;
        dst[i] = v[i];
        dst[i + half] = v[i + half];
    }
}

// Decode attention for one (head, sequence): scores over 0..pos, softmax, weighted V sum.
// grid (NH, B), block D threads (D = 128 -> 4 warps), dynamic smem = max_ctx floats.
__global__ void attn_decode_kernel(const float* __restrict__ qkv, const float* __restrict__ kc,
                                   const float* __restrict__ vc, float* __restrict__ att,
                                   const int* __restrict__ pos_p, int NH, int NKV, int D,
                                   int max_ctx) {
    extern __shared__ float sc[];
    __shared__ float red[32];
    const int head = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int pos = *pos_p, g = head / (NH / NKV);
    const float* q = qkv + (size_t)b * (NH + 2 * NKV) * D + (size_t)head * D;
    const float* kb = kc + ((size_t)b * NKV + g) * max_ctx * D;
    const float* vb = vc + ((size_t)b * NKV + g) * max_ctx * D;
    const float scale = 1.0f / sqrtf((float)D);
    // each lane holds D/32 query elements (strided so that K loads coalesce)
    float qr[8];
    const int per = D / 32;
    for (int j = 0; j < per; ++j) qr[j] = q[lane + 32 * j];
    float mx = -INFINITY;
    for (int p = warp; p <= pos; p += nwarps) {
        const float* kp = kb + (size_t)p * D;
        float d = 0.f;
        for (int j = 0; j < per; ++j) d = fmaf(qr[j], kp[lane + 32 * j], d);
        d = warp_sum(d) * scale;
        if (lane == 0) sc[p] = d;
        mx = fmaxf(mx, d);
    }
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    mx = red[0];
    for (int w = 1; w < nwarps; ++w) mx = fmaxf(mx, red[w]);
    __syncthreads();
    float sum = 0.f;
    for (int p = tid; p <= pos; p += blockDim.x) {
        const float e = expf(sc[p] - mx);
        sc[p] = e;
        sum += e;
    }
    sum = warp_sum(sum);
    if (lane == 0) red[warp] = sum;
    __syncthreads();
    sum = 0.f;
    for (int w = 0; w < nwarps; ++w) sum += red[w];
    const float inv = 1.0f / sum;
    for (int d = tid; d < D; d += blockDim.x) {
        float a = 0.f;
        for (int p = 0; p <= pos; ++p) a = fmaf(sc[p] * inv, vb[(size_t)p * D + d], a);
        att[((size_t)b * NH + head) * D + d] = a;
    }
}

// Tied-embedding logits: f16 table [V, H] x f32 hidden [B, H] -> f32 [B, V].
// Thread t owns the 8 consecutive k = 8t..8t+7 of every row (one 16-byte load per row),
// R rows in flight; same transposing reduction as the ternary GEMV.
template <int MT, int R>
__global__ void __launch_bounds__(1024) logits_f16_kernel(const __half* __restrict__ emb,
                                                          const float* __restrict__ xn,
                                                          float* __restrict__ logits, int V, int H,
                                                          int M, int rows_per_cta) {
    constexpr int NV = R * MT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* red = reinterpret_cast<float*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int nvec = H / 8;
    const bool active = tid < nvec;
    float xv[MT][8];
#pragma unroll
    for (int m = 0; m < MT; ++m)
#pragma unroll
        for (int j = 0; j < 8; ++j) xv[m][j] = (active && m < M) ? xn[(size_t)m * H + tid * 8 + j] : 0.f;
    const int row_begin = blockIdx.x * rows_per_cta;
    const int row_end = min(V, row_begin + rows_per_cta);
    int buf = 0;
    for (int r0 = row_begin; r0 < row_end; r0 += R, buf ^= 1) {
        uint4 w[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int row = r0 + r;
            w[r] = (active && row < row_end) ? ldg_nc_u128(emb + (size_t)row * H + (size_t)tid * 8)
                                             : make_uint4(0, 0, 0, 0);
        }
        float v[NV];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const __half2* hp = reinterpret_cast<const __half2*>(&w[r]);
            float f[8];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float2 t = __half22float2(hp[j]);
                f[2 * j] = t.x;
                f[2 * j + 1] = t.y;
            }
#pragma unroll
            for (int m = 0; m < MT; ++m) {
                float a = 0.f;
#pragma unroll
                for (int j = 0; j < 8; ++j) a = fmaf(f[j], xv[m][j], a);
                v[r * MT + m] = a;
            }
        }
        warp_reduce_transpose<NV, float>(v);
        constexpr int SHIFT = 5 - Log2<NV>::value;
        if ((lane & ((1 << SHIFT) - 1)) == 0) red[(buf * nwarps + warp) * NV + (lane >> SHIFT)] = v[0];
        __syncthreads();
        if (tid < NV) {
            float tot = 0.f;
            for (int wv = 0; wv < nwarps; ++wv) tot += red[(buf * nwarps + wv) * NV + tid];
            const int r = tid / MT, m = tid % MT, row = r0 + r;
            if (row < row_end && m < M) logits[(size_t)m * V + row] = tot;
        }
    }
}

// greedy argmax per sequence (lowest index on ties); feeds the token back and advances pos
__global__ void __launch_bounds__(1024) argmax_kernel(const float* __restrict__ logits, int V,
                                                      int* __restrict__ next, int* __restrict__ tokens,
                                                      int* __restrict__ pos_p, int advance) {
    __shared__ float bv[32];
    __shared__ int bi[32];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* lg = logits + (size_t)b * V;
    float best = -INFINITY;
    int idx = 0x7fffffff;
    for (int i = tid; i < V; i += blockDim.x) {
        const float v = lg[i];
        if (v > best || (v == best && i < idx)) {
            best = v;
            idx = i;
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
        bv[warp] = best;
        bi[warp] = idx;
    }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w)
            if (bv[w] > best || (bv[w] == best && bi[w] < idx)) {
                best = bv[w];
                idx = bi[w];
            }
        if (idx == 0x7fffffff) idx = 0;
        next[b] = idx;
        tokens[b] = idx;
        if (advance && b == 0) *pos_p = *pos_p + 1;
    }
}

}  // namespace b200

using namespace b200;

// ---------------------------------------------------------------------------------------
// the model object
// ---------------------------------------------------------------------------------------
struct LayerW {
    bitnet_b200_weight *wqkv = nullptr, *wo = nullptr, *wgateup = nullptr, *wdown = nullptr;
    float *attn_norm = nullptr, *ffn_norm = nullptr;
};

struct bitnet_b200_model {
    bitnet_b200_ctx* ctx = nullptr;
    bitnet_b200_model_config cfg{};
    std::vector<LayerW> layers;
    __half* emb = nullptr;
    float *final_norm = nullptr, *rope_cos = nullptr, *rope_sin = nullptr;
    float *kc = nullptr, *vc = nullptr;  // [L][B][NKV][max_ctx][D]
    int *pos = nullptr, *tokens = nullptr, *next = nullptr;
    float *x = nullptr, *qkv = nullptr, *att = nullptr, *gu = nullptr, *xn = nullptr, *logits = nullptr;
    float* part_val = nullptr;  // [B, logits CTAs] per-CTA argmax partials (fused path)
    int* part_idx = nullptr;
    int fused = -1;             // -1 undecided, 0 generic kernels, 1 fused PDL path
    int mega = -1;              // -1 undecided, 0 no, 1 the persistent one-kernel step (decode_mega.cu)
    MegaLayer* mega_layers = nullptr;  // device table
    std::vector<uint8_t*> mega_codes;  // fragment-major copies of the code planes (4 per layer), owned
    unsigned* mega_bar = nullptr;      // [2][8] grid-barrier counters, one per 256-byte line
    int mega_parity = 0;
    uint2* mega_ll = nullptr;          // (value, epoch) pair buffers of the phase outputs: x | qkv | att | gu
    size_t mega_ll_pairs = 0;
    uint32_t mega_epoch = 1;           // 0 is what freshly cleared pair buffers hold: never a valid epoch
    size_t mega_smem = 0;
    MegaParams mega_p{};
    int8_t* q8 = nullptr;
    float* s8 = nullptr;
    int* next_host = nullptr;  // pinned
    int* tok_host = nullptr;   // pinned
    cudaGraphExec_t graph = nullptr;
    uint64_t launches_per_step = 0;
    bool have_emb = false, have_rope = false;
    int layers_set = 0;
    int host_pos = 0;  // host mirror of the device-side position: zeroed by reset, advanced per launched step
};

static int dmalloc(void** p, size_t n) {
    B200_CUDA(cudaMalloc(p, n ? n : 1));
    return 0;
}

template <typename T>
static int upload(T** dst, const T* src, size_t n, cudaStream_t st) {
    if (!*dst && dmalloc(reinterpret_cast<void**>(dst), n * sizeof(T))) return -1;
    B200_CUDA(cudaMemcpyAsync(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice, st));
    B200_CUDA(cudaStreamSynchronize(st));
    return 0;
}

static size_t row_bytes(int layout, size_t cols) {
    return layout == BITNET_B200_LAYOUT_QK256 ? div_ceil(cols, 256) * 64 : cols / 32 * 10;
}

// Swaps the context onto a private capturable stream for the lifetime of a graph capture.
struct CaptureStream {
    bitnet_b200_ctx* ctx;
    cudaStream_t s = nullptr, saved = nullptr;
    bool swapped = false;
    explicit CaptureStream(bitnet_b200_ctx* c) : ctx(c) {
        if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) {
            fail("cudaStreamCreateWithFlags failed for graph capture");
            s = nullptr;
            return;
        }
        saved = ctx->stream;
        ctx->stream = s;
        swapped = true;
    }
    bool ok() const { return s != nullptr; }
    void restore() {
        if (swapped) ctx->stream = saved;
        swapped = false;
    }
    ~CaptureStream() {
        restore();
        if (s) cudaStreamDestroy(s);
    }
};

constexpr size_t LOGITS_ROWS_PER_CTA = 64;
// scores [max_ctx] + value partials [ATTN_THREADS / (D/4)][D] = ATTN_THREADS float4
static size_t attn_fused_smem(int max_ctx) { return (((size_t)max_ctx + 3) & ~(size_t)3) * 4 + (size_t)ATTN_THREADS * 16; }

// BITNET_B200_NO_PDL=1 launches the same kernels with ordinary stream serialisation
static bool use_pdl() {
    static const bool v = getenv("BITNET_B200_NO_PDL") == nullptr;
    return v;
}

template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = use_pdl() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

static bool fused_supported(const bitnet_b200_model* m) {
    const auto& c = m->cfg;
    if (c.batch > 8 || c.hidden % 8 || c.hidden / 8 > 1024 || c.head_dim > 256) return false;
    if (getenv("BITNET_B200_NO_FUSED")) return false;
    const LayerW& L = m->layers[0];
    return gemv_fast_supported(m->ctx, L.wqkv, c.batch, true) && gemv_fast_supported(m->ctx, L.wo, c.batch, true) &&
           gemv_fast_supported(m->ctx, L.wgateup, c.batch, true) && gemv_fast_supported(m->ctx, L.wdown, c.batch, true);
}

// The fused step: per layer  [RMSNorm+quant | qkv GEMV] -> [RoPE+KV append+attention] ->
// [quant | o GEMV + residual] -> [RMSNorm+quant | gate/up GEMV | SwiGLU] -> [quant | down GEMV +
// residual]; then [final RMSNorm | logits | argmax partials] -> [argmax]. 5 L + 3 launches.
static int enqueue_step_fused(bitnet_b200_model* m) {
    const auto& c = m->cfg;
    bitnet_b200_ctx* ctx = m->ctx;
    cudaStream_t st = ctx->stream;
    const int B = c.batch, H = c.hidden, D = c.head_dim, NH = c.n_heads, NKV = c.n_kv_heads;
    const uint64_t l0 = g_launches.load();
    // timing diagnostics only (results are wrong when set): bit 0 attention, 1 logits, 2 qkv, 3 o, 4 gate/up, 5 down
    const int skip = getenv("BITNET_B200_DIAG_SKIP") ? atoi(getenv("BITNET_B200_DIAG_SKIP")) : 0;
    B200_CUDA(launch_pdl(embed_pdl_kernel, dim3(B), dim3(256), 0, st, m->emb, m->tokens, m->x, H));
    count_launch();
    for (int l = 0; l < c.n_layers; ++l) {
        const LayerW& L = m->layers[(size_t)l];
        const size_t kvoff = (size_t)l * B * NKV * c.max_ctx * D;
        if (!(skip & 4) && launch_gemv_fast(ctx, L.wqkv, B, PRO_RMS, nullptr, nullptr, m->x, L.attn_norm, c.rms_eps, EPI_STORE,
                               m->qkv, nullptr, nullptr, nullptr, nullptr, nullptr, use_pdl()))
            return fail("fused decode: qkv projection rejected by the streaming GEMV");
        if (!(skip & 1))
        B200_CUDA(launch_pdl(attn_fused_kernel, dim3(NH, B), dim3(ATTN_THREADS), attn_fused_smem(c.max_ctx), st, m->qkv, m->rope_cos,
                             m->rope_sin, m->kc + kvoff, m->vc + kvoff, m->att, m->pos, NH, NKV, D, c.max_ctx));
        count_launch();
        if (!(skip & 8) && launch_gemv_fast(ctx, L.wo, B, PRO_F32, nullptr, nullptr, m->att, nullptr, 0.f, EPI_RESIDUAL, m->x, m->x,
                               nullptr, nullptr, nullptr, nullptr, use_pdl()))
            return fail("fused decode: o projection rejected by the streaming GEMV");
        if (!(skip & 16) && launch_gemv_fast(ctx, L.wgateup, B, PRO_RMS, nullptr, nullptr, m->x, L.ffn_norm, c.rms_eps, EPI_SWIGLU,
                               m->gu, nullptr, nullptr, nullptr, nullptr, nullptr, use_pdl()))
            return fail("fused decode: gate/up projection rejected by the streaming GEMV");
        if (!(skip & 32) && launch_gemv_fast(ctx, L.wdown, B, PRO_F32, nullptr, nullptr, m->gu, nullptr, 0.f, EPI_RESIDUAL, m->x, m->x,
                               nullptr, nullptr, nullptr, nullptr, use_pdl()))
            return fail("fused decode: down projection rejected by the streaming GEMV");
    }
    {
        const int threads = (int)div_ceil((size_t)H / 8, 32) * 32;
        const int rpc = (int)LOGITS_ROWS_PER_CTA;
        const int grid = (int)div_ceil((size_t)c.vocab, (size_t)rpc);
        const size_t smem = 2 * (size_t)(threads / 32) * 16 * 4;
        for (int m0 = 0; m0 < B && !(skip & 2); m0 += 4) {
            const int mc = B - m0 < 4 ? B - m0 : 4;
            const float* xin = m->x + (size_t)m0 * H;
            float* lo = m->logits + (size_t)m0 * c.vocab;
            float* pv = m->part_val + (size_t)m0 * grid;
            int* pi = m->part_idx + (size_t)m0 * grid;
            cudaError_t e;
            if (mc == 1)
                e = launch_pdl(logits_fused_kernel<1, 8>, dim3(grid), dim3(threads), smem, st, m->emb, xin, m->final_norm,
                               c.rms_eps, lo, pv, pi, c.vocab, H, mc, rpc);
            else if (mc == 2)
                e = launch_pdl(logits_fused_kernel<2, 8>, dim3(grid), dim3(threads), smem, st, m->emb, xin, m->final_norm,
                               c.rms_eps, lo, pv, pi, c.vocab, H, mc, rpc);
            else
                e = launch_pdl(logits_fused_kernel<4, 4>, dim3(grid), dim3(threads), smem, st, m->emb, xin, m->final_norm,
                               c.rms_eps, lo, pv, pi, c.vocab, H, mc, rpc);
            B200_CUDA(e);
            count_launch();
        }
        if (!(skip & 2)) {
            B200_CUDA(launch_pdl(argmax_final_kernel, dim3(B), dim3(256), 0, st, m->part_val, m->part_idx, grid, m->next,
                                 m->tokens, m->pos));
            count_launch();
        }
    }
    m->launches_per_step = g_launches.load() - l0;
    return 0;
}

static int enqueue_step(bitnet_b200_model* m) {
    if (m->fused < 0) m->fused = fused_supported(m) ? 1 : 0;
    if (m->fused) return enqueue_step_fused(m);
    const auto& c = m->cfg;
    bitnet_b200_ctx* ctx = m->ctx;
    cudaStream_t st = ctx->stream;
    const int B = c.batch, H = c.hidden, F = c.ffn, D = c.head_dim, NH = c.n_heads, NKV = c.n_kv_heads;
    const int QD = NH * D;
    const uint64_t l0 = g_launches.load();
    embed_kernel<<<B, 256, 0, st>>>(m->emb, m->tokens, m->x, H);
    B200_LAUNCH_CHECK("embed_kernel");
    for (int l = 0; l < c.n_layers; ++l) {
        const LayerW& L = m->layers[(size_t)l];
        const size_t kvoff = (size_t)l * B * NKV * c.max_ctx * D;
        if (launch_actquant(ctx, m->x, L.attn_norm, c.rms_eps, B, H, m->q8, m->s8, nullptr, 1)) return -1;
        if (launch_gemv_i8(ctx, L.wqkv, m->q8, m->s8, B, m->qkv, nullptr, nullptr)) return -1;
        rope_kv_kernel<<<dim3(NH + 2 * NKV, B), D / 2, 0, st>>>(m->qkv, m->rope_cos, m->rope_sin,
                                                                m->kc + kvoff, m->vc + kvoff, m->pos,
                                                                NH, NKV, D, c.max_ctx);
        B200_LAUNCH_CHECK("rope_kv_kernel");
        attn_decode_kernel<<<dim3(NH, B), D, (size_t)c.max_ctx * 4, st>>>(
            m->qkv, m->kc + kvoff, m->vc + kvoff, m->att, m->pos, NH, NKV, D, c.max_ctx);
        B200_LAUNCH_CHECK("attn_decode_kernel");
        if (launch_actquant(ctx, m->att, nullptr, 0.f, B, QD, m->q8, m->s8, nullptr, 0)) return -1;
        if (launch_gemv_i8(ctx, L.wo, m->q8, m->s8, B, m->x, nullptr, m->x)) return -1;
        if (launch_actquant(ctx, m->x, L.ffn_norm, c.rms_eps, B, H, m->q8, m->s8, nullptr, 1)) return -1;
        if (launch_gemv_i8(ctx, L.wgateup, m->q8, m->s8, B, m->gu, nullptr, nullptr)) return -1;
        if (launch_actquant(ctx, m->gu, nullptr, 0.f, B, F, m->q8, m->s8, nullptr, 2)) return -1;
        if (launch_gemv_i8(ctx, L.wdown, m->q8, m->s8, B, m->x, nullptr, m->x)) return -1;
    }
    if (launch_actquant(ctx, m->x, m->final_norm, c.rms_eps, B, H, m->q8, m->s8, m->xn, 1)) return -1;
    {
        const int threads = (int)div_ceil((size_t)H / 8, 32) * 32;
        const int rpc = 64;
        const int grid = (int)div_ceil((size_t)c.vocab, rpc);
        for (int m0 = 0; m0 < B; m0 += 4) {
            const int mc = B - m0 < 4 ? B - m0 : 4;
            const float* xin = m->xn + (size_t)m0 * H;
            float* lo = m->logits + (size_t)m0 * c.vocab;
            if (mc == 1) {
                logits_f16_kernel<1, 8><<<grid, threads, 2 * (threads / 32) * 8 * 4, st>>>(m->emb, xin, lo, c.vocab, H, mc, rpc);
            } else if (mc == 2) {
                logits_f16_kernel<2, 8><<<grid, threads, 2 * (threads / 32) * 16 * 4, st>>>(m->emb, xin, lo, c.vocab, H, mc, rpc);
            } else {
                logits_f16_kernel<4, 4><<<grid, threads, 2 * (threads / 32) * 16 * 4, st>>>(m->emb, xin, lo, c.vocab, H, mc, rpc);
            }
            B200_LAUNCH_CHECK("logits_f16_kernel");
        }
    }
    argmax_kernel<<<B, 1024, 0, st>>>(m->logits, c.vocab, m->next, m->tokens, m->pos, 1);
    B200_LAUNCH_CHECK("argmax_kernel");
    m->launches_per_step = g_launches.load() - l0;
    return 0;
}

// The persistent one-kernel step covers QK256 models with batch <= 4 (decode_mega.cu);
// BITNET_B200_NO_MEGA=1 keeps the per-projection graph.
static int mega_setup(bitnet_b200_model* m) {
    if (m->mega >= 0) return 0;
    m->mega = 0;
    const auto& c = m->cfg;
    if (getenv("BITNET_B200_NO_MEGA")) return 0;
    if (c.layout != BITNET_B200_LAYOUT_QK256 && getenv("BITNET_B200_NO_MEGA_BITNET32")) return 0;
    if (!m->have_emb || !m->have_rope || m->layers_set != c.n_layers) return 0;
    int coop = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, m->ctx->device);
    if (!coop) return 0;
    MegaParams& p = m->mega_p;
    p = MegaParams{};
    p.L = c.n_layers; p.B = c.batch; p.H = c.hidden; p.F = c.ffn; p.NH = c.n_heads; p.NKV = c.n_kv_heads;
    p.D = c.head_dim; p.V = c.vocab; p.max_ctx = c.max_ctx; p.eps = c.rms_eps;
    p.layout = c.layout == BITNET_B200_LAYOUT_QK256 ? 0 : 1;
    m->mega_smem = mega_plan(&p, m->ctx->num_sms);
    if (!m->mega_smem) return 0;
    std::vector<MegaLayer> tab((size_t)c.n_layers);
    for (int l = 0; l < c.n_layers; ++l) {
        const LayerW& L = m->layers[(size_t)l];
        for (const bitnet_b200_weight* w : {L.wqkv, L.wo, L.wgateup, L.wdown})
            if (w->code_stride != w->cols / 4 || w->lut_word != m->layers[0].wqkv->lut_word ||
                (p.layout == 1 && (!w->f16_scales_exact || !w->scales)))  // clamped / overflowing block scales: per-projection graph
                return 0;
    }
    // fragment-major copies of every code plane (decode_mega.cu: mega_repack_kernel), made once
    for (int l = 0; l < c.n_layers; ++l) {
        const LayerW& L = m->layers[(size_t)l];
        uint8_t* fm[4] = {nullptr, nullptr, nullptr, nullptr};
        int j = 0;
        for (const bitnet_b200_weight* w : {L.wqkv, L.wo, L.wgateup, L.wdown}) {
            if (dmalloc((void**)&fm[j], mega_repacked_bytes(p.layout, (int)w->rows, (int)w->cols))) return -1;
            m->mega_codes.push_back(fm[j]);
            if (mega_repack_weight(w->codes, p.layout == 1 ? w->scales : nullptr, w->scale_policy == BITNET_B200_SCALE_ABS_CLAMP,
                                   (int)w->rows, (int)w->cols, fm[j], m->ctx->stream))
                return -1;
            ++j;
        }
        tab[(size_t)l] = MegaLayer{fm[0], fm[1], fm[2], fm[3], L.attn_norm, L.ffn_norm};
    }
    if (dmalloc((void**)&m->mega_layers, tab.size() * sizeof(MegaLayer)) || dmalloc((void**)&m->mega_bar, 4096)) return -1;  // 2 parities x 8 counters on 256-byte lines
    cudaStream_t st = m->ctx->stream;
    B200_CUDA(cudaMemcpyAsync(m->mega_layers, tab.data(), tab.size() * sizeof(MegaLayer), cudaMemcpyHostToDevice, st));
    B200_CUDA(cudaMemsetAsync(m->mega_bar, 0, 4096, st));
    B200_CUDA(cudaStreamSynchronize(st));
    p.layers = m->mega_layers;
    p.emb = m->emb; p.final_norm = m->final_norm; p.rope_cos = m->rope_cos; p.rope_sin = m->rope_sin;
    p.kc = m->kc; p.vc = m->vc; p.pos = m->pos; p.tokens = m->tokens; p.next = m->next;
    {
        const size_t B = (size_t)c.batch, nx = B * c.hidden, nqkv = B * (c.n_heads + 2 * c.n_kv_heads) * c.head_dim,
                     natt = B * c.n_heads * c.head_dim, ngu = B * c.ffn;
        const size_t nattp = B * c.n_heads * 8 * (size_t)(c.head_dim + 2);
        m->mega_ll_pairs = nx + nqkv + natt + ngu + nattp;
        if (dmalloc((void**)&m->mega_ll, m->mega_ll_pairs * sizeof(uint2))) return -1;
        B200_CUDA(cudaMemsetAsync(m->mega_ll, 0, m->mega_ll_pairs * sizeof(uint2), st));
        B200_CUDA(cudaStreamSynchronize(st));
        p.x = m->mega_ll; p.qkv = p.x + nx; p.att = p.qkv + nqkv; p.gu = p.att + natt; p.attp = p.gu + ngu;
    }
    p.att_split = getenv("BITNET_B200_MEGA_NOSPLIT") ? 0 : 1;
    p.logits = m->logits;
    p.barriers = getenv("BITNET_B200_MEGA_BARRIERS") ? 1 : 0;
    p.part_val = m->part_val; p.part_idx = m->part_idx;
    p.q_fixed = m->q8; p.s_fixed = m->s8; p.bar = m->mega_bar;
    p.lut = m->layers[0].wqkv->lut_word;
    p.prefetch = getenv("BITNET_B200_MEGA_NOPF") ? 0 : 1;
    if (getenv("BITNET_B200_MEGA_PROF")) {  // diagnostics: per-phase stamps of every CTA + per-warp trace of one CTA, dumped at destroy
        const size_t n = (size_t)m->ctx->num_sms * 1024 + 16 * 4096;
        if (dmalloc((void**)&p.prof, n * 8)) return -1;
        B200_CUDA(cudaMemset(p.prof, 0, n * 8));
        p.prof_cta = getenv("BITNET_B200_MEGA_PROF_CTA") ? atoi(getenv("BITNET_B200_MEGA_PROF_CTA")) : 0;
    }
    m->mega_parity = 0;
    m->mega = 1;
    m->launches_per_step = 1;
    return 0;
}

static int mega_launch(bitnet_b200_model* m, int mode) {
    MegaParams p = m->mega_p;
    p.mode = mode;
    p.parity = m->mega_parity;
    m->mega_parity ^= 1;
    const uint32_t span = 5u * (uint32_t)p.L + 1u;  // epochs one launch hands out
    if (m->mega_epoch > 0xffffffffu - 2u * span) {   // wrap (after ~28 M steps): start over on cleared flags
        B200_CUDA(cudaMemsetAsync(m->mega_ll, 0, m->mega_ll_pairs * sizeof(uint2), m->ctx->stream));
        m->mega_epoch = 1;
    }
    p.epoch = m->mega_epoch;
    m->mega_epoch += span;
    return launch_decode_mega(p, m->mega_smem, m->ctx->num_sms, m->ctx->stream);
}

// one decode step on the context's stream: the persistent kernel, or a replay of the step graph
static int ensure_graph(bitnet_b200_model* m);
static int step_launch(bitnet_b200_model* m) {
    // Context-length guard: the position lives on the device and every step appends one K/V row at it. Beyond
    // max_ctx the kernels would write past the cache, read the RoPE tables out of bounds and overflow the score array.
    if (m->host_pos >= m->cfg.max_ctx) {
        char buf[128];
        snprintf(buf, sizeof buf, "model_decode: context full (position %d, max_ctx %d): call model_reset", m->host_pos,
                 m->cfg.max_ctx);
        return fail(buf);
    }
    ++m->host_pos;
    if (mega_setup(m)) return -1;
    if (m->mega == 1) return mega_launch(m, 0);
    if (ensure_graph(m)) return -1;
    B200_CUDA(cudaGraphLaunch(m->graph, m->ctx->stream));
    count_launch(m->launches_per_step);
    return 0;
}

static int ensure_graph(bitnet_b200_model* m) {
    if (m->graph) return 0;
    if (!m->have_emb || !m->have_rope || m->layers_set != m->cfg.n_layers)
        return fail("model: weights incomplete (set every layer, the embedding and the RoPE tables first)");
    const int smem = m->cfg.max_ctx * 4;
    if (smem > 48 * 1024)
        B200_CUDA(cudaFuncSetAttribute(attn_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    if (attn_fused_smem(m->cfg.max_ctx) > 48 * 1024)
        B200_CUDA(cudaFuncSetAttribute(attn_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)attn_fused_smem(m->cfg.max_ctx)));
    // capture on a private stream (the legacy default stream cannot be captured), then the
    // instantiated graph is launched on the context's stream
    CaptureStream cap(m->ctx);
    if (!cap.ok()) return -1;
    cudaGraph_t g = nullptr;
    B200_CUDA(cudaStreamBeginCapture(cap.s, cudaStreamCaptureModeThreadLocal));
    const int rc = enqueue_step(m);
    cudaError_t e = cudaStreamEndCapture(cap.s, &g);
    cap.restore();
    if (rc) {
        if (g) cudaGraphDestroy(g);
        return -1;
    }
    if (e != cudaSuccess) return fail_cuda("cudaStreamEndCapture", e);
    e = cudaGraphInstantiate(&m->graph, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) return fail_cuda("cudaGraphInstantiate", e);
    return 0;
}

extern "C" {

int bitnet_b200_model_create(bitnet_b200_ctx* ctx, const bitnet_b200_model_config* cfg,
                             bitnet_b200_model** out) {
    try {
        if (!ctx || !cfg || !out) return fail("model_create: NULL argument");
        *out = nullptr;
        const auto& c = *cfg;
        if (c.hidden <= 0 || c.ffn <= 0 || c.n_layers <= 0 || c.n_heads <= 0 || c.n_kv_heads <= 0 ||
            c.head_dim <= 0 || c.vocab <= 0 || c.max_ctx <= 0 || c.batch <= 0)
            return fail("model_create: all dimensions must be > 0");
        if (c.n_heads % c.n_kv_heads) return fail("model_create: n_heads must be a multiple of n_kv_heads");
        if (c.head_dim % 64 || c.head_dim > 256) return fail("model_create: head_dim must be 64, 128, 192 or 256");
        if (c.hidden % 32 || c.ffn % 32 || c.hidden % 8) return fail("model_create: hidden and ffn must be multiples of 32");
        if ((size_t)c.max_ctx * 4 > 190 * 1024) return fail("model_create: max_ctx too large for the decode attention kernel");
        B200_CUDA(cudaSetDevice(ctx->device));
        std::unique_ptr<bitnet_b200_model, void (*)(bitnet_b200_model*)> m(new bitnet_b200_model(),
                                                                          bitnet_b200_model_destroy);
        m->ctx = ctx;
        m->cfg = c;
        m->layers.resize((size_t)c.n_layers);
        const size_t B = (size_t)c.batch, H = (size_t)c.hidden, F = (size_t)c.ffn, D = (size_t)c.head_dim;
        const size_t NH = (size_t)c.n_heads, NKV = (size_t)c.n_kv_heads;
        const size_t kv = (size_t)c.n_layers * B * NKV * c.max_ctx * D;
        if (dmalloc((void**)&m->kc, kv * 4) || dmalloc((void**)&m->vc, kv * 4)) return -1;
        if (dmalloc((void**)&m->pos, 4) || dmalloc((void**)&m->tokens, B * 4) || dmalloc((void**)&m->next, B * 4)) return -1;
        if (dmalloc((void**)&m->x, B * H * 4) || dmalloc((void**)&m->qkv, B * (NH + 2 * NKV) * D * 4) ||
            dmalloc((void**)&m->att, B * NH * D * 4) || dmalloc((void**)&m->gu, B * 2 * F * 4) ||
            dmalloc((void**)&m->xn, B * H * 4) || dmalloc((void**)&m->logits, B * (size_t)c.vocab * 4))
            return -1;
        const size_t qmax = std::max(std::max(H, F), NH * D);
        if (dmalloc((void**)&m->q8, B * qmax) || dmalloc((void**)&m->s8, B * 4)) return -1;
        const size_t nparts = std::max(div_ceil((size_t)c.vocab, LOGITS_ROWS_PER_CTA), (size_t)ctx->num_sms);
        if (dmalloc((void**)&m->part_val, B * nparts * 4) || dmalloc((void**)&m->part_idx, B * nparts * 4)) return -1;
        B200_CUDA(cudaMallocHost((void**)&m->next_host, B * 4));
        B200_CUDA(cudaMallocHost((void**)&m->tok_host, B * 4));
        B200_CUDA(cudaMemsetAsync(m->kc, 0, kv * 4, ctx->stream));
        B200_CUDA(cudaMemsetAsync(m->vc, 0, kv * 4, ctx->stream));
        B200_CUDA(cudaMemsetAsync(m->pos, 0, 4, ctx->stream));
        B200_CUDA(cudaMemsetAsync(m->tokens, 0, B * 4, ctx->stream));
        B200_CUDA(cudaStreamSynchronize(ctx->stream));
        *out = m.release();
        return 0;
    } catch (const std::exception& e) {
        return fail(e.what());
    }
}

void bitnet_b200_model_destroy(bitnet_b200_model* m) {
    if (!m) return;
    if (m->graph) cudaGraphExecDestroy(m->graph);
    if (m->mega_p.prof) {  // raw u64 dump: [num_sms][1024] phase stamps, then [16][4096] warp trace (tools/megaprof.py reads it)
        const size_t n = (size_t)m->ctx->num_sms * 1024 + 16 * 4096;
        std::vector<unsigned long long> h(n);
        if (cudaMemcpy(h.data(), m->mega_p.prof, n * 8, cudaMemcpyDeviceToHost) == cudaSuccess)
            if (FILE* f = fopen(getenv("BITNET_B200_MEGA_PROF") ? getenv("BITNET_B200_MEGA_PROF") : "/dev/null", "wb")) {
                fwrite(h.data(), 8, n, f);
                fclose(f);
            }
        cudaFree(m->mega_p.prof);
    }
    if (m->mega_layers) cudaFree(m->mega_layers);
    for (uint8_t* q : m->mega_codes) cudaFree(q);
    if (m->mega_bar) cudaFree(m->mega_bar);
    if (m->mega_ll) cudaFree(m->mega_ll);
    for (auto& L : m->layers) {
        bitnet_b200_weight_destroy(L.wqkv);
        bitnet_b200_weight_destroy(L.wo);
        bitnet_b200_weight_destroy(L.wgateup);
        bitnet_b200_weight_destroy(L.wdown);
        cudaFree(L.attn_norm);
        cudaFree(L.ffn_norm);
    }
    void* ptrs[] = {m->emb, m->final_norm, m->rope_cos, m->rope_sin, m->kc, m->vc, m->pos, m->tokens,
                    m->next, m->x, m->qkv, m->att, m->gu, m->xn, m->logits, m->q8, m->s8, m->part_val,
                    m->part_idx};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    if (m->next_host) cudaFreeHost(m->next_host);
    if (m->tok_host) cudaFreeHost(m->tok_host);
    delete m;
}

// One layer's seven projections, host bytes in the reference layout ([out, in] rows):
// wq [NH*D, H], wk/wv [NKV*D, H], wo [H, NH*D], wgate/wup [F, H], wdown [H, F].
// q|k|v are concatenated by rows, gate/up interleaved by rows (row 2i = gate i, 2i+1 = up i):
// both layouts keep a row's bytes contiguous, so fusing is a byte shuffle at upload.
int bitnet_b200_model_set_layer(bitnet_b200_model* m, int layer, const void* wq, const void* wk,
                                const void* wv, const void* wo, const void* wgate, const void* wup,
                                const void* wdown, const float* attn_norm, const float* ffn_norm) {
    try {
        if (!m || !wq || !wk || !wv || !wo || !wgate || !wup || !wdown || !attn_norm || !ffn_norm)
            return fail("model_set_layer: NULL argument");
        const auto& c = m->cfg;
        if (layer < 0 || layer >= c.n_layers) return fail("model_set_layer: layer out of range");
        if (m->graph || m->mega == 1) return fail("model_set_layer: the decode step is already built");
        B200_CUDA(cudaSetDevice(m->ctx->device));
        LayerW& L = m->layers[(size_t)layer];
        const bool fresh = L.wqkv == nullptr;
        const size_t H = (size_t)c.hidden, F = (size_t)c.ffn, QD = (size_t)c.n_heads * c.head_dim,
                     KD = (size_t)c.n_kv_heads * c.head_dim;
        const size_t rbH = row_bytes(c.layout, H), rbQ = row_bytes(c.layout, QD), rbF = row_bytes(c.layout, F);
        std::vector<uint8_t> qkv((QD + 2 * KD) * rbH), gu(2 * F * rbH);
        memcpy(qkv.data(), wq, QD * rbH);
        memcpy(qkv.data() + QD * rbH, wk, KD * rbH);
        memcpy(qkv.data() + (QD + KD) * rbH, wv, KD * rbH);
        for (size_t i = 0; i < F; ++i) {
            memcpy(gu.data() + (2 * i) * rbH, (const uint8_t*)wgate + i * rbH, rbH);
            memcpy(gu.data() + (2 * i + 1) * rbH, (const uint8_t*)wup + i * rbH, rbH);
        }
        bitnet_b200_weight_destroy(L.wqkv);
        bitnet_b200_weight_destroy(L.wo);
        bitnet_b200_weight_destroy(L.wgateup);
        bitnet_b200_weight_destroy(L.wdown);
        L.wqkv = L.wo = L.wgateup = L.wdown = nullptr;
        if (bitnet_b200_weight_create(m->ctx, qkv.data(), qkv.size(), c.layout, QD + 2 * KD, H, nullptr, 0, &L.wqkv)) return -1;
        if (bitnet_b200_weight_create(m->ctx, wo, H * rbQ, c.layout, H, QD, nullptr, 0, &L.wo)) return -1;
        if (bitnet_b200_weight_create(m->ctx, gu.data(), gu.size(), c.layout, 2 * F, H, nullptr, 0, &L.wgateup)) return -1;
        if (bitnet_b200_weight_create(m->ctx, wdown, H * rbF, c.layout, H, F, nullptr, 0, &L.wdown)) return -1;
        if (upload(&L.attn_norm, attn_norm, H, m->ctx->stream)) return -1;
        if (upload(&L.ffn_norm, ffn_norm, H, m->ctx->stream)) return -1;
        if (fresh) m->layers_set++;
        return 0;
    } catch (const std::exception& e) {
        return fail(e.what());
    }
}

int bitnet_b200_model_set_embedding(bitnet_b200_model* m, const uint16_t* emb_f16,
                                    const float* final_norm) {
    if (!m || !emb_f16 || !final_norm) return fail("model_set_embedding: NULL argument");
    B200_CUDA(cudaSetDevice(m->ctx->device));
    if (upload(reinterpret_cast<uint16_t**>(&m->emb), emb_f16, (size_t)m->cfg.vocab * m->cfg.hidden, m->ctx->stream)) return -1;
    if (upload(&m->final_norm, final_norm, (size_t)m->cfg.hidden, m->ctx->stream)) return -1;
    m->have_emb = true;
    return 0;
}

int bitnet_b200_model_set_rope(bitnet_b200_model* m, const float* cos_t, const float* sin_t) {
    if (!m || !cos_t || !sin_t) return fail("model_set_rope: NULL argument");
    B200_CUDA(cudaSetDevice(m->ctx->device));
    const size_t n = (size_t)m->cfg.max_ctx * (m->cfg.head_dim / 2);
    if (upload(&m->rope_cos, cos_t, n, m->ctx->stream)) return -1;
    if (upload(&m->rope_sin, sin_t, n, m->ctx->stream)) return -1;
    m->have_rope = true;
    return 0;
}

int bitnet_b200_model_reset(bitnet_b200_model* m) {
    if (!m) return fail("model_reset: NULL argument");
    B200_CUDA(cudaMemsetAsync(m->pos, 0, 4, m->ctx->stream));
    m->host_pos = 0;
    return 0;
}

// One decode step through host buffers: H2D of the batch's token ids (pinned staging), the
// step graph, D2H of the greedy tokens (and optionally the logits); synchronous.
int bitnet_b200_model_decode(bitnet_b200_model* m, const int32_t* tokens, int32_t* next_tokens,
                             float* logits_out) {
    try {
        if (!m || !tokens || !next_tokens) return fail("model_decode: NULL argument");
        B200_CUDA(cudaSetDevice(m->ctx->device));
        cudaStream_t st = m->ctx->stream;
        const size_t B = (size_t)m->cfg.batch;
        for (size_t b = 0; b < B; ++b) {
            if (tokens[b] < 0 || tokens[b] >= m->cfg.vocab) return fail("model_decode: token id out of range");
            m->tok_host[b] = tokens[b];
        }
        B200_CUDA(cudaMemcpyAsync(m->tokens, m->tok_host, B * 4, cudaMemcpyHostToDevice, st));
        if (step_launch(m)) return -1;
        B200_CUDA(cudaMemcpyAsync(m->next_host, m->next, B * 4, cudaMemcpyDeviceToHost, st));
        if (logits_out)
            B200_CUDA(cudaMemcpyAsync(logits_out, m->logits, B * (size_t)m->cfg.vocab * 4, cudaMemcpyDeviceToHost, st));
        B200_CUDA(cudaStreamSynchronize(st));
        for (size_t b = 0; b < B; ++b) next_tokens[b] = m->next_host[b];
        return 0;
    } catch (const std::exception& e) {
        return fail(e.what());
    }
}

// n greedy steps entirely on device (token fed back by the argmax kernel); asynchronous.
// The current device-side tokens are used as the first input (set them with one decode()
// or model_set_tokens()).
int bitnet_b200_model_decode_device(bitnet_b200_model* m, int n_steps) {
    if (!m || n_steps < 0) return fail("model_decode_device: bad argument");
    if (m->host_pos + n_steps > m->cfg.max_ctx) {  // nothing is launched when the request does not fit
        char buf[160];
        snprintf(buf, sizeof buf, "model_decode_device: %d steps from position %d exceed max_ctx %d", n_steps, m->host_pos,
                 m->cfg.max_ctx);
        return fail(buf);
    }
    B200_CUDA(cudaSetDevice(m->ctx->device));
    for (int i = 0; i < n_steps; ++i)
        if (step_launch(m)) return -1;
    return 0;
}

int bitnet_b200_model_set_tokens(bitnet_b200_model* m, const int32_t* tokens) {
    if (!m || !tokens) return fail("model_set_tokens: NULL argument");
    const size_t B = (size_t)m->cfg.batch;
    for (size_t b = 0; b < B; ++b) {
        if (tokens[b] < 0 || tokens[b] >= m->cfg.vocab) return fail("model_set_tokens: token id out of range");
        m->tok_host[b] = tokens[b];
    }
    B200_CUDA(cudaMemcpyAsync(m->tokens, m->tok_host, B * 4, cudaMemcpyHostToDevice, m->ctx->stream));
    B200_CUDA(cudaStreamSynchronize(m->ctx->stream));
    return 0;
}

int bitnet_b200_model_get_tokens(bitnet_b200_model* m, int32_t* next_tokens) {
    if (!m || !next_tokens) return fail("model_get_tokens: NULL argument");
    const size_t B = (size_t)m->cfg.batch;
    B200_CUDA(cudaMemcpyAsync(m->next_host, m->next, B * 4, cudaMemcpyDeviceToHost, m->ctx->stream));
    B200_CUDA(cudaStreamSynchronize(m->ctx->stream));
    for (size_t b = 0; b < B; ++b) next_tokens[b] = m->next_host[b];
    return 0;
}

// Roofline leg of bench.py: ONLY the ternary linears of one decode step (4 fused GEMVs per
// layer, this model's real weights -- together larger than L2), captured as one graph and
// replayed `reps` times between two events on the context's stream. Reports the mean time
// of one pass and how many GEMV launches a pass holds.
int bitnet_b200_model_time_linears(bitnet_b200_model* m, int reps, float* ms_per_pass,
                                   int* launches_per_pass) {
    try {
        if (!m || reps <= 0 || !ms_per_pass || !launches_per_pass) return fail("model_time_linears: bad argument");
        if (m->layers_set != m->cfg.n_layers) return fail("model_time_linears: weights incomplete");
        B200_CUDA(cudaSetDevice(m->ctx->device));
        const auto& c = m->cfg;
        cudaStream_t st = m->ctx->stream;
        const int B = c.batch;
        if (mega_setup(m)) return -1;
        if (m->mega == 1) {
            // the persistent kernel restricted to its 4L ternary GEMV phases (fixed int8 inputs, no
            // barriers): one launch streams every packed ternary weight of the model once
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0);
            cudaEventCreate(&e1);
            int rc = mega_launch(m, 1);  // warm-up pass
            cudaEventRecord(e0, st);
            for (int i = 0; i < reps && !rc; ++i) rc = mega_launch(m, 1);
            cudaEventRecord(e1, st);
            cudaError_t e = cudaStreamSynchronize(st);
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
            if (rc) return -1;
            if (e != cudaSuccess) return fail_cuda("model_time_linears", e);
            // The pass that was timed must have computed the linears: what the LAST layer's qkv and down phases left in
            // the hand-off buffers is compared with the stand-alone GEMV on the same fixed int8 rows (QK256: the same
            // one-multiply epilogue on the same int32 sums, bit for bit; BitNet32-F16: f32 block sums in another order).
            {
                const LayerW& L = m->layers[(size_t)c.n_layers - 1];
                const bitnet_b200_weight* ws[2] = {L.wqkv, L.wdown};
                const uint2* outs[2] = {m->mega_p.qkv, m->mega_p.x};
                for (int k = 0; k < 2; ++k) {
                    const size_t n = (size_t)B * ws[k]->rows;
                    float* ref = nullptr;
                    B200_CUDA(cudaMalloc(&ref, n * sizeof(float)));
                    std::vector<float> h_ref(n);
                    std::vector<uint2> h_out(n);
                    int rc2 = launch_gemv_i8(m->ctx, ws[k], m->q8, m->s8, B, ref, nullptr, nullptr);
                    if (!rc2 && cudaMemcpyAsync(h_ref.data(), ref, n * sizeof(float), cudaMemcpyDeviceToHost, st) != cudaSuccess) rc2 = -1;
                    if (!rc2 && cudaMemcpyAsync(h_out.data(), outs[k], n * sizeof(uint2), cudaMemcpyDeviceToHost, st) != cudaSuccess) rc2 = -1;
                    if (!rc2 && cudaStreamSynchronize(st) != cudaSuccess) rc2 = -1;
                    cudaFree(ref);
                    if (rc2) return fail("model_time_linears: could not verify the GEMV-phase pass");
                    float amax = 0.f;
                    for (size_t i = 0; i < n; ++i) amax = fmaxf(amax, fabsf(h_ref[i]));
                    const float tol = m->mega_p.layout == 0 ? 0.f : 1e-4f * amax;
                    for (size_t i = 0; i < n; ++i) {
                        float v;
                        memcpy(&v, &h_out[i].x, 4);
                        if (!(fabsf(v - h_ref[i]) <= tol)) {
                            char buf[200];
                            snprintf(buf, sizeof buf, "model_time_linears: the GEMV-phase pass disagrees with linear_i8 (%s, element %zu: %g vs %g)",
                                     k == 0 ? "qkv" : "down", i, (double)v, (double)h_ref[i]);
                            return fail(buf);
                        }
                    }
                }
            }
            *ms_per_pass = ms / (float)reps;
            *launches_per_pass = 1;
            return 0;
        }
        cudaGraph_t g = nullptr;
        cudaGraphExec_t ge = nullptr;
        const uint64_t l0 = g_launches.load();
        CaptureStream cap(m->ctx);
        if (!cap.ok()) return -1;
        B200_CUDA(cudaStreamBeginCapture(cap.s, cudaStreamCaptureModeThreadLocal));
        int rc = 0;
        for (int l = 0; l < c.n_layers && !rc; ++l) {
            const LayerW& L = m->layers[(size_t)l];
            const bitnet_b200_weight* ws[4] = {L.wqkv, L.wo, L.wgateup, L.wdown};
            float* outs[4] = {m->qkv, m->xn, m->gu, m->xn};
            for (int i = 0; i < 4 && !rc; ++i) {
                int r = 1;
                if (m->fused != 0)
                    r = launch_gemv_fast(m->ctx, ws[i], B, PRO_Q8, m->q8, m->s8, nullptr, nullptr, 0.f, EPI_STORE,
                                           outs[i], nullptr, nullptr, nullptr, nullptr, nullptr, use_pdl());
                if (r == 1) r = launch_gemv_i8(m->ctx, ws[i], m->q8, m->s8, B, outs[i], nullptr, nullptr);
                rc = r;
            }
        }
        cudaError_t e = cudaStreamEndCapture(cap.s, &g);
        cap.restore();
        const int per_pass = (int)(g_launches.load() - l0);
        if (rc) {
            if (g) cudaGraphDestroy(g);
            return -1;
        }
        if (e != cudaSuccess) return fail_cuda("cudaStreamEndCapture", e);
        e = cudaGraphInstantiate(&ge, g, 0);
        cudaGraphDestroy(g);
        if (e != cudaSuccess) return fail_cuda("cudaGraphInstantiate", e);
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaGraphLaunch(ge, st);  // warm-up pass
        cudaEventRecord(e0, st);
        for (int i = 0; i < reps; ++i) cudaGraphLaunch(ge, st);
        cudaEventRecord(e1, st);
        e = cudaStreamSynchronize(st);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        cudaGraphExecDestroy(ge);
        if (e != cudaSuccess) return fail_cuda("model_time_linears", e);
        count_launch((uint64_t)per_pass * (uint64_t)reps);
        *ms_per_pass = ms / (float)reps;
        *launches_per_pass = per_pass;
        return 0;
    } catch (const std::exception& e) {
        return fail(e.what());
    }
}

uint64_t bitnet_b200_model_launches_per_step(const bitnet_b200_model* m) { return m ? m->launches_per_step : 0; }
uint64_t bitnet_b200_model_weight_bytes(const bitnet_b200_model* m) {
    if (!m) return 0;
    uint64_t n = 0;
    for (const auto& L : m->layers)
        for (const bitnet_b200_weight* w : {L.wqkv, L.wo, L.wgateup, L.wdown})
            if (w) n += w->ref_nbytes;
    return n;
}

}  // extern "C"
