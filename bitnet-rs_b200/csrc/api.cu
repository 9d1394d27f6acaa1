// api.cu -- the C ABI of libbitnet_b200.so (include/bitnet_b200.h).
#include <cstdio>
#include <cstring>
#include <memory>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace b200 {

static thread_local std::string t_err;
std::atomic<uint64_t> g_launches{0};
thread_local const char* g_last_linear_kernel = "none";

void set_error(const std::string& msg) { t_err = msg; }
void clear_error() { t_err.clear(); }
int fail(const std::string& msg) {
    t_err = msg;
    return -1;
}
int fail_cuda(const char* what, cudaError_t e) {
    t_err = std::string("CUDA error in ") + what + ": " + cudaGetErrorName(e) + " (" +
            cudaGetErrorString(e) + ")";
    return -1;
}

// kernels (gemv.cu, quant.cu, gemm_tcgen05.cu)
int launch_gemv_i8(const bitnet_b200_ctx*, const bitnet_b200_weight*, const int8_t*, const float*,
                   int, float*, int32_t*, const float* residual);
int launch_gemv_f32(const bitnet_b200_ctx*, const bitnet_b200_weight*, const float*, int, float*);
int launch_gemv_fast(const bitnet_b200_ctx*, const bitnet_b200_weight*, int M, int pro, const int8_t* q,
                       const float* act_scale, const float* x, const float* gamma, float eps, int epi, float* y,
                       const float* residual, int32_t* acc_out, int8_t* q_out, float* scale_out, float* normed_out,
                       bool pdl);
bool gemm_small_supported(const bitnet_b200_ctx*, const bitnet_b200_weight*, int M);
int launch_gemm_small(const bitnet_b200_ctx*, const bitnet_b200_weight*, const int8_t*, const float*, int M, float* y,
                      int32_t* acc_out, bool pdl);
int launch_gemm_i8(const bitnet_b200_ctx*, const bitnet_b200_weight*, const int8_t*, const float*,
                   int, float*, int32_t*, float*);
bool gemm_i8_supported(const bitnet_b200_weight*, int M);
int launch_gemm_f16(const bitnet_b200_ctx*, const bitnet_b200_weight*, const int8_t*, const float*, int, float*);
bool gemm_f16_supported(const bitnet_b200_weight*, int M);
int launch_actquant(const bitnet_b200_ctx*, const float*, const float*, float, int, int, int8_t*,
                    float*, float*, int mode);
int launch_quantize_weights(cudaStream_t, const float*, int, uint8_t*, int, float*, int);
int launch_matmul_u8(cudaStream_t, const int8_t*, const uint8_t*, float*, int, int, int);
int launch_qk256_legacy(cudaStream_t, const uint8_t*, const float*, const float*, float*, int, int);
int launch_i2s_matmul_f32(cudaStream_t, const float*, const uint8_t*, const float*, float*, int,
                          int, int, int);
int launch_split_bitnet32(cudaStream_t, const uint8_t*, size_t, uint8_t*, __half*, float*, int);
int launch_repack_codes(cudaStream_t, uint8_t*, size_t);

// ---- process-wide default context for the slice API (Part 1/2) --------------------------
static std::mutex g_mu;
static int g_refcount = 0;
static int g_device = -1;  // -1 = not probed, -2 = unavailable

static int probe_device_locked() {
    if (g_device != -1) return g_device;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        cudaGetLastError();
        g_device = -2;
        return g_device;
    }
    int dev = 0;
    if (const char* e = getenv("BITNET_B200_DEVICE")) dev = atoi(e);
    cudaDeviceProp prop;
    if (dev < 0 || dev >= n || cudaGetDeviceProperties(&prop, dev) != cudaSuccess || prop.major != 10) {
        cudaGetLastError();
        g_device = -2;
        return g_device;
    }
    g_device = dev;
    return g_device;
}

static int default_device() {
    std::lock_guard<std::mutex> lk(g_mu);
    return probe_device_locked();
}

// RAII device buffer for the synchronous slice API
struct DevBuf {
    void* p = nullptr;
    ~DevBuf() {
        if (p) cudaFree(p);
    }
    int alloc(size_t n) {
        B200_CUDA(cudaMalloc(&p, n ? n : 1));
        return 0;
    }
    template <typename T>
    T* as() {
        return reinterpret_cast<T*>(p);
    }
};

struct StreamGuard {
    cudaStream_t s = nullptr;
    ~StreamGuard() {
        if (s) cudaStreamDestroy(s);
    }
    int create() {
        B200_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        return 0;
    }
};

static int need_device(int* dev) {
    const int d = default_device();
    if (d < 0) return fail("no CUDA device: libbitnet_b200 needs an sm_100 (B200) GPU and has no CPU fallback");
    B200_CUDA(cudaSetDevice(d));
    *dev = d;
    return 0;
}

}  // namespace b200

using namespace b200;

extern "C" {

// ======================================================================================
// Part 1 -- the reference's kernel bridge
// ======================================================================================
int bitnet_cpp_init(void) {
    try {
        clear_error();
        std::lock_guard<std::mutex> lk(g_mu);
        if (probe_device_locked() < 0) {
            set_error("BitNet B200 kernels not available: no sm_100 CUDA device");
            return -1;
        }
        ++g_refcount;
        return 0;
    } catch (...) {
        set_error("Unknown initialization error");
        return -1;
    }
}

void bitnet_cpp_cleanup(void) {
    try {
        clear_error();
        std::lock_guard<std::mutex> lk(g_mu);
        if (g_refcount > 0) --g_refcount;
    } catch (...) {
    }
}

int bitnet_cpp_is_available(void) { return default_device() >= 0 ? 1 : 0; }

const char* bitnet_cpp_get_last_error(void) { return t_err.empty() ? nullptr : t_err.c_str(); }

int bitnet_cpp_matmul_i2s(const int8_t* a, const uint8_t* b, float* c, int m, int n, int k) {
    try {
        if (!a || !b || !c) return fail("Null pointer passed to matmul_i2s");
        if (m <= 0 || n <= 0 || k <= 0) return fail("Invalid matrix dimensions");
        int dev;
        if (need_device(&dev)) return -1;
        StreamGuard sg;
        if (sg.create()) return -1;
        DevBuf da, db, dc;
        const size_t na = (size_t)m * k, nb = (size_t)k * n, nc = (size_t)m * n * sizeof(float);
        if (da.alloc(na) || db.alloc(nb) || dc.alloc(nc)) return -1;
        B200_CUDA(cudaMemcpyAsync(da.p, a, na, cudaMemcpyHostToDevice, sg.s));
        B200_CUDA(cudaMemcpyAsync(db.p, b, nb, cudaMemcpyHostToDevice, sg.s));
        if (launch_matmul_u8(sg.s, da.as<int8_t>(), db.as<uint8_t>(), dc.as<float>(), m, n, k)) return -1;
        B200_CUDA(cudaMemcpyAsync(c, dc.p, nc, cudaMemcpyDeviceToHost, sg.s));
        B200_CUDA(cudaStreamSynchronize(sg.s));
        return 0;
    } catch (const std::exception& e) {
        return fail(e.what());
    } catch (...) {
        return fail("Unknown matmul error");
    }
}

int bitnet_cpp_quantize(const float* in, int in_len, uint8_t* out, int out_len, float* scales,
                        int scales_len, int qtype) {
    try {
        if (!in || !out || !scales) return fail("Null pointer passed to quantize");
        if (in_len <= 0 || out_len <= 0 || scales_len <= 0) return fail("Invalid buffer lengths");
        if (qtype < 0 || qtype > 2) return fail("Invalid quantization type");
        const int block = qtype == 0 ? 32 : (qtype == 1 ? 64 : 128);
        const char* nm = qtype == 0 ? "I2_S" : (qtype == 1 ? "TL1" : "TL2");
        const int nblocks = (in_len + block - 1) / block;
        char buf[160];
        if (out_len < in_len / 4) {  // cpu/fallback.rs:106-114
            snprintf(buf, sizeof buf, "Output buffer too small for %s: expected %d, got %d", nm,
                     in_len / 4, out_len);
            return fail(buf);
        }
        if (scales_len < nblocks) {  // cpu/fallback.rs:116-124
            snprintf(buf, sizeof buf, "Scales buffer too small: expected %d, got %d", nblocks,
                     scales_len);
            return fail(buf);
        }
        int dev;
        if (need_device(&dev)) return -1;
        StreamGuard sg;
        if (sg.create()) return -1;
        DevBuf di, dout, ds;
        if (di.alloc((size_t)in_len * 4) || dout.alloc((size_t)out_len) || ds.alloc((size_t)nblocks * 4))
            return -1;
        B200_CUDA(cudaMemcpyAsync(di.p, in, (size_t)in_len * 4, cudaMemcpyHostToDevice, sg.s));
        // OR semantics: start from the caller's bytes
        B200_CUDA(cudaMemcpyAsync(dout.p, out, (size_t)out_len, cudaMemcpyHostToDevice, sg.s));
        if (launch_quantize_weights(sg.s, di.as<float>(), in_len, dout.as<uint8_t>(), out_len,
                                    ds.as<float>(), qtype))
            return -1;
        B200_CUDA(cudaMemcpyAsync(out, dout.p, (size_t)out_len, cudaMemcpyDeviceToHost, sg.s));
        B200_CUDA(cudaMemcpyAsync(scales, ds.p, (size_t)nblocks * 4, cudaMemcpyDeviceToHost, sg.s));
        B200_CUDA(cudaStreamSynchronize(sg.s));
        return 0;
    } catch (const std::exception& e) {
        return fail(e.what());
    } catch (...) {
        return fail("Unknown quantization error");
    }
}

// ======================================================================================
// Part 3 -- contexts and weights
// ======================================================================================
int bitnet_b200_ctx_create(int device, void* stream, bitnet_b200_ctx** out) {
    try {
        if (!out) return fail("ctx_create: out is NULL");
        *out = nullptr;
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
            cudaGetLastError();
            return fail("no CUDA device: libbitnet_b200 needs an sm_100 (B200) GPU and has no CPU fallback");
        }
        if (device < 0 || device >= n) return fail("ctx_create: invalid device ordinal");
        cudaDeviceProp prop;
        B200_CUDA(cudaGetDeviceProperties(&prop, device));
        if (prop.major != 10)
            return fail(std::string("ctx_create: device is sm_") + std::to_string(prop.major) +
                        std::to_string(prop.minor) + ", this library is built for sm_100a only");
        B200_CUDA(cudaSetDevice(device));
        auto* c = new bitnet_b200_ctx();
        c->device = device;
        c->num_sms = prop.multiProcessorCount;
        // NULL is a real stream: the legacy default stream (what torch uses by default)
        c->stream = reinterpret_cast<cudaStream_t>(stream);
        cudaError_t e = cudaMalloc(&c->workspace, 4096);
        if (e == cudaSuccess) e = cudaMemsetAsync(c->workspace, 0, 4096, c->stream);
        if (e != cudaSuccess) {
            delete c;
            return fail_cuda("cudaMalloc(workspace)", e);
        }
        *out = c;
        return 0;
    } catch (...) {
        return fail("ctx_create: unknown error");
    }
}

int bitnet_b200_ctx_set_stream(bitnet_b200_ctx* ctx, void* stream) {
    if (!ctx) return fail("ctx is NULL");
    ctx->stream = reinterpret_cast<cudaStream_t>(stream);
    return 0;
}

int bitnet_b200_ctx_synchronize(bitnet_b200_ctx* ctx) {
    if (!ctx) return fail("ctx is NULL");
    B200_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

void bitnet_b200_ctx_destroy(bitnet_b200_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->workspace) cudaFree(ctx->workspace);
    if (ctx->gemm_scratch) cudaFree(ctx->gemm_scratch);
    if (ctx->gemm_counters) cudaFree(ctx->gemm_counters);
    delete ctx;
}

int bitnet_b200_weight_create(bitnet_b200_ctx* ctx, const void* host_bytes, size_t nbytes,
                              int layout, size_t rows, size_t cols, const int8_t* lut,
                              int scale_policy, bitnet_b200_weight** out) {
    try {
        if (!ctx || !host_bytes || !out) return fail("weight_create: NULL argument");
        *out = nullptr;
        if (rows == 0 || cols == 0) return fail("weight_create: rows and cols must be > 0");
        if (scale_policy != BITNET_B200_SCALE_ABS_CLAMP && scale_policy != BITNET_B200_SCALE_RAW)
            return fail("weight_create: invalid scale policy");
        B200_CUDA(cudaSetDevice(ctx->device));
        auto w = new bitnet_b200_weight();
        std::unique_ptr<bitnet_b200_weight, void (*)(bitnet_b200_weight*)> guard(w, bitnet_b200_weight_destroy);
        w->device = ctx->device;
        w->layout = layout;
        w->rows = rows;
        w->cols = cols;
        w->scale_policy = scale_policy;
        static const int8_t MAP_A[4] = {-2, -1, 1, 2};
        const int8_t* l = lut ? lut : MAP_A;
        w->lut_word = 0;
        for (int i = 0; i < 4; ++i) {
            w->lut[i] = l[i];
            w->lut_word |= (uint32_t)(uint8_t)l[i] << (8 * i);
        }
        char buf[256];
        if (layout == BITNET_B200_LAYOUT_QK256) {
            const size_t stride = div_ceil(cols, 256) * 64;
            const size_t expected = rows * stride;
            const size_t diff = nbytes > expected ? nbytes - expected : expected - nbytes;
            if (diff > 128) {  // I2SQk256NoScale::new, i2s_qk256.rs:85-104
                snprintf(buf, sizeof buf,
                         "I2SQk256NoScale: data size mismatch: got %zu bytes, expected %zu for %zux%zu matrix. "
                         "Check tensor orientation: QK256 requires [out_dim, in_dim] layout.",
                         nbytes, expected, rows, cols);
                return fail(buf);
            }
            w->code_stride = stride;
            w->ref_nbytes = expected;
            B200_CUDA(cudaMalloc(&w->codes, expected));
            const size_t ncopy = nbytes < expected ? nbytes : expected;
            if (ncopy < expected) B200_CUDA(cudaMemsetAsync(w->codes, 0, expected, ctx->stream));
            B200_CUDA(cudaMemcpyAsync(w->codes, host_bytes, ncopy, cudaMemcpyHostToDevice, ctx->stream));
            if (launch_repack_codes(ctx->stream, w->codes, expected)) return -1;  // stride % 64 == 0
        } else if (layout == BITNET_B200_LAYOUT_BITNET32_F16) {
            if (cols % 32 != 0) return fail("weight_create: BitNet32-F16 requires cols % 32 == 0");
            const size_t nblk = rows * (cols / 32);
            if (nbytes != nblk * 10) {  // quant/i2s.rs:631-633
                snprintf(buf, sizeof buf, "I2_S: internal size mismatch for block=32 (got %zu bytes, expected %zu)",
                         nbytes, nblk * 10);
                return fail(buf);
            }
            w->code_stride = cols / 4;
            w->nblk32 = cols / 32;
            w->ref_nbytes = nbytes;
            DevBuf staging;
            if (staging.alloc(nbytes)) return -1;
            B200_CUDA(cudaMalloc(&w->codes, nblk * 8));
            B200_CUDA(cudaMalloc(&w->scales, nblk * sizeof(__half)));
            B200_CUDA(cudaMalloc(&w->scales_f32, nblk * sizeof(float)));
            B200_CUDA(cudaMemcpyAsync(staging.p, host_bytes, nbytes, cudaMemcpyHostToDevice, ctx->stream));
            {   // host scan of the f16 scale bits (the bytes are here anyway): see f16_scales_exact
                const uint8_t* hb = static_cast<const uint8_t*>(host_bytes);
                int amax = 0;
                for (int i = 0; i < 4; ++i) amax = std::max(amax, std::abs((int)w->lut[i]));
                bool exact = true;
                for (size_t b = 0; b < nblk && exact; ++b) {
                    const uint16_t bits = (uint16_t)(hb[b * 10 + 8] | (hb[b * 10 + 9] << 8));
                    const uint16_t mag = bits & 0x7fffu;  // f16 magnitudes order like their bit patterns
                    if (scale_policy == BITNET_B200_SCALE_ABS_CLAMP) {
                        // s = clamp(|h|, 1e-3, 1e3): exact in f16 iff unclamped (1e-3 is not an f16; 1e3 = 0x63D0 is)
                        if (mag < 0x1419u /* first f16 >= 1e-3 */ || mag > 0x63D0u /* 1000 */) exact = false;
                    } else {
                        if (mag >= 0x7C00u) exact = false;  // inf / NaN
                        else {                               // |s| * max|LUT| must stay finite in f16
                            const int e = mag >> 10, f = mag & 0x3ff;
                            const float hv = e ? ldexpf(1.0f + f / 1024.0f, e - 15) : ldexpf(f / 1024.0f, -14);
                            if (hv * (float)amax > 65504.0f) exact = false;
                        }
                    }
                }
                w->f16_scales_exact = exact;
            }
            if (launch_split_bitnet32(ctx->stream, staging.as<uint8_t>(), nblk, w->codes, w->scales,
                                      w->scales_f32, scale_policy == BITNET_B200_SCALE_ABS_CLAMP))
                return -1;
            B200_CUDA(cudaStreamSynchronize(ctx->stream));
        } else {
            return fail("weight_create: unknown layout");
        }
        B200_CUDA(cudaStreamSynchronize(ctx->stream));
        *out = guard.release();
        return 0;
    } catch (const std::exception& e) {
        return fail(e.what());
    } catch (...) {
        return fail("weight_create: unknown error");
    }
}

void bitnet_b200_weight_destroy(bitnet_b200_weight* w) {
    if (!w) return;
    cudaSetDevice(w->device);
    if (w->codes) cudaFree(w->codes);
    if (w->scales) cudaFree(w->scales);
    if (w->scales_f32) cudaFree(w->scales_f32);
    delete w;
}

size_t bitnet_b200_weight_rows(const bitnet_b200_weight* w) { return w ? w->rows : 0; }
size_t bitnet_b200_weight_cols(const bitnet_b200_weight* w) { return w ? w->cols : 0; }
size_t bitnet_b200_weight_nbytes(const bitnet_b200_weight* w) { return w ? w->ref_nbytes : 0; }

// ======================================================================================
// Part 3 -- device-pointer compute
// ======================================================================================
int bitnet_b200_actquant_i8(bitnet_b200_ctx* ctx, const float* x, int M, int K, int8_t* q,
                            float* scale) {
    if (!ctx || !x || !q || !scale) return fail("actquant_i8: NULL argument");
    B200_CUDA(cudaSetDevice(ctx->device));
    return launch_actquant(ctx, x, nullptr, 0.f, M, K, q, scale, nullptr, 0);
}

int bitnet_b200_rmsnorm_actquant(bitnet_b200_ctx* ctx, const float* x, const float* gamma,
                                 float eps, int M, int K, int8_t* q, float* scale, float* normed) {
    if (!ctx || !x || !gamma || !q || !scale) return fail("rmsnorm_actquant: NULL argument");
    if (!(eps > 0.f) || !isfinite(eps))  // validate_rms_norm_args, cpu/layer_norm.rs
        return fail("rmsnorm_actquant: eps must be positive and finite");
    B200_CUDA(cudaSetDevice(ctx->device));
    return launch_actquant(ctx, x, gamma, eps, M, K, q, scale, normed, 1);
}

int bitnet_b200_linear_i8(bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q,
                          const float* act_scale, int M, float* y, int32_t* acc_out) {
    if (!ctx || !w || !q || !act_scale || !y) return fail("linear_i8: NULL argument");
    if (M <= 0) return fail("linear_i8: M must be > 0");
    if (w->device != ctx->device) return fail("linear_i8: weight lives on another device");
    B200_CUDA(cudaSetDevice(ctx->device));
    // per-32-block accumulators are only materialised by the GEMV
    const bool force_gemv = acc_out && w->layout == BITNET_B200_LAYOUT_BITNET32_F16;
    if (!force_gemv && M <= 64) {
        // batched-decode regime: mma.sync kernel without the tcgen05 pipeline's fixed cost (gemm_small.cu). 1 = not its
        // shape, or the other kernel of this M range (tcgen05 GEMM / GEMV) measured faster on it: fall through
        const int rc = launch_gemm_small(ctx, w, q, act_scale, M, y, acc_out, getenv("BITNET_B200_PDL_API") != nullptr);
        if (rc != 1) return rc;  // g_last_linear_kernel set by the launcher (two variants)
    }
    if (!force_gemv && gemm_i8_supported(w, M)) {
        g_last_linear_kernel = "gemm_tcgen05";
        return launch_gemm_i8(ctx, w, q, act_scale, M, y, acc_out, nullptr);
    }
    if (!acc_out && gemm_f16_supported(w, M)) {  // BitNet32-F16, M >= 9: tcgen05 kind::f16 on s * LUT[code]
        g_last_linear_kernel = "gemm_f16_tcgen05";
        return launch_gemm_f16(ctx, w, q, act_scale, M, y);
    }
    if (M <= 8 && !getenv("BITNET_B200_NO_STREAM")) {  // decode fast path (gemv_stream.cu); 1 = shape not covered
        const int rc = launch_gemv_fast(ctx, w, M, /*PRO_Q8*/ 0, q, act_scale, nullptr, nullptr, 0.f, /*EPI_STORE*/ 0,
                                          y, nullptr, acc_out, nullptr, nullptr, nullptr,
                                          getenv("BITNET_B200_PDL_API") != nullptr);
        if (rc != 1) {
            g_last_linear_kernel = "gemv_fast";
            return rc;
        }
    }
    g_last_linear_kernel = "gemv";
    return launch_gemv_i8(ctx, w, q, act_scale, M, y, acc_out, nullptr);
}

int bitnet_b200_linear_i8_swiglu(bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const int8_t* q, const float* act_scale,
                                 int M, float* h) {
    if (!ctx || !w || !q || !act_scale || !h) return fail("linear_i8_swiglu: NULL argument");
    if (M <= 0) return fail("linear_i8_swiglu: M must be > 0");
    if (w->device != ctx->device) return fail("linear_i8_swiglu: weight lives on another device");
    if (w->rows & 1) return fail("linear_i8_swiglu: the weight needs interleaved (gate, up) rows, i.e. an even row count");
    B200_CUDA(cudaSetDevice(ctx->device));
    if (!gemm_i8_supported(w, M))
        return fail("linear_i8_swiglu: batched / prefill path only (QK256, M >= 9, cols % 128 == 0); decode uses fused_linear(EPI_SWIGLU)");
    return launch_gemm_i8(ctx, w, q, act_scale, M, nullptr, nullptr, h);
}

int bitnet_b200_fused_linear(bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const float* x,
                             const float* gamma, float eps, int M, int epilogue, float* y,
                             const float* residual, int8_t* q_out, float* scale_out) {
    if (!ctx || !w || !x || !y) return fail("fused_linear: NULL argument");
    if (M <= 0 || M > 8) return fail("fused_linear: M must be in 1..8 (decode regime)");
    if (epilogue < 0 || epilogue > 2) return fail("fused_linear: invalid epilogue");
    if (epilogue == BITNET_B200_EPI_RESIDUAL && !residual) return fail("fused_linear: residual is NULL");
    if (gamma && (!(eps > 0.f) || !isfinite(eps))) return fail("fused_linear: eps must be positive and finite");
    if (w->device != ctx->device) return fail("fused_linear: weight lives on another device");
    B200_CUDA(cudaSetDevice(ctx->device));
    const int rc = launch_gemv_fast(ctx, w, M, gamma ? 2 : 1, nullptr, nullptr, x, gamma, eps, epilogue, y, residual,
                                      nullptr, q_out, scale_out, nullptr, false);
    if (rc == 1) return fail("fused_linear: shape not covered by the streaming GEMV (need K % 64 == 0, BitNet32-F16 K % 256 == 0, aligned pointers)");
    return rc;
}

int bitnet_b200_linear_f32(bitnet_b200_ctx* ctx, const bitnet_b200_weight* w, const float* x,
                           int M, float* y) {
    if (!ctx || !w || !x || !y) return fail("linear_f32: NULL argument");
    if (M <= 0) return fail("linear_f32: M must be > 0");
    if (w->device != ctx->device) return fail("linear_f32: weight lives on another device");
    B200_CUDA(cudaSetDevice(ctx->device));
    return launch_gemv_f32(ctx, w, x, M, y);
}

// ======================================================================================
// Part 2 -- gemv_qk256 on host slices
// ======================================================================================
int bitnet_b200_gemv_qk256(const uint8_t* qs, size_t qs_len, const float* x, size_t x_len,
                           float* y, size_t y_len, size_t rows, size_t cols,
                           size_t row_stride_bytes) {
    try {
        char buf[160];
        if (y_len != rows) {  // i2s_qk256.rs:301-303
            snprintf(buf, sizeof buf, "I2S_QK256: y_out length %zu != rows %zu", y_len, rows);
            return fail(buf);
        }
        if (x_len < cols) {  // :304-306
            snprintf(buf, sizeof buf, "I2S_QK256: x length %zu < cols %zu", x_len, cols);
            return fail(buf);
        }
        if (qs_len < rows * row_stride_bytes) {  // :308-311
            snprintf(buf, sizeof buf, "I2S_QK256: data too short: %zu < %zu", qs_len,
                     rows * row_stride_bytes);
            return fail(buf);
        }
        if (rows == 0) return 0;
        if (!qs || !x || !y) return fail("I2S_QK256: NULL pointer");
        if (row_stride_bytes != div_ceil(cols, 256) * 64) {  // debug_assert :203-210
            snprintf(buf, sizeof buf, "I2S_QK256: row bytes mismatch: got %zu, expected %zu for %zu cols",
                     row_stride_bytes, div_ceil(cols, 256) * 64, cols);
            return fail(buf);
        }
        int dev;
        if (need_device(&dev)) return -1;
        StreamGuard sg;  // own non-blocking stream: concurrent host threads do not serialise
        if (sg.create()) return -1;
        bitnet_b200_ctx* ctx = nullptr;
        if (bitnet_b200_ctx_create(dev, sg.s, &ctx)) return -1;
        std::unique_ptr<bitnet_b200_ctx, void (*)(bitnet_b200_ctx*)> cg(ctx, bitnet_b200_ctx_destroy);
        bitnet_b200_weight* w = nullptr;
        if (bitnet_b200_weight_create(ctx, qs, rows * row_stride_bytes, BITNET_B200_LAYOUT_QK256, rows,
                                      cols, nullptr, 0, &w))
            return -1;
        std::unique_ptr<bitnet_b200_weight, void (*)(bitnet_b200_weight*)> wg(w, bitnet_b200_weight_destroy);
        DevBuf dx, dy;
        if (dx.alloc(cols * 4) || dy.alloc(rows * 4)) return -1;
        B200_CUDA(cudaMemcpyAsync(dx.p, x, cols * 4, cudaMemcpyHostToDevice, ctx->stream));
        if (launch_gemv_f32(ctx, w, dx.as<float>(), 1, dy.as<float>())) return -1;
        B200_CUDA(cudaMemcpyAsync(y, dy.p, rows * 4, cudaMemcpyDeviceToHost, ctx->stream));
        B200_CUDA(cudaStreamSynchronize(ctx->stream));
        return 0;
    } catch (const std::exception& e) {
        return fail(e.what());
    } catch (...) {
        return fail("I2S_QK256: unknown error");
    }
}

int bitnet_b200_i2s_matmul_f32(const float* act, const uint8_t* packed, const float* scales,
                               float* out, int m, int n, int k, int block_size) {
    try {
        if (!act || !packed || !scales || !out) return fail("i2s_matmul: NULL pointer");
        if (block_size <= 0) return fail("block_size must be > 0");  // quantized_matmul.rs:214-218
        if (m <= 0 || n <= 0 || k <= 0) {
            char buf[96];
            snprintf(buf, sizeof buf, "dimensions must be > 0: m=%d, n=%d, k=%d", m, n, k);
            return fail(buf);
        }
        int dev;
        if (need_device(&dev)) return -1;
        StreamGuard sg;
        if (sg.create()) return -1;
        const size_t packed_k = ((size_t)k + 3) / 4, nbk = ((size_t)k + block_size - 1) / block_size;
        DevBuf da, dw, ds, dout;
        if (da.alloc((size_t)m * k * 4) || dw.alloc(packed_k * n) || ds.alloc((size_t)n * nbk * 4) ||
            dout.alloc((size_t)m * n * 4))
            return -1;
        B200_CUDA(cudaMemcpyAsync(da.p, act, (size_t)m * k * 4, cudaMemcpyHostToDevice, sg.s));
        B200_CUDA(cudaMemcpyAsync(dw.p, packed, packed_k * n, cudaMemcpyHostToDevice, sg.s));
        B200_CUDA(cudaMemcpyAsync(ds.p, scales, (size_t)n * nbk * 4, cudaMemcpyHostToDevice, sg.s));
        if (launch_i2s_matmul_f32(sg.s, da.as<float>(), dw.as<uint8_t>(), ds.as<float>(), dout.as<float>(), m,
                                  n, k, block_size))
            return -1;
        B200_CUDA(cudaMemcpyAsync(out, dout.p, (size_t)m * n * 4, cudaMemcpyDeviceToHost, sg.s));
        B200_CUDA(cudaStreamSynchronize(sg.s));
        return 0;
    } catch (const std::exception& e) {
        return fail(e.what());
    } catch (...) {
        return fail("i2s_matmul: unknown error");
    }
}

// ---- legacy QK256 entry points (crates/bitnet-quantization/src/qk256_dispatch.rs) ----------------
static int legacy_check(size_t out_len, size_t rows, size_t cols, size_t packed_len, size_t scales_len, size_t act_len) {
    // the reference asserts (panics) with these messages, qk256_dispatch.rs:28-38; here they are errors
    if (out_len != rows) return fail("Output length mismatch");
    if (act_len != cols) return fail("Activation length mismatch");
    if (cols % 256) return fail("Cols must be multiple of QK256=256");
    if (packed_len != rows * cols / 4) return fail("Packed weight size mismatch");
    if (scales_len != rows * (cols / 256)) return fail("Scales length mismatch");
    return 0;
}

int bitnet_b200_qk256_gemv(float* output, size_t out_len, size_t rows, size_t cols, const uint8_t* packed,
                           size_t packed_len, const float* scales, size_t scales_len, const float* activations,
                           size_t act_len) {
    if (legacy_check(out_len, rows, cols, packed_len, scales_len, act_len)) return -1;
    (void)scales;  // validated, not consumed: the GGML no-scale format (qk256_dispatch.rs:14-16)
    return bitnet_b200_gemv_qk256(packed, packed_len, activations, act_len, output, out_len, rows, cols, cols / 256 * 64);
}

int bitnet_b200_qk256_gemv_scalar(float* output, size_t out_len, size_t rows, size_t cols, const uint8_t* packed,
                                  size_t packed_len, const float* scales, size_t scales_len, const float* activations,
                                  size_t act_len) {
    try {
        if (legacy_check(out_len, rows, cols, packed_len, scales_len, act_len)) return -1;
        if (rows == 0) return 0;
        if (!output || !packed || !scales || !activations) return fail("qk256_gemv_scalar: NULL pointer");
        if (rows > 0x7fffffffu || cols / 256 > 0x7fffffffu) return fail("qk256_gemv_scalar: dimensions too large");
        int dev;
        if (need_device(&dev)) return -1;
        StreamGuard sg;
        if (sg.create()) return -1;
        DevBuf dw, ds, dx, dy;
        if (dw.alloc(packed_len) || ds.alloc(scales_len * 4) || dx.alloc(cols * 4) || dy.alloc(rows * 4)) return -1;
        B200_CUDA(cudaMemcpyAsync(dw.p, packed, packed_len, cudaMemcpyHostToDevice, sg.s));
        B200_CUDA(cudaMemcpyAsync(ds.p, scales, scales_len * 4, cudaMemcpyHostToDevice, sg.s));
        B200_CUDA(cudaMemcpyAsync(dx.p, activations, cols * 4, cudaMemcpyHostToDevice, sg.s));
        if (launch_qk256_legacy(sg.s, dw.as<uint8_t>(), ds.as<float>(), dx.as<float>(), dy.as<float>(), (int)rows,
                                (int)(cols / 256)))
            return -1;
        B200_CUDA(cudaMemcpyAsync(output, dy.p, rows * 4, cudaMemcpyDeviceToHost, sg.s));
        B200_CUDA(cudaStreamSynchronize(sg.s));
        return 0;
    } catch (const std::exception& e) {
        return fail(e.what());
    } catch (...) {
        return fail("qk256_gemv_scalar: unknown error");
    }
}

const char* bitnet_b200_version(void) { return "bitnet_b200 0.1.0 (sm_100a)"; }
uint64_t bitnet_b200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
const char* bitnet_b200_last_linear_kernel(void) { return b200::g_last_linear_kernel; }

}  // extern "C"
