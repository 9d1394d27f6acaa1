/*
 * model.cpp -- CPU restatement of a BitNet-b1.58-2B-4T-shaped decoder built from the oracle's
 * pinned pieces. TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * The block graph is the reference's (crates/bitnet-transformer/src/lib.rs:977-1134 pre-norm
 * residual blocks, :134-184 split-half RoPE, :410-508 GQA attention with 1/sqrt(head_dim)
 * scaling and softmax over keys, :756-789 down(silu(gate)*up), :1599-1630 tied-embedding
 * logits) with the two substitutions SURVEY.md §3.1 prescribes for this build: true RMSNorm
 * (cpu::rms_norm, crates/bitnet-kernels/src/cpu/layer_norm.rs:118-137) in place of candle's
 * LayerNorm, and every linear computed as
 *     quantize_symmetric_i8 (cpu/quantize.rs:32-49) -> packed I2_S matmul -> scale,
 * i.e. the composed int8 linear of oracle.h. The reference holds no golden tokens for
 * random-init weights (SURVEY §4 "End-to-end"), so parity here is CPU-oracle-model vs
 * GPU-model on the same graph: PARITY UNPINNED at model level, pinned piecewise.
 */
#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

#include "oracle.h"

namespace {

struct Layer {
    std::vector<uint8_t> wq, wk, wv, wo, wgate, wup, wdown;
    std::vector<float> attn_norm, ffn_norm;
};

struct Model {
    oracle_model_config cfg;
    std::vector<Layer> layers;
    std::vector<uint16_t> emb;  // [vocab, hidden] f16 bits
    std::vector<float> final_norm;
    std::vector<float> rope_cos, rope_sin;  // [max_ctx, head_dim/2]
    // per sequence
    std::vector<std::vector<float>> kc, vc;  // [n_layers * n_kv * max_ctx * head_dim]
    std::vector<int> pos;
};

template <typename F>
void parallel_rows(size_t rows, int threads, F fn) {
    if (threads <= 1 || rows < 64) {
        fn(0, rows);
        return;
    }
    std::vector<std::thread> pool;
    size_t per = (rows + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        size_t r0 = std::min(rows, (size_t)t * per), r1 = std::min(rows, r0 + per);
        if (r0 < r1) pool.emplace_back([=] { fn(r0, r1); });
    }
    for (auto& th : pool) th.join();
}

// y[r] = act_scale * sum_blk s_blk * (int dot), the composed int8 linear (oracle.h)
void linear_i8(const Model& m, const std::vector<uint8_t>& w, size_t rows, size_t cols,
               const int8_t* q, float act_scale, float* y, int threads) {
    if (m.cfg.layout == 0) {
        const size_t stride = (cols + 255) / 256 * 64;
        std::vector<float> x(cols);
        for (size_t c = 0; c < cols; ++c) x[c] = (float)q[c];
        parallel_rows(rows, threads, [&](size_t r0, size_t r1) {
            // the reference kernel itself; exact integers under the F5 bound
            ref_gemv_qk256(w.data() + r0 * stride, (r1 - r0) * stride, x.data(), cols, y + r0,
                           r1 - r0, r1 - r0, cols, stride);
            for (size_t r = r0; r < r1; ++r) y[r] = act_scale * (float)(int32_t)y[r];
        });
    } else {
        static const int8_t MAP_A[4] = {-2, -1, 1, 2};
        parallel_rows(rows, threads, [&](size_t r0, size_t r1) {
            oracle_linear_bitnet32_i8(w.data() + r0 * (cols / 32) * 10, r1 - r0, cols, q, &act_scale,
                                      1, MAP_A, 0, nullptr, y + r0);
        });
    }
}

const float* f16_table() {
    static std::vector<float> t = [] {
        std::vector<float> v(65536);
        for (uint32_t i = 0; i < 65536; ++i) v[i] = ref_f16_to_f32((uint16_t)i);
        return v;
    }();
    return t.data();
}

void quant(const float* x, size_t n, std::vector<int8_t>& q, float& s) {
    q.resize(n);
    ref_quantize_symmetric_i8(x, n, 8, q.data(), &s);
}

}  // namespace

extern "C" {

void* oracle_model_create(const oracle_model_config* cfg, int n_seqs) {
    auto* m = new Model();
    m->cfg = *cfg;
    m->layers.resize((size_t)cfg->n_layers);
    const size_t kv = (size_t)cfg->n_layers * cfg->n_kv_heads * cfg->max_ctx * cfg->head_dim;
    m->kc.assign((size_t)n_seqs, std::vector<float>(kv, 0.f));
    m->vc.assign((size_t)n_seqs, std::vector<float>(kv, 0.f));
    m->pos.assign((size_t)n_seqs, 0);
    return m;
}

void oracle_model_destroy(void* h) { delete static_cast<Model*>(h); }

static size_t wbytes(const Model& m, size_t rows, size_t cols) {
    return m.cfg.layout == 0 ? rows * ((cols + 255) / 256 * 64) : rows * (cols / 32) * 10;
}

int oracle_model_set_layer(void* h, int l, const uint8_t* wq, const uint8_t* wk, const uint8_t* wv,
                           const uint8_t* wo, const uint8_t* wgate, const uint8_t* wup,
                           const uint8_t* wdown, const float* attn_norm, const float* ffn_norm) {
    auto& m = *static_cast<Model*>(h);
    const auto& c = m.cfg;
    if (l < 0 || l >= c.n_layers) return -1;
    auto& L = m.layers[(size_t)l];
    const size_t H = (size_t)c.hidden, F = (size_t)c.ffn, QD = (size_t)c.n_heads * c.head_dim,
                 KD = (size_t)c.n_kv_heads * c.head_dim;
    L.wq.assign(wq, wq + wbytes(m, QD, H));
    L.wk.assign(wk, wk + wbytes(m, KD, H));
    L.wv.assign(wv, wv + wbytes(m, KD, H));
    L.wo.assign(wo, wo + wbytes(m, H, QD));
    L.wgate.assign(wgate, wgate + wbytes(m, F, H));
    L.wup.assign(wup, wup + wbytes(m, F, H));
    L.wdown.assign(wdown, wdown + wbytes(m, H, F));
    L.attn_norm.assign(attn_norm, attn_norm + H);
    L.ffn_norm.assign(ffn_norm, ffn_norm + H);
    return 0;
}

int oracle_model_set_embedding(void* h, const uint16_t* emb_f16, const float* final_norm) {
    auto& m = *static_cast<Model*>(h);
    m.emb.assign(emb_f16, emb_f16 + (size_t)m.cfg.vocab * m.cfg.hidden);
    m.final_norm.assign(final_norm, final_norm + m.cfg.hidden);
    return 0;
}

int oracle_model_set_rope(void* h, const float* cos_t, const float* sin_t) {
    auto& m = *static_cast<Model*>(h);
    const size_t n = (size_t)m.cfg.max_ctx * (m.cfg.head_dim / 2);
    m.rope_cos.assign(cos_t, cos_t + n);
    m.rope_sin.assign(sin_t, sin_t + n);
    return 0;
}

void oracle_model_reset(void* h) {
    auto& m = *static_cast<Model*>(h);
    std::fill(m.pos.begin(), m.pos.end(), 0);
}

// One decode step of sequence `seq`; returns the greedy token (lowest index on ties),
// optionally the logits [vocab] and the final hidden state [hidden].
int oracle_model_step(void* h, int seq, int token, float* logits_out, float* hidden_out,
                      int threads) {
    auto& m = *static_cast<Model*>(h);
    const auto& c = m.cfg;
    const size_t H = (size_t)c.hidden, F = (size_t)c.ffn, D = (size_t)c.head_dim,
                 NH = (size_t)c.n_heads, NKV = (size_t)c.n_kv_heads, QD = NH * D, KD = NKV * D;
    const size_t group = NH / NKV;
    const int pos = m.pos[(size_t)seq];
    if (pos >= c.max_ctx || token < 0 || token >= c.vocab) return -1;

    std::vector<float> x(H), hn(H), qv(QD), kv(KD), vv(KD), att(QD), o(H), gate(F), up(F), act(F),
        down(H), scores((size_t)pos + 1);
    std::vector<int8_t> q8;
    float s;
    for (size_t i = 0; i < H; ++i) x[i] = ref_f16_to_f32(m.emb[(size_t)token * H + i]);

    for (int l = 0; l < c.n_layers; ++l) {
        const Layer& L = m.layers[(size_t)l];
        ref_rms_norm(x.data(), H, L.attn_norm.data(), H, c.rms_eps, hn.data());
        quant(hn.data(), H, q8, s);
        linear_i8(m, L.wq, QD, H, q8.data(), s, qv.data(), threads);
        linear_i8(m, L.wk, KD, H, q8.data(), s, kv.data(), threads);
        linear_i8(m, L.wv, KD, H, q8.data(), s, vv.data(), threads);
        // RoPE, split-half layout (transformer lib.rs:134-184), tables shared with the GPU
        const float* cs = m.rope_cos.data() + (size_t)pos * (D / 2);
        const float* sn = m.rope_sin.data() + (size_t)pos * (D / 2);
        auto rope = [&](float* v, size_t heads) {
            for (size_t hd = 0; hd < heads; ++hd)
                for (size_t i = 0; i < D / 2; ++i) {
                    float a = v[hd * D + i], b = v[hd * D + i + D / 2];
                    v[hd * D + i] = a * cs[i] - b * sn[i];
                    v[hd * D + i + D / 2] = a * sn[i] + b * cs[i];
                }
        };
        rope(qv.data(), NH);
        rope(kv.data(), NKV);
        float* kc = m.kc[(size_t)seq].data() + (size_t)l * NKV * c.max_ctx * D;
        float* vc = m.vc[(size_t)seq].data() + (size_t)l * NKV * c.max_ctx * D;
        for (size_t g = 0; g < NKV; ++g) {
            memcpy(kc + (g * c.max_ctx + pos) * D, kv.data() + g * D, D * 4);
            memcpy(vc + (g * c.max_ctx + pos) * D, vv.data() + g * D, D * 4);
        }
        const float scale = 1.0f / sqrtf((float)D);
        for (size_t hd = 0; hd < NH; ++hd) {
            const size_t g = hd / group;
            float mx = -INFINITY;
            for (int p = 0; p <= pos; ++p) {
                const float* kp = kc + (g * c.max_ctx + p) * D;
                float d = 0.f;
                for (size_t i = 0; i < D; ++i) d += qv[hd * D + i] * kp[i];
                scores[(size_t)p] = d * scale;
                mx = fmaxf(mx, scores[(size_t)p]);
            }
            float sum = 0.f;
            for (int p = 0; p <= pos; ++p) {
                scores[(size_t)p] = expf(scores[(size_t)p] - mx);
                sum += scores[(size_t)p];
            }
            const float inv = 1.0f / sum;
            for (size_t i = 0; i < D; ++i) {
                float a = 0.f;
                for (int p = 0; p <= pos; ++p) a += (scores[(size_t)p] * inv) * vc[(g * c.max_ctx + p) * D + i];
                att[hd * D + i] = a;
            }
        }
        quant(att.data(), QD, q8, s);
        linear_i8(m, L.wo, H, QD, q8.data(), s, o.data(), threads);
        for (size_t i = 0; i < H; ++i) x[i] = x[i] + o[i];

        ref_rms_norm(x.data(), H, L.ffn_norm.data(), H, c.rms_eps, hn.data());
        quant(hn.data(), H, q8, s);
        linear_i8(m, L.wgate, F, H, q8.data(), s, gate.data(), threads);
        linear_i8(m, L.wup, F, H, q8.data(), s, up.data(), threads);
        for (size_t i = 0; i < F; ++i) {
            const float g = gate[i];
            act[i] = (g / (1.0f + expf(-g))) * up[i];  // silu(gate) * up
        }
        quant(act.data(), F, q8, s);
        linear_i8(m, L.wdown, H, F, q8.data(), s, down.data(), threads);
        for (size_t i = 0; i < H; ++i) x[i] = x[i] + down[i];
    }
    ref_rms_norm(x.data(), H, m.final_norm.data(), H, c.rms_eps, hn.data());
    if (hidden_out) memcpy(hidden_out, hn.data(), H * 4);

    std::vector<float> logits((size_t)c.vocab);
    const float* h2f = f16_table();
    parallel_rows((size_t)c.vocab, threads, [&](size_t r0, size_t r1) {
        for (size_t r = r0; r < r1; ++r) {
            const uint16_t* row = m.emb.data() + r * H;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;  // tied, pre-transposed embedding in f32
            for (size_t i = 0; i < H; i += 4) {
                a0 += h2f[row[i]] * hn[i];
                a1 += h2f[row[i + 1]] * hn[i + 1];
                a2 += h2f[row[i + 2]] * hn[i + 2];
                a3 += h2f[row[i + 3]] * hn[i + 3];
            }
            logits[r] = (a0 + a1) + (a2 + a3);
        }
    });
    int best = 0;
    for (int r = 1; r < c.vocab; ++r)
        if (logits[(size_t)r] > logits[(size_t)best]) best = r;
    if (logits_out) memcpy(logits_out, logits.data(), (size_t)c.vocab * 4);
    m.pos[(size_t)seq] = pos + 1;
    return best;
}

}  // extern "C"
