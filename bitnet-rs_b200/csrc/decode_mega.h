// decode_mega.h -- parameters of the persistent decode-step kernel (decode_mega.cu), shared with
// the model object in decode.cu.
#pragma once
#include "common.cuh"

namespace b200 {

struct MegaLayer {
    const uint8_t *wqkv, *wo, *wgu, *wdown;  // fragment-major code planes [N/16][K/32][32 lanes] words (mega_repack_weight)
    const float *attn_norm, *ffn_norm;
};

// one ternary projection of a layer (qkv, o, gate/up, down), shape-derived constants filled by mega_plan
struct MegaOp {
    int N, K, KC;      // rows, columns, k-chunks (K / 32)
    int T;             // consecutive 16-row tiles per ring slot (one bulk copy, one barrier pair)
    int tile_bytes;    // bytes of one 16-row tile in the fragment-major copy (codes [+ f16 block scales])
    int tq, tr;        // tiles per CTA: the first tr CTAs own tq + 1 contiguous tiles, the others tq
};

struct MegaParams {
    MegaOp op[4];
    const MegaLayer* layers;
    const __half* emb;
    const float *final_norm, *rope_cos, *rope_sin;
    float *kc, *vc;
    int *pos, *tokens, *next;
    uint2 *x, *qkv, *att, *gu;  // phase outputs as (f32 bits, epoch) pairs: each 8-byte store carries its own ready flag
    uint2* attp;             // attention partials of split contexts: [B, NH, 8 splits, D + 2] pairs (sum e*v | max | sum e)
    int att_split;           // 1 = split long contexts over several CTAs per head (knob BITNET_B200_MEGA_NOSPLIT)
    float *logits, *part_val;
    int* part_idx;
    const int8_t* q_fixed;   // mode 1: fixed int8 rows [B, max(H, F, NH*D)]
    const float* s_fixed;    // mode 1: their scales
    unsigned* bar;           // [2] grid-barrier counters, used alternately per launch
    int parity;
    int L, B, H, F, NH, NKV, D, V, max_ctx;
    float eps;
    uint32_t lut;
    uint32_t epoch;          // epoch of this launch's first phase output (advances by 5L + 1 per launch)
    int barriers;            // 1 = keep the grid barrier after every phase as well (diagnostics knob BITNET_B200_MEGA_BARRIERS)
    int prefetch;            // 1 = L2 prefetches inside the grid barriers (diagnostics knob BITNET_B200_MEGA_NOPF)
    int mode;                // 0 = full decode step; 1 = the 4L ternary GEMVs only, no dependencies
    int n_slots, act_bytes, attn_floats;
    int layout;              // 0 = QK256, 1 = BitNet32-F16 (per-32-column f16 block scales travel with the codes)
    int slot_bytes;          // ring slot size (mega_plan)
    int off_red, off_act, off_attn, off_stage;  // shared-memory offsets of the shape-dependent pieces (mega_plan)
    int red_ld;              // ints per warp in the partial-sum buffer: (tiles per CTA at most) * 16 * B
    unsigned long long* prof;  // optional diagnostics buffer (decode_mega.cu: prof_stamp / wtrace), or nullptr
    int prof_cta;              // CTA whose warps record the finer trace
};

// decode_mega.cu
size_t mega_plan(MegaParams* p, int num_sms);  // 0 = shape not covered
int mega_repack_weight(const uint8_t* codes, const __half* scales, int scale_abs_clamp, int N, int K, uint8_t* out, cudaStream_t st);
size_t mega_repacked_bytes(int layout, int N, int K);
int launch_decode_mega(const MegaParams& p, size_t smem, int num_sms, cudaStream_t st);

}  // namespace b200
