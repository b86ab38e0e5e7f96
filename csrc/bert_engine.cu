// Native step executor for the fused BERT trainer: the forward / backward launch sequences as two C calls.
// Python fills a `fedb200_bert_plan` with raw pointers once; per step it passes only (T, B, max_len, seed).
// This is the host-side runtime around the kernels ("executors ... native where it matters"):
// ~270 launches per training step issued back to back from C++ instead of through the interpreter, which is
// what keeps small batches (the 4-sample tail batch of the reference's 100-sample epoch) GPU-bound.
// Sequence = HF BertForSequenceClassification forward + its exact backward (see fedb200/models/fused_bert.py
// for the op-by-op description; that file keeps a Python twin of this sequence for debugging).
#include "common.cuh"
#include "../include/fedb200.h"

namespace fedb {

static inline unsigned long long op_seed(unsigned long long step_seed, int layer, int which) {
    unsigned long long x = step_seed * 0x9E3779B97F4A7C15ULL + (unsigned long long)(layer * 8 + which) * 0xD1B54A32D192ED03ULL;
    x ^= x >> 29;
    return x & 0x7FFFFFFFFFFFFFFFULL;
}

// split-K factor for a weight-gradient GEMM: fill the persistent grid with little wave-quantisation loss
static int wgrad_splits(int out_rows, int out_cols, int tokens, int sms) {
    const int bn = out_cols >= 256 ? 256 : (out_cols >= 128 ? 128 : 64);
    const int tiles = ((out_rows + 127) / 128) * ((out_cols + bn - 1) / bn);
    const int kmax = (tokens + 63) / 64 > 0 ? (tokens + 63) / 64 : 1;
    int best = 1;
    double best_eff = 0.0;
    for (int s = 1; s <= 32 && s <= kmax; ++s) {
        const int total = tiles * s;
        const double eff = (double)total / (double)(((total + sms - 1) / sms) * sms);
        if (total >= (int)(0.9 * sms) && eff >= 0.9) return s;
        if (eff > best_eff) { best = s; best_eff = eff; }
    }
    return best;
}

}  // namespace fedb

using namespace fedb;


extern "C" int fedb200_bert_forward(const fedb200_bert_plan* p, int T, int B, int max_len, int train,
                                    unsigned long long step_seed, int sm_count_hint, void* stream) {
    FEDB_REQUIRE(p != nullptr && T > 0 && B > 0, "bert_forward: empty batch");
    FEDB_REQUIRE(p->L >= 1 && p->L <= FEDB200_MAX_LAYERS, "bert_forward: %d layers (max %d)", p->L, FEDB200_MAX_LAYERS);
    const int H = p->H, I = p->I;
    const int E = p->E > 0 ? p->E : H;
    const float ph = train ? p->p_hidden : 0.f, pa = train ? p->p_attn : 0.f;
    const float pe = train ? p->p_embed : 0.f, phead = train ? p->p_head : 0.f;
    const float scale = 1.0f / sqrtf((float)(H / p->heads));
    void* emb_dst = E == H ? p->layer[0].x_in : p->emb_out;
    FEDB_TRY(fedb200_embed_fwd(p->ids, p->pos_ids, p->tt, p->word, p->pos, p->type, p->emb_lnw, p->emb_lnb, emb_dst,
                               p->emb_mean, p->emb_rstd, T, E, p->ln_eps, op_seed(step_seed, 99, 0), pe, stream));
    if (E != H)      // ALBERT: embedding_hidden_mapping_in
        FEDB_TRY(fedb200_gemm_bf16(0, p->emb_out, E, p->map_w, E, T, H, E, FEDB200_EPI_BIAS, p->layer[0].x_in, nullptr, H, p->map_b,
                                   nullptr, 0, p->act, 0, 0.f, 1, 0, stream));
    for (int l = 0; l < p->L; ++l) {
        const fedb200_bert_layer& a = p->layer[l];
        FEDB_TRY(fedb200_gemm_bf16(0, a.x_in, H, a.wqkv, H, T, 3 * H, H, FEDB200_EPI_BIAS, a.qkv, nullptr, 3 * H, a.bqkv,
                                   nullptr, 0, p->act, 0, 0.f, 1, 0, stream));
        FEDB_TRY(fedb200_attn_fwd(a.qkv, a.ctx, a.lse, p->cu_seqlens, B, T, p->heads, max_len, scale, op_seed(step_seed, l, 1),
                                  pa, train ? a.drop_bits : nullptr, stream));
        FEDB_TRY(fedb200_gemm_bf16(0, a.ctx, H, a.wo, H, T, H, H, FEDB200_EPI_BIAS_DROP_RES, a.z1, nullptr, H, a.bo, a.x_in,
                                   H, p->act, op_seed(step_seed, l, 2), ph, 1, 0, stream));
        FEDB_TRY(fedb200_layernorm_fwd(a.z1, a.ln1w, a.ln1b, a.a1, a.mean1, a.rstd1, T, H, p->ln_eps, stream));
        FEDB_TRY(fedb200_gemm_bf16(0, a.a1, H, a.w1, H, T, I, H, FEDB200_EPI_BIAS_GELU, train ? a.u : nullptr, a.g, I, a.b1, nullptr, 0, p->act, 0,
                                   0.f, 1, 0, stream));
        FEDB_TRY(fedb200_gemm_bf16(0, a.g, I, a.w2, I, T, H, I, FEDB200_EPI_BIAS_DROP_RES, a.z2, nullptr, H, a.b2, a.a1, H,
                                   p->act, op_seed(step_seed, l, 3), ph, 1, 0, stream));
        FEDB_TRY(fedb200_layernorm_fwd(a.z2, a.ln2w, a.ln2b, a.x_out, a.mean2, a.rstd2, T, H, p->ln_eps, stream));
    }
    FEDB_TRY(fedb200_head_fwd(p->layer[p->L - 1].x_out, p->cu_seqlens, p->Wp, p->bp, p->Wc, p->bc, p->labels, p->pooled,
                              p->pooled_drop, p->logits, p->loss, p->pred, p->dlogits, B, H, p->C, 1.0f / (float)B,
                              op_seed(step_seed, 98, 0), phead, stream));
    return 0;
}

extern "C" int fedb200_bert_backward(const fedb200_bert_plan* p, int T, int B, int max_len,
                                     unsigned long long step_seed, int sm_count_hint, void* stream) {
    FEDB_REQUIRE(p != nullptr && T > 0 && B > 0, "bert_backward: empty batch");
    const int H = p->H, I = p->I;
    const int sms = sm_count_hint > 0 ? sm_count_hint : sm_count();
    const int E = p->E > 0 ? p->E : H;
    const float ph = p->p_hidden, pa = p->p_attn;
    const float scale = 1.0f / sqrtf((float)(H / p->heads));
    cudaStream_t s = (cudaStream_t)stream;
    void* df = ph > 0.f ? p->dfb : p->dzb;          // dense-output gradient seen through its dropout mask
    FEDB_CHECK(cudaMemsetAsync(p->dx, 0, (size_t)T * H * 2, s));
    FEDB_TRY(fedb200_head_bwd(p->layer[p->L - 1].x_out, p->cu_seqlens, p->Wp, p->Wc, p->pooled, p->pooled_drop, p->dlogits,
                              p->dpre, p->g_Wp, p->g_bp, p->g_Wc, p->g_bc, p->dx, B, H, p->C, op_seed(step_seed, 98, 0),
                              p->p_head, stream));
    for (int l = p->L - 1; l >= 0; --l) {
        const fedb200_bert_layer& a = p->layer[l];
        // ---- FFN block ----
        FEDB_TRY(fedb200_layernorm_bwd(p->dx, a.z2, a.mean2, a.rstd2, a.ln2w, p->dzb, p->dfb, a.g_ln2w, a.g_ln2b, a.g_b2, T, H,
                                       op_seed(step_seed, l, 3), ph, stream));
        FEDB_TRY(fedb200_gemm_bf16(1, df, H, a.w2, I, T, I, H, FEDB200_EPI_MUL_GELU_GRAD, p->du, nullptr, I, nullptr, a.u, I,
                                   p->act, 0, 0.f, 1, 0, stream));
        FEDB_TRY(fedb200_gemm_bf16(2, df, H, a.g, I, H, I, T, FEDB200_EPI_ATOMIC_F32, a.g_w2, nullptr, I, nullptr, nullptr, 0,
                                   p->act, 0, 0.f, wgrad_splits(H, I, T, sms), 0, stream));
        FEDB_TRY(fedb200_colsum_bf16(p->du, a.g_b1, T, I, I, stream));
        FEDB_TRY(fedb200_gemm_bf16(1, p->du, I, a.w1, H, T, H, I, FEDB200_EPI_ADD, p->dx, nullptr, H, nullptr, p->dzb, H, p->act,
                                   0, 0.f, 1, 0, stream));
        FEDB_TRY(fedb200_gemm_bf16(2, p->du, I, a.a1, H, I, H, T, FEDB200_EPI_ATOMIC_F32, a.g_w1, nullptr, H, nullptr, nullptr, 0,
                                   p->act, 0, 0.f, wgrad_splits(I, H, T, sms), 0, stream));
        // ---- attention block ----
        FEDB_TRY(fedb200_layernorm_bwd(p->dx, a.z1, a.mean1, a.rstd1, a.ln1w, p->dzb, p->dfb, a.g_ln1w, a.g_ln1b, a.g_bo, T, H,
                                       op_seed(step_seed, l, 2), ph, stream));
        FEDB_TRY(fedb200_gemm_bf16(1, df, H, a.wo, H, T, H, H, FEDB200_EPI_ADD, p->dctx, nullptr, H, nullptr, nullptr, 0, p->act,
                                   0, 0.f, 1, 0, stream));
        FEDB_TRY(fedb200_gemm_bf16(2, df, H, a.ctx, H, H, H, T, FEDB200_EPI_ATOMIC_F32, a.g_wo, nullptr, H, nullptr, nullptr, 0,
                                   p->act, 0, 0.f, wgrad_splits(H, H, T, sms), 0, stream));
        FEDB_TRY(fedb200_attn_bwd(a.qkv, a.ctx, p->dctx, a.lse, p->delta, p->dqkv, p->dq_accum, a.g_bqkv, p->cu_seqlens, B, T,
                                  p->heads, max_len, scale, op_seed(step_seed, l, 1), pa, a.drop_bits, stream));
        FEDB_TRY(fedb200_gemm_bf16(1, p->dqkv, 3 * H, a.wqkv, H, T, H, 3 * H, FEDB200_EPI_ADD, p->dx, nullptr, H, nullptr, p->dzb,
                                   H, p->act, 0, 0.f, 1, 0, stream));
        FEDB_TRY(fedb200_gemm_bf16(2, p->dqkv, 3 * H, a.x_in, H, 3 * H, H, T, FEDB200_EPI_ATOMIC_F32, a.g_wqkv, nullptr, H, nullptr,
                                   nullptr, 0, p->act, 0, 0.f, wgrad_splits(3 * H, H, T, sms), 0, stream));
    }
    const void* d_emb = p->dx;
    if (E != H) {    // back through the ALBERT embedding projection: bias, weight and input gradients
        FEDB_TRY(fedb200_colsum_bf16(p->dx, p->g_map_b, T, H, H, stream));
        FEDB_TRY(fedb200_gemm_bf16(2, p->dx, H, p->emb_out, E, H, E, T, FEDB200_EPI_ATOMIC_F32, p->g_map_w, nullptr, E, nullptr,
                                   nullptr, 0, p->act, 0, 0.f, wgrad_splits(H, E, T, sms), 0, stream));
        FEDB_TRY(fedb200_gemm_bf16(1, p->dx, H, p->map_w, E, T, E, H, FEDB200_EPI_ADD, p->d_emb, nullptr, E, nullptr, nullptr, 0,
                                   p->act, 0, 0.f, 1, 0, stream));
        d_emb = p->d_emb;
    }
    FEDB_TRY(fedb200_embed_bwd(d_emb, p->ids, p->pos_ids, p->tt, p->word, p->pos, p->type, p->emb_lnw, p->emb_mean, p->emb_rstd,
                               p->g_word, p->g_pos, p->g_type, p->g_emb_lnw, p->g_emb_lnb, T, E, p->type_vocab,
                               op_seed(step_seed, 99, 0), p->p_embed, stream));
    return 0;
}
