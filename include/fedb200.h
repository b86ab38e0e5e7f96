/* fedb200 — C ABI of the hand-written sm_100a kernels (libfedb200.so).
 *
 * Conventions: every function returns 0 on success or a cudaError_t value; fedb200_last_error() gives the
 * message.  Pointers are device pointers unless noted.  `stream` is a cudaStream_t passed as void*.  Nothing
 * here allocates device memory unless its comment says so: callers own all buffers (PyTorch tensors in the
 * Python package).  "Reference" citations point into the research repository this framework re-implements
 * (paths relative to its root); most kernels have no counterpart there because it ships no native code.
 */
#ifndef FEDB200_H
#define FEDB200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FEDB200_ABI_VERSION 1
#define FEDB200_MAX_MIX_SOURCES 64
#define FEDB200_MAX_RANKS 8
#define FEDB200_FLAG_WORDS 128      /* uint32 words of the per-rank flag buffer of the peer-memory collectives */

/* GEMM epilogues */
#define FEDB200_EPI_BIAS 0          /* out = acc + bias                                  (QKV projection)        */
#define FEDB200_EPI_BIAS_DROP_RES 1 /* out = dropout(acc + bias) + aux                   (attn-out, FFN-down)    */
#define FEDB200_EPI_BIAS_GELU 2     /* out = acc + bias ; out2 = gelu(out)               (FFN-up)                */
#define FEDB200_EPI_MUL_GELU_GRAD 3 /* out = acc * gelu'(aux)                            (dgrad through GELU)    */
#define FEDB200_EPI_ADD 4           /* out = acc (+ aux)                                 (dgrad + residual grad) */
#define FEDB200_EPI_ATOMIC_F32 5    /* out_f32 += acc (split-K)                          (wgrad)                 */
#define FEDB200_EPI_DISCARD 6       /* drain TMEM, store nothing (profiling only: isolates the mainloop)          */

const char* fedb200_last_error(void);
int fedb200_abi_version(void);
int fedb200_sm_count(void);
/* programmatic dependent launch between this library's kernels (default on; FEDB200_PDL=0 also disables) */
int fedb200_set_pdl(int on);
/* Reproducible-reduction mode (csrc/determinism.cu): with a workspace registered (256-byte aligned device memory owned by
 * the caller, >= 1 MiB; 256 MiB covers BERT-large), split-K weight gradients, bias / LayerNorm column sums and the
 * embedding scatter store partials there and add them in a fixed order instead of with fp32 atomics, so a training step
 * is bit-reproducible for a given input, seed and SM count.  ws == NULL switches it off (the default). */
int fedb200_set_deterministic(void* ws, size_t bytes);
int fedb200_get_deterministic(void);
/* TMA tensor maps are memoised per (pointer, shape, box): counters of the per-thread cache (host pointers, may be NULL). */
int fedb200_tmap_cache_stats(unsigned long long* hits, unsigned long long* misses);

/* ---- flat-arena kernels ------------------------------------------------------------------------------------ */
/* K13. One pass of HF-semantics AdamW over the whole flat arena (reference: transformers.AdamW python loop,
 * src/Servercase/server_IID_IMDB.py:109,117): p/m/v update, optional bf16 shadow write, optional grad zeroing. */
int fedb200_adamw_flat(float* p, float* g, float* m, float* v, void* shadow_bf16, size_t n, int step, float lr,
                       float beta1, float beta2, float eps, float weight_decay, float grad_scale, int zero_grad,
                       void* stream);
/* K1 (local stage) / K2 / K3 combine: out = sum_j weights[j] * srcs[j]; srcs may be peer-mapped pointers.
 * `srcs` and `weights` are HOST arrays.  Reference: `sum(param)/len(param)`,
 * src/Serverlesscase/serverless_IID_IMDB.py:269, and Flower's FedAvg aggregate (server_IID_IMDB.py:205-209). */
int fedb200_mix_n(const float* const* srcs, const float* weights, int n_src, float* out, void* shadow_bf16, size_t n,
                  void* stream);
int fedb200_colsum_bf16(const void* x, float* out, int M, int N, int ld, void* stream);
int fedb200_cast_f32_bf16(const float* in, void* out, size_t n, void* stream);
int fedb200_delta_bf16(const float* a, const float* b, void* out, size_t n, void* stream);

/* ---- local-trainer kernels ----------------------------------------------------------------------------------- */
/* K8. bf16 GEMM on tcgen05/TMEM fed by TMA. layout 0: A[M,K] B[N,K]; 1: A[M,K] B[K,N]; 2: A[K,M] B[K,N]. */
int fedb200_gemm_bf16(int layout, const void* A, int lda, const void* B, int ldb, int M, int N, int K, int epi,
                      void* out, void* out2, int ldo, const float* bias, const void* aux, int ld_aux, int act,
                      unsigned long long seed, float dropout_p, int splits, int max_ctas, void* stream);
/* GEMM kernel selection: 0 auto (CTA pairs, cta_group::2, when N >= 128), 1 force the 1-CTA kernel, 2 force pairs. */
int fedb200_set_gemm_mode(int mode);
/* K10. LayerNorm over bf16 rows (eps 1e-12 in BERT). bwd also emits dgamma/dbeta and the producing layer's dbias. */
int fedb200_layernorm_fwd(const void* z, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                          int T, int H, float eps, void* stream);
int fedb200_layernorm_bwd(const void* dy, const void* z, const float* mean, const float* rstd, const float* gamma,
                          void* dz, void* df, float* dgamma, float* dbeta, float* dbias, int T, int H,
                          unsigned long long seed, float dropout_p, void* stream);
/* K11. word + position + token-type gather, LayerNorm, dropout; and the scatter backward. */
int fedb200_embed_fwd(const int* ids, const int* pos, const int* tt, const float* word, const float* posw,
                      const float* typew, const float* gamma, const float* beta, void* out, float* mean, float* rstd,
                      int T, int H, float eps, unsigned long long seed, float dropout_p, void* stream);
int fedb200_embed_bwd(const void* dx, const int* ids, const int* pos, const int* tt, const float* word,
                      const float* posw, const float* typew, const float* gamma, const float* mean, const float* rstd,
                      float* gword, float* gpos, float* gtype, float* dgamma, float* dbeta, int T, int H,
                      int type_vocab, unsigned long long seed, float dropout_p, void* stream);
/* K12. pooler + dropout + classifier + softmax cross-entropy, forward and backward. */
int fedb200_head_fwd(const void* x, const int* first_tok, const float* Wp, const float* bp, const float* Wc,
                     const float* bc, const int* labels, float* pooled, float* pooled_drop, float* logits,
                     float* loss, int* pred, float* dlogits, int B, int H, int C, float grad_scale,
                     unsigned long long seed, float dropout_p, void* stream);
int fedb200_head_bwd(const void* x, const int* first_tok, const float* Wp, const float* Wc, const float* pooled,
                     const float* pooled_drop, const float* dlogits, float* dpre, float* gWp, float* gbp, float* gWc,
                     float* gbc, void* dx, int B, int H, int C, unsigned long long seed, float dropout_p,
                     void* stream);
/* K9. fused masked-softmax attention over packed variable-length sequences (head_dim 64), tcgen05 + TMA.
 * qkv is [T, 3H] bf16 (Q | K | V), ctx/dctx [T, H], lse [heads, T] fp32; cu_seqlens is int32 [B+1] on device. */
/* drop_bits (optional, uint32 [heads, T, ceil(max_seqlen / 32)]): the forward stores its dropout keep decisions there and
 * the backward, given the same buffer and max_seqlen, reads them instead of re-hashing (NULL = regenerate from seed). */
int fedb200_attn_fwd(const void* qkv, void* ctx, float* lse, const int* cu_seqlens, int B, int T, int heads,
                     int max_seqlen, float scale, unsigned long long seed, float dropout_p, void* drop_bits, void* stream);
/* dq_accum: scratch of at least ceil(max_seqlen / 128) * T * H * 2 bytes (per-kv-tile bf16 dQ partials; no initialisation).
 * dbias_qkv (optional, fp32 [3H]): the q/k/v bias
 * gradients (column sums of dqkv) are accumulated into it by the finishing pass. */
int fedb200_attn_bwd(const void* qkv, const void* ctx, const void* dctx, const float* lse, float* delta, void* dqkv,
                     float* dq_accum, float* dbias_qkv, const int* cu_seqlens, int B, int T, int heads, int max_seqlen,
                     float scale, unsigned long long seed, float dropout_p, const void* drop_bits, void* stream);

/* ---- native step executor ------------------------------------------------------------------------------------- */
/* One call = the whole forward (or backward) launch sequence of the fused BERT trainer.  The plan holds raw
 * device pointers into the caller's arenas / workspace (filled once by fedb200/models/fused_bert.py), so a
 * training step costs one host call instead of ~270 Python-level launches.  Reference: the body of
 * `train()` / `test()`, src/Servercase/server_IID_IMDB.py:108-135. */
#define FEDB200_MAX_LAYERS 48
typedef struct {
    /* bf16 shadow weights, fp32 master biases / LN params, fp32 gradients */
    const void *wqkv, *wo, *w1, *w2;
    const float *bqkv, *bo, *b1, *b2, *ln1w, *ln1b, *ln2w, *ln2b;
    float *g_wqkv, *g_bqkv, *g_wo, *g_bo, *g_w1, *g_b1, *g_w2, *g_b2, *g_ln1w, *g_ln1b, *g_ln2w, *g_ln2b;
    /* saved activations */
    void *x_in, *x_out, *qkv, *ctx, *z1, *a1, *u, *g, *z2;
    float *mean1, *rstd1, *mean2, *rstd2, *lse;
    void* drop_bits;                 /* attention dropout keep-bits (NULL: backward re-hashes) */
} fedb200_bert_layer;

typedef struct {
    int H, I, L, heads, C, type_vocab, act;   /* act: 0 erf GELU (BERT), 1 tanh GELU (ALBERT gelu_new) */
    int E;                                    /* embedding width; E != H adds the ALBERT 128 -> H projection */
    float ln_eps, p_hidden, p_attn, p_embed, p_head;
    /* embeddings + head (fp32 master) and their gradients */
    const float *word, *pos, *type, *emb_lnw, *emb_lnb, *Wp, *bp, *Wc, *bc;
    float *g_word, *g_pos, *g_type, *g_emb_lnw, *g_emb_lnb, *g_Wp, *g_bp, *g_Wc, *g_bc;
    float *emb_mean, *emb_rstd;
    /* factorised-embedding projection (ALBERT): bf16 shadow weight [H, E], fp32 bias, grads, [T, E] buffers */
    const void* map_w;
    const float* map_b;
    float *g_map_w, *g_map_b;
    void *emb_out, *d_emb;
    /* inputs (packed tokens) */
    const int *ids, *pos_ids, *tt, *cu_seqlens, *labels;
    /* head buffers */
    float *pooled, *pooled_drop, *dpre, *logits, *dlogits, *loss;
    int* pred;
    /* backward workspace */
    void *dx, *dzb, *dfb, *dctx, *du, *dqkv;
    float *dq_accum, *delta;
    fedb200_bert_layer layer[FEDB200_MAX_LAYERS];
} fedb200_bert_plan;

/* `plan` is a HOST pointer.  T packed tokens in B sequences, longest max_len.  `train` enables dropout and saves
 * what backward needs; `step_seed` must be the same for the forward and the backward of one step. */
int fedb200_bert_forward(const fedb200_bert_plan* plan, int T, int B, int max_len, int train,
                         unsigned long long step_seed, int sm_count, void* stream);
int fedb200_bert_backward(const fedb200_bert_plan* plan, int T, int B, int max_len, unsigned long long step_seed,
                          int sm_count, void* stream);

/* ---- federated kernels --------------------------------------------------------------------------------------- */
/* K4. Gram matrix of client updates over a column slice, tcgen05 + TMA straight from (peer) buffers. */
int fedb200_gram_bf16(const void* const* rows_ptrs, int n_ranks, int rows_per_rank, size_t D, size_t ld,
                      size_t col_begin, size_t col_end, float* out, int max_ctas, void* stream);
/* Few clients: segment-stacked form (rows cut into segments that fill the 128-row tile).  Buffers are [rows_per_rank, ld]
 * with ld = nseg_total * seg_len, zero padded; reduces segments [seg_begin, seg_begin + S) of every client. */
int fedb200_gram_bf16_seg(const void* const* rows_ptrs, int n_ranks, int rows_per_rank, size_t seg_len, int nseg_total,
                          int seg_begin, int S, float* out, int max_ctas, void* stream);
/* K4, single launch: the same kernels with everything a multi-GPU call needs inside the launch — cross-rank start barrier
 * (all delta rows written), clearing of the partial buffer, the column-slice Gram, and the one-shot N x N cross-rank sum in
 * rank order (bit-identical result on every rank).  part_ptrs[p]: rank p's partial buffer, fp32 [2][n * n] zero-initialised
 * once (halves alternate by epoch parity); flag_ptrs[p]: rank p's FEDB200_FLAG_WORDS flag words; result: this rank's [n, n]
 * fp32 output; every rank calls with the same epoch 1, 2, 3, ...  n_ranks == 1 is allowed (single GPU). */
int fedb200_gram_fused(const void* const* rows_ptrs, int n_ranks, int rows_per_rank, size_t D, size_t ld, size_t col_begin,
                       size_t col_end, const float* const* part_ptrs, unsigned int* const* flag_ptrs, float* result,
                       int rank, unsigned int epoch, int timeout_ms, int max_ctas, void* stream);
int fedb200_gram_fused_seg(const void* const* rows_ptrs, int n_ranks, int rows_per_rank, size_t seg_len, int nseg_total,
                           int seg_begin, int S, const float* const* part_ptrs, unsigned int* const* flag_ptrs,
                           float* result, int rank, unsigned int epoch, int timeout_ms, int max_ctas, void* stream);
/* K5. PageRank power iteration + mean±k·sigma mask (reference: All_graphs_IMDB_dataset.ipynb:168-176). */
int fedb200_pagerank(const float* mat, int n, int mode, double alpha, double tol, int max_iter, double k_sigma,
                     int low_only, double* ranks, int* mask, double* info, void* stream);
/* K6. SHA-256 leaves of `src` in chunk_bytes pieces while copying it to `dst` (dst may be NULL or peer memory). */
int fedb200_sha256_leaves(const void* src, void* dst, size_t nbytes, unsigned int chunk_bytes, void* leaves,
                          void* stream);
/* Merkle reduction of n 32-byte nodes (clobbers `nodes`; `scratch` >= ceil(n/2)*32 bytes) -> 32-byte root. */
int fedb200_merkle_root(void* nodes, void* scratch, unsigned int n, void* root_out, void* stream);
/* Merkle root of `nbytes` at `src` -> 32 bytes at `root_out`, with NO workspace argument and no allocation:
 * leaves and tree levels live in a static device array inside the library (6 MiB per GPU, held until exit),
 * so at most FEDB200_STATIC_MERKLE_MAX_LEAVES chunks are accepted and larger inputs return an error.  For
 * callers that cannot allocate (async gossip receive path, budget-capped ledger auditors).  First call per
 * device resolves the array address — issue it before any stream capture.  Calls share the array: overlapping
 * calls on one GPU race.  No counterpart in the reference (it has no ledger code and no native code). */
#define FEDB200_STATIC_MERKLE_MAX_LEAVES 131072
int fedb200_sha256_root_static(const void* src, size_t nbytes, unsigned int chunk_bytes, void* root_out, void* stream);
/* K7. Re-hash n 128-byte block headers and check index / prev-hash links; first_bad = 0xFFFFFFFF if intact. */
int fedb200_block_validate(const void* headers, const void* recorded_hashes, unsigned int n, void* hashes_out,
                           unsigned int* first_bad, void* stream);

/* ---- peer-memory runtime (NVLink / NVSwitch) ------------------------------------------------------------------- */
/* Symmetric allocation: cudaMalloc + CUDA IPC export; the 64-byte handle goes to every peer through the host. */
int fedb200_peer_alloc(size_t bytes, void** dev_ptr, void* ipc_handle_out64);
int fedb200_peer_open(const void* ipc_handle64, void** mapped_ptr);
int fedb200_peer_close(void* mapped_ptr);
int fedb200_peer_free(void* dev_ptr);
/* K1. Two-shot fused FedAvg: rank r reduces slice r of every rank's `acc` buffer (acc_r = sum of n_k * w_k over
 * its clients), scales by 1 / sum over ranks of `my_weight` (exchanged through the flag words; pass
 * inv_total_weight > 0 to override) and writes the result to every rank's `out` buffer (and optional bf16
 * shadow); start/end barriers are flag words in peer memory.  acc_ptrs / out_ptrs / flag_ptrs: HOST arrays of
 * `world` peer-mapped device pointers (own rank included); flag buffers need FEDB200_FLAG_WORDS zero-initialised words. */
int fedb200_fedavg_reduce(const float* const* acc_ptrs, float* const* out_ptrs, void* const* shadow_ptrs,
                          unsigned int* const* flag_ptrs, int rank, int world, size_t n, float my_weight,
                          float inv_total_weight, unsigned int epoch, int timeout_ms, void* stream);
/* Same, with two more knobs.  oneshot != 0: every rank reduces the WHOLE range from all peers and writes only its own
 * `out` (one pass; for payloads of a few KB such as the N x N Gram partials or scalar metrics, where the second hop of the
 * two-shot form dominates); its flag words are separate, so one-shot and two-shot calls may interleave (each family counts
 * its own epochs).  max_ctas > 0 caps the grid: no CTA waits for another CTA of the same launch, so a small grid (e.g. 32)
 * can run on a side stream beside compute kernels; 0 = one CTA per SM.  If the summed weight is 0 (every contributor masked
 * out) or the start barrier timed out, `out` is left untouched.  Reference: serverless_IID_IMDB.py:269-270. */
int fedb200_fedavg_reduce_ex(const float* const* acc_ptrs, float* const* out_ptrs, void* const* shadow_ptrs,
                             unsigned int* const* flag_ptrs, int rank, int world, size_t n, float my_weight,
                             float inv_total_weight, unsigned int epoch, int timeout_ms, int oneshot, int max_ctas,
                             void* stream);
/* K1, NVLS form: the buffers are bound to a multicast object (host-side CUDA VMM set-up; fedb200/parallel/peer_mem.py uses
 * torch's symmetric-memory allocator for it); mc_acc / mc_out are the MULTICAST addresses, acc_ptrs / out_ptrs the unicast peer
 * pointers of the same buffers (scalar tail only).  The sum is formed inside the NVSwitch (multimem.ld_reduce) and the result is
 * written to every replica with one multimem.st, so each GPU receives D / world reduced elements instead of (world - 1) D /
 * world operands.  Same barriers, weights, masked / timed-out behaviour and epoch sequence as the two-shot P2P form. */
int fedb200_fedavg_reduce_nvls(const float* const* acc_ptrs, float* const* out_ptrs, const float* mc_acc, float* mc_out,
                               unsigned int* const* flag_ptrs, int rank, int world, size_t n, float my_weight,
                               float inv_total_weight, unsigned int epoch, int timeout_ms, int max_ctas, void* stream);
/* Slot exchange (device-side all-gather of small per-round records): `table_ptrs` are the `world` peer-mapped copies of a
 * [world, slot_bytes] table; this rank's slot is pushed into every peer's copy, then a cross-rank barrier.  Replaces the host
 * all_gather_object of roots / counts / metrics (no counterpart in the reference, which has no communication layer). */
int fedb200_peer_exchange(void* const* table_ptrs, unsigned int* const* flag_ptrs, int rank, int world, int slot_bytes,
                          unsigned int epoch, int timeout_ms, void* stream);
/* K2. Fused gossip mix, one launch per rank and round: start barrier -> (first-hop relay copies -> mid barrier) ->
 * out[r] = sum_j weights[r][j] * srcs[j] for all local rows r at once, every distinct source read once -> end barrier.
 * srcs / relay_srcs may be peer-mapped; weights is a HOST [n_rows][n_src] row-major array; any_relay must be the same on
 * every rank (it enables the mid barrier).  Limits: 48 sources, 8 rows, 8 relay copies.  Reference: the shared-model hand-off
 * + numpy mean, src/Serverlesscase/serverless_IID_IMDB.py:257-270; routed paths Medical_Transcriptions_All_graphs.ipynb:984-996. */
int fedb200_gossip_mix(const float* const* srcs, const float* weights, int n_src, float* const* outs, int n_rows,
                       const float* const* relay_srcs, float* const* relay_dsts, int n_relay, int any_relay,
                       unsigned int* const* flag_ptrs, int rank, int world, size_t n, unsigned int epoch, int timeout_ms,
                       void* stream);
/* K3. One asynchronous, staleness-weighted mix from OBSERVED versions, applied chunk by chunk as the peers' publishes land:
 * w_i <- (1 - sum a_j) w_i + sum a_j w_j(v_j), a_j = stale_w[e_i - v_j] / n_live.  rings / chunk_flags / pubs: per source, the
 * (peer-mapped) outbox ring [R][D], its chunk flags [R][n_chunks] and its publish-started version word; the first CTA fixes
 * the versions (whatever the pub words hold: nobody waits for a peer to reach an epoch) and writes them to log[0 .. n_src),
 * log[n_src] counts chunks whose producer lapped the reader (must stay 0); mask (device int32 [num_peers] or NULL) zeroes
 * eliminated peers; decision: >= 2 + n_src zeroed local words, launch_id increasing by 1 per launch; stale_w: HOST float[32].
 * Formula only in the reference: Medical_Transcriptions_All_graphs.ipynb:979-980, 997-999. */
int fedb200_gossip_mix_async(const float* self_src, float* out, const float* const* rings,
                             const unsigned int* const* chunk_flags, const unsigned int* const* pubs, const int* src_peers,
                             int n_src, const int* mask, unsigned int* decision, unsigned int* log, unsigned int* err, int R,
                             size_t D, size_t chunk_elems, unsigned int my_epoch, unsigned int launch_id,
                             const float* stale_w, int max_ctas, int timeout_ms, void* stream);
/* Publish copy for K3: CTA c copies chunk c of src into dst (an outbox slot, local or peer-mapped) and releases flags[c] = value. */
int fedb200_copy_flagged(const float* src, float* dst, size_t n, size_t chunk_elems, unsigned int* flags, unsigned int value,
                         void* stream);
/* flags[0 .. n) = value with a system-scope release (marks a slot written by another kernel, e.g. the fused hash + copy). */
int fedb200_set_flags(unsigned int* flags, unsigned int value, int n, void* stream);
/* Device-side barrier across ranks (flag words in peer memory), usable between any two kernels. */
int fedb200_peer_barrier(unsigned int* const* flag_ptrs, int rank, int world, unsigned int epoch, int timeout_ms,
                         void* stream);

/* K3 signalling: monotone progress words in peer memory (`flag` may be a peer-mapped pointer for the wait).  The wait
 * spins on the device with a clock-bounded deadline; on timeout it sets *err (optional) and returns so the stream
 * never hangs.  Used by the schedule-driven asynchronous gossip executor (fedb200/runtime.py). */
int fedb200_flag_set(unsigned int* flag, unsigned int value, void* stream);
int fedb200_flag_wait(const unsigned int* flag, unsigned int value, int timeout_ms, unsigned int* err, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FEDB200_H */
