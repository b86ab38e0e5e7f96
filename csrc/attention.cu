// Fused masked-softmax attention, forward and backward (K9), on tcgen05 / TMEM with TMA-fed operands.
// Replaces HF BertSelfAttention (QK^T/sqrt(d) + padding mask -> softmax -> dropout -> PV; module dump at
// serverless_cancer_classification_with_BioBERT.ipynb:537-545).  Sequences are PACKED (no pad tokens):
// qkv is [T, 3H] bf16 (Q | K | V column blocks, head h at columns h*64), cu_seqlens gives each sequence's
// token range, so the padding mask of the reference becomes "kv index < sequence length".
//
// forward (`attn_fwd4_kernel`, the default), one CTA per (128-row q tile, head, sequence), two CTAs per SM; 576 threads:
//   warp 0 TMA   : Q tile once, then K_j / V_j tiles (single stage each; two CTAs per SM overlap instead)
//   warp 1 MMA   : S = Q K_j^T  (M128 N128 K64, both operands K-major from smem) -> TMEM;
//                  O += P_j V_j with the A operand P_j read FROM TMEM (bf16 pairs) and B = V tile MN-major from smem
//   warps 2-17   : four threads per q row, 32 kv columns each: one sweep over S per tile (row maximum, exp2 against the
//                  reference maximum known so far, row sum, dropout decisions), P packed to bf16 and stored back to TMEM.
//                  The running output never leaves TMEM; it is rescaled only when a row moves its reference by > 2^8.
//   TMEM per CTA : S 128 | O 64 | P 64 columns.  Shared memory carries only Q, K, V (profiles/attn_fwd_trace.md: the P round
//                  trip through shared memory was the bottleneck of the two-thread kernel, which is kept as
//                  `attn_fwd_kernel`, FEDB200_ATTN_FWD=2, for comparison).
// backward, one CTA per (128-row kv tile, head, sequence), looping over q tiles:
//   S^T = K Q_i^T and dP^T = V dO_i^T land in TMEM with the kv index on the lanes, so the thread that owns a kv
//   row produces its rows of P^T and dS^T directly in the A-operand layout: P^T goes back to TMEM as bf16 pairs and is
//   the TMEM A operand of dV += P^T dO_i; dS^T goes to shared memory, where the same buffer is read
//   K-major for dK += dS^T Q_i and MN-major for dQ_i = dS K (no transposes anywhere).  dV/dK accumulate in
//   TMEM across the q loop; every kv CTA stores its dQ partial into slot `kv tile` of `dq_accum` (scratch,
//   [ceil(max_seqlen/128), T, H] bf16, no initialisation needed) and the finishing kernel sums the slots in fp32.
// Dropout uses the library-wide counter RNG: row = head * T + q token, col = kv position in the sequence.  The forward
// can also store its keep decisions as bit words ([heads, T, ceil(max_seqlen/32)] uint32); given those, the backward
// reads one word per q row instead of re-hashing every element (13 -> 4 instructions per element).
// lse is stored in base-2 units of the scaled scores: lse2 = max2 + log2(sum), p = exp2(s*scale*log2e - lse2).
#include "common.cuh"
#include <type_traits>
#include "../include/fedb200.h"

namespace fedb {

constexpr int AT_SOFTMAX_THREADS = 256;      // two threads per q row
constexpr int AT_THREADS = 64 + AT_SOFTMAX_THREADS;   // warp 0 TMA, warp 1 MMA, warps 2-9 softmax
constexpr int AT_TILE = 128 * 64 * 2;        // 16 KB: [128 rows x 64 bf16]
constexpr float LOG2E = 1.4426950408889634f;

// write 8 consecutive bf16 (one 16-byte chunk `cc` of the 128-byte row `r`) into a 128B-swizzled tile
__device__ __forceinline__ void st_swizzled_chunk(uint8_t* tile, int r, int cc, const float* v) {
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
    o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(tile + r * 128 + ((cc ^ (r & 7)) << 4)) = o;
}

__device__ __forceinline__ uint64_t desc_kmajor(uint32_t tile_addr, int kstep) {     // 16 k-elements per step
    return umma_smem_desc(tile_addr + kstep * 32, 16, 1024);
}
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t tile_addr, int kstep, uint32_t lbo) {   // 16 k-rows per step
    return umma_smem_desc(tile_addr + kstep * 2048, lbo, 1024);
}

// =================================================================================================
// forward
// =================================================================================================
__global__ void __launch_bounds__(AT_THREADS, 2) attn_fwd_kernel(const __grid_constant__ CUtensorMap tm_qkv,
                                                                 __nv_bfloat16* __restrict__ ctx, float* __restrict__ lse,
                                                                 const int* __restrict__ cu_seqlens, int T, int H,
                                                                 float scale_log2, uint64_t seed, uint32_t keep_thresh,
                                                                 float inv_keep, int use_dropout,
                                                                 uint32_t* __restrict__ drop_bits, int bits_stride,
                                                                 long long* __restrict__ trace) {
    // trace (diagnostics, normally null): 32 clock64() stamps per CTA written by the first softmax thread
    long long* tr = trace ? trace + ((size_t)(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 32 : nullptr;
    const bool tracer = tr != nullptr && threadIdx.x == 64;
    if (tracer) { tr[0] = clock64(); unsigned sm; asm volatile("mov.u32 %0, %%smid;" : "=r"(sm)); tr[31] = sm; }
    pdl_sync();
    if (tracer) tr[1] = clock64();
    const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * 128;
    const int tok0 = cu_seqlens[b], len = cu_seqlens[b + 1] - tok0;
    if (q0 >= len) return;
    const int n_kv = (len + 127) / 128;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + AT_TILE;              // single stage: K_{j+1} streams in while the softmax of block j runs
    uint8_t* sV = sK + AT_TILE;
    uint8_t* sP = sV + AT_TILE;              // 2 tiles (kv 0-63 | 64-127)
    __shared__ __align__(8) uint64_t q_full, k_full, v_full, k_empty, v_empty, s_full, p_ready, o_full;
    __shared__ uint32_t tmem_slot;
    __shared__ float s_xchg[2][2][128];      // [tile parity][column half][row]: row maxima, then (slot 0) the row sums

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_qkv);
        mbar_init(&q_full, 1);
        mbar_init(&k_full, 1); mbar_init(&v_full, 1); mbar_init(&k_empty, 1); mbar_init(&v_empty, 1);
        mbar_init(&s_full, 1); mbar_init(&p_ready, AT_SOFTMAX_THREADS); mbar_init(&o_full, 1);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(&tmem_slot, 256);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_s = tmem_slot, tmem_o = tmem_slot + 128;

    if (warp == 0) {
        // warp-uniform loops, one elected lane issues (operands stay in uniform registers; see gemm2_tcgen05.cu)
        if (elect_one()) {
            mbar_expect_tx(&q_full, AT_TILE);
            tma_load_2d(sQ, &tm_qkv, &q_full, h * 64, tok0 + q0);
        }
        __syncwarp();
        for (int j = 0; j < n_kv; ++j) {
            const uint32_t ph = ((uint32_t)j & 1u) ^ 1u;
            mbar_wait(&k_empty, ph);
            if (elect_one()) {
                mbar_expect_tx(&k_full, AT_TILE);
                tma_load_2d(sK, &tm_qkv, &k_full, H + h * 64, tok0 + j * 128);
            }
            __syncwarp();
            mbar_wait(&v_empty, ph);
            if (elect_one()) {
                mbar_expect_tx(&v_full, AT_TILE);
                tma_load_2d(sV, &tm_qkv, &v_full, 2 * H + h * 64, tok0 + j * 128);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);
        constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);
        const uint32_t aQ = smem_u32(sQ), aP = smem_u32(sP);
        const uint32_t aK = smem_u32(sK), aV = smem_u32(sV);
        mbar_wait(&q_full, 0);
        mbar_wait(&k_full, 0);
        tc_fence_after_sync();
        if (elect_one()) {
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16(tmem_s, desc_kmajor(aQ, k), desc_kmajor(aK, k), idesc_s, k > 0);
            umma_commit(&s_full);
            umma_commit(&k_empty);
        }
        __syncwarp();
        for (int j = 0; j < n_kv; ++j) {
            mbar_wait(&p_ready, (uint32_t)j & 1u);            // S(j) has been consumed (and P(j) is in shared memory)
            // S(j+1) goes first: the softmax warps run their max pass over it while P(j) V(j) executes, so neither the
            // QK^T nor the PV latency sits on their critical path any more
            if (j + 1 < n_kv) {
                mbar_wait(&k_full, (uint32_t)(j + 1) & 1u);
                tc_fence_after_sync();
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(tmem_s, desc_kmajor(aQ, k), desc_kmajor(aK, k), idesc_s, k > 0);
                    umma_commit(&s_full);
                    umma_commit(&k_empty);
                }
                __syncwarp();
            }
            mbar_wait(&v_full, (uint32_t)j & 1u);
            tc_fence_after_sync();
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    umma_bf16(tmem_o, desc_kmajor(aP + (k >> 2) * AT_TILE, k & 3), desc_mnmajor(aV, k, 8192), idesc_o, k > 0);
                umma_commit(&o_full);
                umma_commit(&v_empty);
            }
            __syncwarp();
        }
    } else {
        // ---------------- softmax / output warps: TWO threads per q row ----------------
        // Warps 2-5 take kv columns 0-63 of every S tile, warps 6-9 columns 64-127 (a warp may only touch the TMEM lanes
        // of its quadrant, warp % 4, so the two threads of a row sit in warps w and w + 4).  The pair exchanges its row
        // maximum through shared memory once per tile (one named barrier); row sums stay per thread until the end; the
        // running output is split the same way (head-dim columns 0-31 | 32-63).  Round 1 had ONE thread per row: 8 softmax
        // warps per SM (two CTAs), i.e. two per scheduler, each a long dependent chain — the kernel was issue-latency
        // bound at 10.7 % tensor-pipe utilisation.
        const int quad = warp & 3;
        const int half = (warp - 2) >> 2;                     // 0: kv columns 0-63 / d columns 0-31, 1: the rest
        const int r = quad * 32 + lane;                       // row within the q tile == TMEM lane
        const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
        const int q_tok = tok0 + q0 + r;
        const bool q_ok = q0 + r < len;
        const uint32_t drop_row = (uint32_t)(h * T + q_tok);
        const uint32_t drop_key = dropout_key(seed);
        const float log2_inv_keep = use_dropout ? log2f(inv_keep) : 0.f;
        const bool keep_all = keep_thresh >= 65536u;                  // p == 0 through the dropout path
        const uint32_t thr_hi = keep_thresh << 16;                    // lane < thresh16  <=>  (lane << 16) < (thresh16 << 16)
        float o[32];
#pragma unroll
        for (int d = 0; d < 32; ++d) o[d] = 0.f;
        float m_run = -INFINITY, l_part = 0.f, alpha_prev = 1.f;
        uint8_t* tile = sP + half * AT_TILE;                  // this thread's 64 kv columns are exactly one P tile
        // running output <- running output * alpha + P(jj) V(jj) (this thread's 32 head-dim columns); also the point after which
        // the P buffer may be overwritten (the PV product has been read out of shared memory)
        auto fold_o = [&](int jj, float a) {
            mbar_wait(&o_full, (uint32_t)jj & 1u);
            tc_fence_after_sync();
            if (tracer) tr[6 + 5 * jj] = clock64();
            uint32_t raw[32];
            tmem_ld_32x32(tmem_o + lane_addr + half * 32, raw);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = o[i] * a + __uint_as_float(raw[i]);
            tc_fence_before_sync();
            if (tracer) tr[7 + 5 * jj] = clock64();
        };
        if (tracer) tr[2] = clock64();
        for (int j = 0; j < n_kv; ++j) {
            mbar_wait(&s_full, (uint32_t)j & 1u);
            tc_fence_after_sync();
            if (tracer) tr[3 + 5 * j] = clock64();
            const int kv_valid = min(128, len - j * 128) - half * 64;      // valid columns among this thread's 64
            const bool full = kv_valid >= 64;             // interior: no per-element masking (warp-uniform)
            // pass 1: maximum over this thread's 64 columns
            float mx = -INFINITY;
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                uint32_t raw[32];
                tmem_ld_32x32(tmem_s + lane_addr + half * 64 + c * 32, raw);
                tmem_ld_wait();
                if (full) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(raw[i]));
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (c * 32 + i < kv_valid) mx = fmaxf(mx, __uint_as_float(raw[i]));
                }
            }
            s_xchg[j & 1][half][r] = mx;
            asm volatile("bar.sync 1, %0;" ::"n"(AT_SOFTMAX_THREADS) : "memory");
            mx = fmaxf(mx, s_xchg[j & 1][half ^ 1][r]);
            if (tracer) tr[4 + 5 * j] = clock64();
            const float m_new = fmaxf(m_run, mx * scale_log2);
            const float alpha = exp2f(m_run - m_new);
            const float m_eff = m_new - log2_inv_keep;
            l_part *= alpha;
            if (j > 0) fold_o(j - 1, alpha_prev);            // P(j-1) V(j-1) finished while the max pass above ran
            alpha_prev = alpha;
            // pass 2: probabilities -> smem (bf16, swizzled K-major), partial row sum.  16 columns at a time: with the running
            // output (32 registers) live, a 32-wide chunk left the compiler no registers to interleave the independent
            // exp2 / hash chains (the kernel is capped at 96 registers by two 320-thread CTAs per SM)
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                const int chunk = j * 4 + half * 2 + c;                 // 32-column dropout chunk
                const uint32_t base = use_dropout ? attn_drop_base(drop_key, drop_row, (uint32_t)chunk) : 0u;
                uint32_t bits = 0u;
#pragma unroll
                for (int sub = 0; sub < 2; ++sub) {
                    uint32_t raw[16];
                    tmem_ld_32x16(tmem_s + lane_addr + half * 64 + c * 32 + sub * 16, raw);
                    tmem_ld_wait();
                    float p[16];
                    float rs = 0.f;
                    // with dropout the survivors' scale 1/(1-p) is folded into the exponent (m_eff = m - log2(1/(1-p))): every
                    // p[i], and so the row sum, carries that factor; it cancels in O / l and is subtracted from lse at the end
                    if (full) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            p[i] = exp2f(fmaf(__uint_as_float(raw[i]), scale_log2, -m_eff));
                            rs += p[i];
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            p[i] = (c * 32 + sub * 16 + i < kv_valid) ? exp2f(fmaf(__uint_as_float(raw[i]), scale_log2, -m_eff)) : 0.f;
                            rs += p[i];
                        }
                    }
                    l_part += rs;
                    if (use_dropout) {
                        // keep decisions: one full hash per (row, 32-column chunk), a cheap mixing step per column pair
                        // (common.cuh: attn_drop_*); they are also packed into one word that the backward re-reads
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const int kk = sub * 8 + k;
                            const uint32_t w = attn_drop_word(base, (uint32_t)kk);
                            const bool keep_lo = (w << 16) < thr_hi || keep_all;     // (w & 0xFFFF) < thresh16
                            const bool keep_hi = w < thr_hi || keep_all;             // (w >> 16)   < thresh16
                            if (keep_lo) bits |= 1u << (2 * kk); else p[2 * k] = 0.f;
                            if (keep_hi) bits |= 2u << (2 * kk); else p[2 * k + 1] = 0.f;
                        }
                    }
#pragma unroll
                    for (int g = 0; g < 2; ++g) st_swizzled_chunk(tile, r, c * 4 + sub * 2 + g, p + g * 8);
                }
                if (use_dropout && drop_bits != nullptr && q_ok && chunk < bits_stride)      // chunks past max_seqlen are fully masked
                    drop_bits[((size_t)h * T + q_tok) * bits_stride + chunk] = bits;
            }
            fence_async_smem();
            tc_fence_before_sync();
            mbar_arrive(&p_ready);
            m_run = m_new;
            if (tracer) tr[5 + 5 * j] = clock64();
        }
        fold_o(n_kv - 1, alpha_prev);
        // row sum = both halves' partial sums (both were rescaled by the same alphas)
        s_xchg[n_kv & 1][half][r] = l_part;
        asm volatile("bar.sync 1, %0;" ::"n"(AT_SOFTMAX_THREADS) : "memory");
        const float l_run = l_part + s_xchg[n_kv & 1][half ^ 1][r];
        if (q_ok) {
            const float inv = (use_dropout ? inv_keep : 1.f) / l_run;      // l_run carries the folded 1/(1-p): true sum = l_run * (1-p)
            __nv_bfloat16* dst = ctx + (size_t)q_tok * H + h * 64 + half * 32;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                uint4 pk;
                pk.x = pack_bf16x2(o[g * 8 + 0] * inv, o[g * 8 + 1] * inv); pk.y = pack_bf16x2(o[g * 8 + 2] * inv, o[g * 8 + 3] * inv);
                pk.z = pack_bf16x2(o[g * 8 + 4] * inv, o[g * 8 + 5] * inv); pk.w = pack_bf16x2(o[g * 8 + 6] * inv, o[g * 8 + 7] * inv);
                *reinterpret_cast<uint4*>(dst + g * 8) = pk;
            }
            if (half == 0) lse[(size_t)h * T + q_tok] = m_run + log2f(l_run) - log2_inv_keep;   // true lse (l carries 1/(1-p))
        }
        if (tracer) { tr[29] = n_kv; tr[30] = clock64(); }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_slot, 256);
}

// -------------------------------------------------------------------------------------------------
// forward, four threads per q row (the default; FEDB200_ATTN_FWD=2 selects the kernel above)
//
// What the two-thread kernel measured (profiles/attn_fwd_trace.md): per kv tile 144 KB of shared-memory traffic per CTA
// (MMA operand reads 80 KB, P stores 32 KB, TMA writes 32 KB) on a 128 B/clk port shared by two CTAs, two TMEM sweeps over S
// and a fold of the P V product into 32 output registers.  Here:
//   * 16 softmax warps per CTA (32 per SM): thread (row, quarter) owns 32 kv columns of its row and sweeps them ONCE per tile,
//     8 columns at a time with the next eight in flight: maximum, exponentials against the reference maximum known so far,
//     dropout decisions and bf16 packing in one pass.  If a row then turns out to exceed the reference by more than 2^8 the
//     warp repeats the sweep against the true maximum (S is released only after that decision), so the result is exact.
//   * P goes back to TMEM (one tcgen05.st per thread) and P_j V_j takes it from there as its A operand: no shared-memory
//     round trip, no proxy fence, and the P buffer (32 KB) is gone.
//   * the running output stays in TMEM: P_j V_j accumulates onto it, and it is rescaled (tcgen05.ld / st of the thread's
//     16 head-dim columns) only on that same rare path.
// fedb200/ops/attn_model.py restates the reference-maximum rule in NumPy (tests/test_attention_algorithm_cpu.py).
// -------------------------------------------------------------------------------------------------
constexpr int AT4_SOFTMAX_WARPS = 16;
constexpr int AT4_SOFTMAX_THREADS = AT4_SOFTMAX_WARPS * 32;
constexpr int AT4_THREADS = 64 + AT4_SOFTMAX_THREADS;      // warp 0 TMA, warp 1 MMA, warps 2-17 softmax
constexpr float AT4_RESCALE_LOG2 = 8.f;

template <bool TRACE>
__global__ void __launch_bounds__(AT4_THREADS, 2) attn_fwd4_kernel(const __grid_constant__ CUtensorMap tm_qkv,
                                                                  __nv_bfloat16* __restrict__ ctx, float* __restrict__ lse,
                                                                  const int* __restrict__ cu_seqlens, int T, int H,
                                                                  float scale_log2, uint64_t seed, uint32_t keep_thresh,
                                                                  float inv_keep, int use_dropout,
                                                                  uint32_t* __restrict__ drop_bits, int bits_stride,
                                                                  long long* __restrict__ trace, int diag) {
    // diag (TRACE build only; bench/attn_trace.py): 1 = no exp2 (p = its argument), 2 = no TMEM loads in the sweep, 4 = no P stores
    const bool d_noexp = TRACE && (diag & 1), d_nold = TRACE && (diag & 2), d_nost = TRACE && (diag & 4);
    long long* tr = (TRACE && trace) ? trace + ((size_t)(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * 32 : nullptr;
    const bool tracer = TRACE && tr != nullptr && threadIdx.x == 64;
    if (tracer) { tr[0] = clock64(); unsigned sm; asm volatile("mov.u32 %0, %%smid;" : "=r"(sm)); tr[31] = sm; }
    pdl_sync();
    if (tracer) tr[1] = clock64();
    const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * 128;
    const int tok0 = cu_seqlens[b], len = cu_seqlens[b + 1] - tok0;
    if (q0 >= len) return;
    const int n_kv = (len + 127) / 128;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + AT_TILE;
    uint8_t* sV = sK + AT_TILE;              // (P lives in TMEM: three 16 KB tiles per CTA)
    __shared__ __align__(8) uint64_t q_full, k_full, v_full, k_empty, v_empty, s_full, s_free, p_ready, o_full;
    __shared__ uint32_t tmem_slot;
    __shared__ float s_xchg[2][4][128];      // [tile parity][column quarter][row]: row maxima, at the end the row sums

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_qkv);
        mbar_init(&q_full, 1);
        mbar_init(&k_full, 1); mbar_init(&v_full, 1); mbar_init(&k_empty, 1); mbar_init(&v_empty, 1);
        mbar_init(&s_full, 1); mbar_init(&s_free, AT4_SOFTMAX_WARPS); mbar_init(&p_ready, AT4_SOFTMAX_WARPS);
        mbar_init(&o_full, 1);
        mbar_fence_init();
        if (diag & 8) {          // the first three tiles are requested before the TMEM allocation and the CTA-wide barrier
            mbar_expect_tx(&q_full, AT_TILE);
            tma_load_2d(sQ, &tm_qkv, &q_full, h * 64, tok0 + q0);
            mbar_expect_tx(&k_full, AT_TILE);
            tma_load_2d(sK, &tm_qkv, &k_full, H + h * 64, tok0);
            mbar_expect_tx(&v_full, AT_TILE);
            tma_load_2d(sV, &tm_qkv, &v_full, 2 * H + h * 64, tok0);
        }
    }
    if (warp == 1) tmem_alloc(&tmem_slot, 256);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_s = tmem_slot, tmem_o = tmem_slot + 128, tmem_p = tmem_slot + 192;   // S 128 | O 64 | P 64 (bf16 pairs)

    if (warp == 0) {
        const bool early = (diag & 8) != 0;
        if (!early && elect_one()) {
            mbar_expect_tx(&q_full, AT_TILE);
            tma_load_2d(sQ, &tm_qkv, &q_full, h * 64, tok0 + q0);
        }
        __syncwarp();
        for (int j = early ? 1 : 0; j < n_kv; ++j) {
            const uint32_t ph = ((uint32_t)j & 1u) ^ 1u;
            mbar_wait(&k_empty, ph);
            if (elect_one()) {
                mbar_expect_tx(&k_full, AT_TILE);
                tma_load_2d(sK, &tm_qkv, &k_full, H + h * 64, tok0 + j * 128);
            }
            __syncwarp();
            mbar_wait(&v_empty, ph);
            if (elect_one()) {
                mbar_expect_tx(&v_full, AT_TILE);
                tma_load_2d(sV, &tm_qkv, &v_full, 2 * H + h * 64, tok0 + j * 128);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);
        constexpr uint32_t idesc_o = umma_idesc_bf16(128, 64, 0, 1);
        const uint32_t aQ = smem_u32(sQ);
        const uint32_t aK = smem_u32(sK), aV = smem_u32(sV);
        mbar_wait(&q_full, 0);
        mbar_wait(&k_full, 0);
        tc_fence_after_sync();
        if (elect_one()) {
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16(tmem_s, desc_kmajor(aQ, k), desc_kmajor(aK, k), idesc_s, k > 0);
            umma_commit(&s_full);
            umma_commit(&k_empty);
        }
        __syncwarp();
        for (int j = 0; j < n_kv; ++j) {
            if (j + 1 < n_kv) {
                mbar_wait(&k_full, (uint32_t)(j + 1) & 1u);
                mbar_wait(&s_free, (uint32_t)j & 1u);             // every softmax thread holds its part of S(j) in registers
                tc_fence_after_sync();
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(tmem_s, desc_kmajor(aQ, k), desc_kmajor(aK, k), idesc_s, k > 0);
                    umma_commit(&s_full);
                    umma_commit(&k_empty);
                }
                __syncwarp();
            }
            mbar_wait(&v_full, (uint32_t)j & 1u);
            mbar_wait(&p_ready, (uint32_t)j & 1u);                // P(j) is in shared memory, O carries the right scale
            tc_fence_after_sync();
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < 8; ++k)        // A = P(j) from TMEM: 16 k-elements = 8 packed columns per step
                    umma_bf16_ts(tmem_o, tmem_p + k * 8, desc_mnmajor(aV, k, 8192), idesc_o, (j > 0 || k > 0) ? 1u : 0u);
                umma_commit(&o_full);
                umma_commit(&v_empty);
            }
            __syncwarp();
        }
    } else {
        const int quad = warp & 3;                            // TMEM lane quadrant this warp may touch
        const int quarter = (warp - 2) >> 2;                  // kv columns quarter*32 .. +31 of every tile; head-dim columns quarter*16 .. +15
        const int r = quad * 32 + lane;                       // row within the q tile == TMEM lane
        const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
        const int q_tok = tok0 + q0 + r;
        const bool q_ok = q0 + r < len;
        const uint32_t drop_row = (uint32_t)(h * T + q_tok);
        const uint32_t drop_key = dropout_key(seed);
        const float log2_inv_keep = use_dropout ? log2f(inv_keep) : 0.f;
        const bool keep_all = keep_thresh >= 65536u;
        const uint32_t thr_hi = keep_thresh << 16;
        // The row maximum starts as a GUESS (0): the probabilities of a tile are computed against the maximum known so far while
        // the tile's own maximum is collected in the same sweep, so S is read from TMEM once per tile (TMEM reads run at
        // 64 B/clk per SM — a second sweep over the 64 KB tile costs ~1 000 clk).  If the tile then turns out to exceed the
        // reference by more than 2^8 (or, on the first tile, to lie more than 2^8 below the guess) the sweep is repeated
        // against the true maximum.  Any reference value gives the same result as long as nothing over- or underflows:
        // numerator, row sum and lse all carry it.
        float m_run = 0.f, l_part = 0.f;
        if (tracer) tr[2] = clock64();
        for (int j = 0; j < n_kv; ++j) {
            mbar_wait(&s_full, (uint32_t)j & 1u);
            tc_fence_after_sync();
            if (tracer) tr[3 + 5 * j] = clock64();
            const int kv_valid = min(128, len - j * 128) - quarter * 32;      // valid columns among this thread's 32 (may be <= 0)
            const bool full = kv_valid >= 32;                 // warp-uniform
            const int chunk = j * 4 + quarter;                // 32-column dropout chunk == this thread's columns
            const uint32_t base = use_dropout ? attn_drop_base(drop_key, drop_row, (uint32_t)chunk) : 0u;
            if (tracer) tr[6 + 5 * j] = clock64();
            float rs, mx;
            uint32_t bits;
            uint32_t pk[16];                                  // this thread's 32 probabilities as bf16 pairs: 16 TMEM columns of P
            // one sweep over this thread's 32 columns, 16 at a time: raw maximum, probabilities against `m_eff` (the survivors'
            // dropout scale 1/(1-p) is folded into it: every p[i], and so the row sum, carries that factor; it cancels in O / l
            // and is subtracted from lse at the end), keep decisions, P packed to bf16 pairs in registers
            auto sweep = [&](const float m_eff) {
                rs = 0.f; mx = -INFINITY; bits = 0u;
                // 8 columns at a time, the next eight in flight while these are processed (the loads' latency, ~250-400 clk under
                // two CTAs, was fully exposed with 16-column chunks; wider prefetching does not fit the 56 registers)
                uint32_t buf[2][8];
                const uint32_t s_addr = tmem_s + lane_addr + quarter * 32;
                if (!d_nold) {
                    tmem_ld_32x8(s_addr, buf[0]);
                    tmem_ld_wait();
                }
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    uint32_t (&raw)[8] = buf[c & 1];
                    if (d_nold) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) raw[i] = (uint32_t)(i + lane) << 20;
                    } else if (c < 3) {
                        tmem_ld_32x8(s_addr + (c + 1) * 8, buf[(c + 1) & 1]);
                    }
                    if (tracer && j == 1 && (c & 1) == 0) { tr[23 + c] = clock64() + (__uint_as_float(raw[0]) == 1.2345f ? 1 : 0); }
                    float p[8];
                    if (full) {
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            mx = fmaxf(mx, __uint_as_float(raw[i]));
                            const float arg = fmaf(__uint_as_float(raw[i]), scale_log2, -m_eff);
                            p[i] = d_noexp ? arg : exp2f(arg);
                            rs += p[i];
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const bool ok = c * 8 + i < kv_valid;
                            if (ok) mx = fmaxf(mx, __uint_as_float(raw[i]));
                            p[i] = ok ? exp2f(fmaf(__uint_as_float(raw[i]), scale_log2, -m_eff)) : 0.f;
                            rs += p[i];
                        }
                    }
                    if (use_dropout) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const int kk = c * 4 + k;
                            const uint32_t w = attn_drop_word(base, (uint32_t)kk);
                            const bool keep_lo = (w << 16) < thr_hi || keep_all;     // (w & 0xFFFF) < thresh16
                            const bool keep_hi = w < thr_hi || keep_all;             // (w >> 16)   < thresh16
                            if (keep_lo) bits |= 1u << (2 * kk); else p[2 * k] = 0.f;
                            if (keep_hi) bits |= 2u << (2 * kk); else p[2 * k + 1] = 0.f;
                        }
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) pk[c * 4 + q] = pack_bf16x2(p[2 * q], p[2 * q + 1]);
                    if (!d_nold && c < 3) tmem_ld_wait();
                    if (tracer && j == 1 && (c & 1) == 1) tr[23 + c] = clock64();
                }
            };
            sweep(m_run - log2_inv_keep);
            s_xchg[j & 1][quarter][r] = mx;
            if (tracer && j == 1) tr[27] = clock64();
            asm volatile("bar.sync 1, %0;" ::"n"(AT4_SOFTMAX_THREADS) : "memory");
            const float mrow = fmaxf(fmaxf(s_xchg[j & 1][0][r], s_xchg[j & 1][1][r]), fmaxf(s_xchg[j & 1][2][r], s_xchg[j & 1][3][r])) *
                               scale_log2;
            if (tracer) tr[4 + 5 * j] = clock64();
            // decided per warp (the TMEM accesses are warp-collective); the four warps that share these 32 rows see the same
            // maxima and take the same branch
            const bool redo = mrow > m_run + AT4_RESCALE_LOG2 || (j == 0 && mrow < m_run - AT4_RESCALE_LOG2);
            bool waited_o = false;
            if (__any_sync(0xffffffffu, redo)) {
                const float m_new = j == 0 ? mrow : fmaxf(m_run, mrow);
                if (j > 0) {                                  // bring the row sum and the running output (TMEM) to the new reference
                    mbar_wait(&o_full, (uint32_t)(j - 1) & 1u);
                    tc_fence_after_sync();
                    waited_o = true;
                    const float alpha = exp2f(m_run - m_new);
                    l_part *= alpha;
                    uint32_t o[16];
                    tmem_ld_32x16(tmem_o + lane_addr + quarter * 16, o);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                    tmem_st_32x16(tmem_o + lane_addr + quarter * 16, o);
                    tmem_st_wait();
                }
                m_run = m_new;
                sweep(m_run - log2_inv_keep);
            }
            l_part += rs;
            if (use_dropout && drop_bits != nullptr && q_ok && chunk < bits_stride)     // chunks past max_seqlen are fully masked
                drop_bits[((size_t)h * T + q_tok) * bits_stride + chunk] = bits;
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_free);              // S(j+1) may overwrite the buffer: its MMAs run under the rest of this tile
            // P(j) goes to TMEM, where P(j) V(j) reads it as its A operand: no shared-memory round trip (P was 64 KB of the
            // 144 KB of shared-memory traffic per tile).  The columns are free once P(j-1) V(j-1) has completed — issued a
            // whole sweep ago, so this wait is short
            if (j > 0 && !waited_o) {
                mbar_wait(&o_full, (uint32_t)(j - 1) & 1u);
                tc_fence_after_sync();
            }
            if (!d_nost) {
                tmem_st_32x16(tmem_p + lane_addr + quarter * 16, pk);
                tmem_st_wait();
            }
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(&p_ready);             // P(j) is in TMEM, O carries the right reference
            if (tracer) tr[5 + 5 * j] = clock64();
        }
        mbar_wait(&o_full, (uint32_t)(n_kv - 1) & 1u);
        tc_fence_after_sync();
        s_xchg[n_kv & 1][quarter][r] = l_part;
        asm volatile("bar.sync 1, %0;" ::"n"(AT4_SOFTMAX_THREADS) : "memory");
        const float l_run = (s_xchg[n_kv & 1][0][r] + s_xchg[n_kv & 1][1][r]) + (s_xchg[n_kv & 1][2][r] + s_xchg[n_kv & 1][3][r]);
        uint32_t o[16];
        tmem_ld_32x16(tmem_o + lane_addr + quarter * 16, o);
        tmem_ld_wait();
        if (q_ok) {
            const float inv = (use_dropout ? inv_keep : 1.f) / l_run;      // l_run carries the folded 1/(1-p): true sum = l_run * (1-p)
            __nv_bfloat16* dst = ctx + (size_t)q_tok * H + h * 64 + quarter * 16;
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                uint4 pk;
                pk.x = pack_bf16x2(__uint_as_float(o[g * 8 + 0]) * inv, __uint_as_float(o[g * 8 + 1]) * inv);
                pk.y = pack_bf16x2(__uint_as_float(o[g * 8 + 2]) * inv, __uint_as_float(o[g * 8 + 3]) * inv);
                pk.z = pack_bf16x2(__uint_as_float(o[g * 8 + 4]) * inv, __uint_as_float(o[g * 8 + 5]) * inv);
                pk.w = pack_bf16x2(__uint_as_float(o[g * 8 + 6]) * inv, __uint_as_float(o[g * 8 + 7]) * inv);
                *reinterpret_cast<uint4*>(dst + g * 8) = pk;
            }
            if (quarter == 0) lse[(size_t)h * T + q_tok] = m_run + log2f(l_run) - log2_inv_keep;   // true lse (l carries 1/(1-p))
        }
        if (tracer) { tr[29] = n_kv; tr[30] = clock64(); }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_slot, 256);
}

// =================================================================================================
// backward
// =================================================================================================
// delta[h*T + tok] = sum_d dctx[tok, h*64 + d] * ctx[tok, h*64 + d]
__global__ void __launch_bounds__(256) attn_delta_kernel(const __nv_bfloat16* __restrict__ ctx, const __nv_bfloat16* __restrict__ dctx,
                                                         float* __restrict__ delta, int T, int heads) {
    pdl_sync();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= T * heads) return;
    const int tok = idx / heads, h = idx % heads;
    const size_t off = (size_t)tok * heads * 64 + h * 64;
    float s = 0.f;
#pragma unroll
    for (int g = 0; g < 8; ++g) {
        const uint4 a = *reinterpret_cast<const uint4*>(ctx + off + g * 8);
        const uint4 b = *reinterpret_cast<const uint4*>(dctx + off + g * 8);
        const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), a2 = unpack_bf16x2(a.z), a3 = unpack_bf16x2(a.w);
        const float2 b0 = unpack_bf16x2(b.x), b1 = unpack_bf16x2(b.y), b2 = unpack_bf16x2(b.z), b3 = unpack_bf16x2(b.w);
        s += a0.x * b0.x + a0.y * b0.y + a1.x * b1.x + a1.y * b1.y + a2.x * b2.x + a2.y * b2.y + a3.x * b3.x + a3.y * b3.y;
    }
    delta[(size_t)h * T + tok] = s;
}

// Finishing pass over dqkv [T, 3H]: the Q third is the sum of the per-kv-tile dQ partials written by the main kernel
// (`dq_part` is [n_kv_tiles, T, H] bf16; plain stores there instead of atomics: fp32 vector REDs into one [T, H]
// buffer measured ~200 us per layer of back-pressure, the partial round trip costs ~35 us), and — when `dbias` is
// given — the column sums of all three thirds (the q/k/v bias gradients) are accumulated in the same sweep, so the
// backward needs no separate column-sum kernel for dqkv.  One block column per 64 columns, one block row per sequence;
// blockDim = (32, 8): a warp-row covers 64 contiguous columns, the 8 rows of threads stride over the sequence's tokens.
__global__ void __launch_bounds__(256) attn_bwd_finish_kernel(const __nv_bfloat16* __restrict__ dq_part, __nv_bfloat16* __restrict__ dqkv,
                                                              float* __restrict__ dbias, const int* __restrict__ cu_seqlens,
                                                              int T, int H, float* __restrict__ part) {
    pdl_sync();
    // blockDim = (8, 32): 8 threads x 8 columns = one 64-column slab, 32 token rows per sweep; one block per
    // (slab, sequence) so each bias-gradient column receives one atomic per sequence
    __shared__ float red[32][65];
    const int col = blockIdx.x * 64 + threadIdx.x * 8;            // column within [0, 3H)
    const int s0 = cu_seqlens[blockIdx.y], s1 = cu_seqlens[blockIdx.y + 1];
    const int n_kv = (s1 - s0 + 127) / 128;
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    auto add8 = [](float (&a)[8], const uint4& u) {
        const float2 p0 = unpack_bf16x2(u.x), p1 = unpack_bf16x2(u.y), p2 = unpack_bf16x2(u.z), p3 = unpack_bf16x2(u.w);
        a[0] += p0.x; a[1] += p0.y; a[2] += p1.x; a[3] += p1.y; a[4] += p2.x; a[5] += p2.y; a[6] += p3.x; a[7] += p3.y;
    };
    if (col < 3 * H) {
        if (col < H) {
            for (int r = s0 + threadIdx.y; r < s1; r += 32) {
                float v[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] = 0.f;
                for (int j = 0; j < n_kv; ++j)
                    add8(v, *reinterpret_cast<const uint4*>(dq_part + ((size_t)j * T + r) * H + col));
                uint4 pk;
                pk.x = pack_bf16x2(v[0], v[1]); pk.y = pack_bf16x2(v[2], v[3]);
                pk.z = pack_bf16x2(v[4], v[5]); pk.w = pack_bf16x2(v[6], v[7]);
                *reinterpret_cast<uint4*>(dqkv + (size_t)r * 3 * H + col) = pk;
                add8(acc, pk);                                    // sum what the GEMMs will consume
            }
        } else if (dbias != nullptr) {
            for (int r = s0 + threadIdx.y; r < s1; r += 32)
                add8(acc, *reinterpret_cast<const uint4*>(dqkv + (size_t)r * 3 * H + col));
        }
    }
    if (dbias == nullptr) return;
#pragma unroll
    for (int k = 0; k < 8; ++k) red[threadIdx.y][threadIdx.x * 8 + k] = acc[k];
    __syncthreads();
    const int t = threadIdx.y * 8 + threadIdx.x;
    if (t < 64 && blockIdx.x * 64 + t < 3 * H) {
        float sum = 0.f;
#pragma unroll 8
        for (int y = 0; y < 32; ++y) sum += red[y][t];
        if (part != nullptr) part[(size_t)blockIdx.y * 3 * H + blockIdx.x * 64 + t] = sum;   // reproducible mode: per-sequence partials
        else atomicAdd(dbias + blockIdx.x * 64 + t, sum);
    }
}

constexpr int ATB_CWARPS = 16;            // compute warps: four per TMEM lane quadrant, 32 q-columns each
constexpr int ATB_CTHREADS = ATB_CWARPS * 32;
constexpr int ATB_THREADS = 64 + ATB_CTHREADS;   // warp 0 TMA, warp 1 MMA, warps 2-17 compute

template <bool PT_TMEM>
__global__ void __launch_bounds__(ATB_THREADS, 1) attn_bwd_kernel(const __grid_constant__ CUtensorMap tm_qkv,
                                                                 const __grid_constant__ CUtensorMap tm_do,
                                                                 const float* __restrict__ lse, const float* __restrict__ delta,
                                                                 __nv_bfloat16* __restrict__ dqkv, float* __restrict__ dq_accum,
                                                                 const int* __restrict__ cu_seqlens, int T, int H,
                                                                 float scale, float scale_log2, uint64_t seed,
                                                                 uint32_t keep_thresh, float inv_keep, int use_dropout,
                                                                 const uint32_t* __restrict__ drop_bits, int bits_stride) {
    pdl_sync();
    const int b = blockIdx.z, h = blockIdx.y, kv0 = blockIdx.x * 128;
    const int tok0 = cu_seqlens[b], len = cu_seqlens[b + 1] - tok0;
    if (kv0 >= len) return;
    const int n_q = (len + 127) / 128;

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sK = smem;
    uint8_t* sV = sK + AT_TILE;
    uint8_t* sQ = sV + AT_TILE;              // 2 stages
    uint8_t* sdO = sQ + 2 * AT_TILE;         // 2 stages
    // PT_TMEM: Pd^T lives in TMEM (64 columns of bf16 pairs, written with tcgen05.st and read by the dV MMA as its A operand)
    // instead of two shared-memory buffers: 64 KB less shared-memory traffic per q tile and 64 KB less shared memory
    uint8_t* sdS = sdO + 2 * AT_TILE;        // dS^T : 2 buffers x 2 tiles (q 0-63 | 64-127), rows = kv
    uint8_t* sP = sdS + 4 * AT_TILE;         // Pd^T : 2 buffers x 2 tiles (only when it is not kept in TMEM)
    __shared__ __align__(8) uint64_t kv_full, qdo_full[2], qdo_empty[2], sdp_full, pds_ready, dq_full, dq_empty, dkv_full, sdp_free,
        dv_done;
    __shared__ uint32_t tmem_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm_qkv);
        tma_prefetch_desc(&tm_do);
        mbar_init(&kv_full, 1);
        for (int s = 0; s < 2; ++s) { mbar_init(&qdo_full[s], 1); mbar_init(&qdo_empty[s], 1); }
        mbar_init(&sdp_full, 1); mbar_init(&pds_ready, ATB_CTHREADS); mbar_init(&dq_full, 1); mbar_init(&dq_empty, ATB_CTHREADS);
        mbar_init(&dkv_full, 1); mbar_init(&sdp_free, ATB_CTHREADS); mbar_init(&dv_done, 1);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(&tmem_slot, 512);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tm_st = tmem_slot, tm_dpt = tmem_slot + 128, tm_dv = tmem_slot + 256, tm_dk = tmem_slot + 320,
                   tm_dq = tmem_slot + 384, tm_pt = tmem_slot + 448;

    if (warp == 0) {
        // warp-uniform loops, one elected lane issues (operands stay in uniform registers; see gemm2_tcgen05.cu)
        if (elect_one()) {
            mbar_expect_tx(&kv_full, 2 * AT_TILE);
            tma_load_2d(sK, &tm_qkv, &kv_full, H + h * 64, tok0 + kv0);
            tma_load_2d(sV, &tm_qkv, &kv_full, 2 * H + h * 64, tok0 + kv0);
        }
        __syncwarp();
        for (int i = 0; i < n_q; ++i) {
            const int s = i & 1;
            mbar_wait(&qdo_empty[s], ((uint32_t)(i >> 1) & 1u) ^ 1u);
            if (elect_one()) {
                mbar_expect_tx(&qdo_full[s], 2 * AT_TILE);
                tma_load_2d(sQ + s * AT_TILE, &tm_qkv, &qdo_full[s], h * 64, tok0 + i * 128);
                tma_load_2d(sdO + s * AT_TILE, &tm_do, &qdo_full[s], h * 64, tok0 + i * 128);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        constexpr uint32_t idesc_nn = umma_idesc_bf16(128, 128, 0, 0);     // S^T, dP^T
        constexpr uint32_t idesc_kmn = umma_idesc_bf16(128, 64, 0, 1);     // dV, dK
        constexpr uint32_t idesc_mnmn = umma_idesc_bf16(128, 64, 1, 1);    // dQ
        const uint32_t aK = smem_u32(sK), aV = smem_u32(sV), aP = smem_u32(sP), aS = smem_u32(sdS);
        // Software pipeline: S^T/dP^T of q tile i+1 are issued BEFORE the dV/dK/dQ products of tile i, so the compute
        // warps start on tile i+1 after ~512 tensor cycles while the 24 gradient MMAs of tile i run underneath.  That is
        // safe because Pd^T/dS^T are double-buffered in shared memory (buffer i & 1) and every later commit covers the
        // earlier MMAs (in-order completion): sdp_full(i+2) implies the MMAs that read buffer i & 1 have finished.
        mbar_wait(&kv_full, 0);
        mbar_wait(&qdo_full[0], 0);
        tc_fence_after_sync();
        if (elect_one()) {
            const uint32_t aQ = smem_u32(sQ), aO = smem_u32(sdO);
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16(tm_st, desc_kmajor(aK, k), desc_kmajor(aQ, k), idesc_nn, k > 0);
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16(tm_dpt, desc_kmajor(aV, k), desc_kmajor(aO, k), idesc_nn, k > 0);
            umma_commit(&sdp_full);
        }
        __syncwarp();
        for (int i = 0; i < n_q; ++i) {
            const int s = i & 1;
            const uint32_t aQ = smem_u32(sQ + s * AT_TILE), aO = smem_u32(sdO + s * AT_TILE);
            const uint32_t aPb = aP + s * 2 * AT_TILE, aSb = aS + s * 2 * AT_TILE;
            mbar_wait(&sdp_free, (uint32_t)i & 1u);           // tile i: S^T/dP^T are in the compute warps' registers
            if (i + 1 < n_q) {
                const int s1 = s ^ 1;
                mbar_wait(&qdo_full[s1], (uint32_t)((i + 1) >> 1) & 1u);
                tc_fence_after_sync();
                if (elect_one()) {
                    const uint32_t aQ1 = smem_u32(sQ + s1 * AT_TILE), aO1 = smem_u32(sdO + s1 * AT_TILE);
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(tm_st, desc_kmajor(aK, k), desc_kmajor(aQ1, k), idesc_nn, k > 0);
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_bf16(tm_dpt, desc_kmajor(aV, k), desc_kmajor(aO1, k), idesc_nn, k > 0);
                    umma_commit(&sdp_full);
                }
                __syncwarp();
            }
            mbar_wait(&pds_ready, (uint32_t)i & 1u);          // tile i: Pd^T/dS^T written
            tc_fence_after_sync();
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < 8; ++k) {    // dV += Pd^T dO_i   (reduction over q)
                    if (PT_TMEM) umma_bf16_ts(tm_dv, tm_pt + k * 8, desc_mnmajor(aO, k, 8192), idesc_kmn, (i | k) != 0);
                    else umma_bf16(tm_dv, desc_kmajor(aPb + (k >> 2) * AT_TILE, k & 3), desc_mnmajor(aO, k, 8192), idesc_kmn, (i | k) != 0);
                }
                if (PT_TMEM) umma_commit(&dv_done);      // Pd^T(i) has been consumed: the compute warps may store tile i+1's
#pragma unroll
                for (int k = 0; k < 8; ++k)      // dK += dS^T Q_i
                    umma_bf16(tm_dk, desc_kmajor(aSb + (k >> 2) * AT_TILE, k & 3), desc_mnmajor(aQ, k, 8192), idesc_kmn, (i | k) != 0);
            }
            __syncwarp();
            mbar_wait(&dq_empty, ((uint32_t)i & 1u) ^ 1u);    // dQ of tile i-1 drained (the compute warps do it after tile i's math)
            tc_fence_after_sync();
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < 8; ++k)      // dQ_i = dS K   (A = dS^T buffer read MN-major, reduction over kv)
                    umma_bf16(tm_dq, desc_mnmajor(aSb, k, AT_TILE), desc_mnmajor(aK, k, 8192), idesc_mnmn, k > 0);
                umma_commit(&dq_full);
                umma_commit(&qdo_empty[s]);
            }
            __syncwarp();
        }
        if (elect_one()) umma_commit(&dkv_full);
        __syncwarp();
    } else {
        // Sixteen compute warps: the four warps of a TMEM lane quadrant (rows) take one 32-column group of
        // S^T / dP^T each; column groups 0-1 / 2-3 are the two 64-column smem tiles of Pd^T / dS^T.
        const int quad = warp & 3;
        const int cg = (warp - 2) >> 2;                       // q-column group 0..3
        const int r = quad * 32 + lane;                       // kv row for S^T/dP^T/dK/dV, q row for dQ
        const int tid = (warp - 2) * 32 + lane;               // 0..511
        const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
        const int kv = kv0 + r;
        const bool kv_ok = kv < len;
        const uint32_t drop_key_b = dropout_key(seed);
        const uint32_t drop_lane = (uint32_t)kv & 1u;
        const uint32_t drop_row0 = (uint32_t)(h * T + tok0);
        const int kv_chunk = kv >> 5;                         // warp-uniform (kv0 and quad*32 are multiples of 32)
        // Per-q-row inputs of this warp's 32 q columns live in registers (lane l <-> q = qb + cg*32 + l) and are broadcast
        // with shuffles; they are fetched one tile ahead so their global-memory latency is off the critical path.
        auto fetch_rows = [&](int qb, float& lse_q, float& delta_q, uint32_t& bits_q) {
            const int q = qb + cg * 32 + lane;
            lse_q = 0.f; delta_q = 0.f; bits_q = 0xFFFFFFFFu;
            if (q < len) {
                lse_q = __ldg(lse + (size_t)h * T + tok0 + q);
                delta_q = __ldg(delta + (size_t)h * T + tok0 + q);
                if (use_dropout && drop_bits != nullptr && kv_chunk < bits_stride)
                    bits_q = __ldg(drop_bits + ((size_t)h * T + tok0 + q) * bits_stride + kv_chunk);
            }
        };
        // dQ tile of q tile `it`: lanes are q rows now; each column group drains 16 of the 64 head-dim columns into
        // slot blockIdx.x of the partial buffer (plain bf16 stores, summed by attn_bwd_finish_kernel)
        auto drain_dq = [&](int it) {
            mbar_wait(&dq_full, (uint32_t)it & 1u);
            tc_fence_after_sync();
            const int q = it * 128 + r;
            uint32_t raw[16];
            tmem_ld_32x16(tm_dq + lane_addr + cg * 16, raw);
            tmem_ld_wait();
            if (q < len) {
                __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(dq_accum) + ((size_t)blockIdx.x * T + tok0 + q) * H + h * 64 + cg * 16;
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    uint4 pk;
                    pk.x = pack_bf16x2(__uint_as_float(raw[g * 8 + 0]), __uint_as_float(raw[g * 8 + 1]));
                    pk.y = pack_bf16x2(__uint_as_float(raw[g * 8 + 2]), __uint_as_float(raw[g * 8 + 3]));
                    pk.z = pack_bf16x2(__uint_as_float(raw[g * 8 + 4]), __uint_as_float(raw[g * 8 + 5]));
                    pk.w = pack_bf16x2(__uint_as_float(raw[g * 8 + 6]), __uint_as_float(raw[g * 8 + 7]));
                    *reinterpret_cast<uint4*>(dst + g * 8) = pk;      // default policy: the finish kernel re-reads it from L2
                }
            }
            tc_fence_before_sync();
            mbar_arrive(&dq_empty);
        };
        float my_lse, my_delta;
        uint32_t my_bits;
        fetch_rows(0, my_lse, my_delta, my_bits);
        const int drop_mode = !use_dropout ? 0 : (drop_bits != nullptr ? 1 : 2);
        for (int i = 0; i < n_q; ++i) {
            const int qb = i * 128;
            uint8_t* tp = sP + ((i & 1) * 2 + (cg >> 1)) * AT_TILE;
            uint8_t* td = sdS + ((i & 1) * 2 + (cg >> 1)) * AT_TILE;
            float nx_lse = 0.f, nx_delta = 0.f;
            uint32_t nx_bits = 0xFFFFFFFFu;
            if (i + 1 < n_q) fetch_rows(qb + 128, nx_lse, nx_delta, nx_bits);
            mbar_wait(&sdp_full, (uint32_t)i & 1u);
            tc_fence_after_sync();
            const bool full_tile = kv0 + 128 <= len && qb + 128 <= len;       // no masking needed (block-uniform)
#pragma unroll 1
            for (int hc = 0; hc < 2; ++hc) {                                   // two 16-column halves of the group
                const int c0 = cg * 32 + hc * 16;
                uint32_t rs[16], rp[16];
                tmem_ld_32x16(tm_st + lane_addr + c0, rs);
                tmem_ld_32x16(tm_dpt + lane_addr + c0, rp);
                tmem_ld_wait();
                if (hc == 1) {                   // S^T/dP^T fully read: the MMA warp may overwrite them with tile i+1 now
                    tc_fence_before_sync();
                    mbar_arrive(&sdp_free);
                }
                float pd[16], ds[16];
                // The element loop is instantiated per (dropout source, masking) combination and selected by one
                // warp-uniform switch: with the conditions inside the loop every element carried its own branches
                // (and the unused hash path), which serialised the 16 independent chains.
                auto elems = [&](auto drop_c, auto full_c) {
                    constexpr int DROP = decltype(drop_c)::value;       // 0 none, 1 stored keep-bits, 2 recompute the hash
                    constexpr bool FULL = decltype(full_c)::value;
#pragma unroll
                    for (int jq = 0; jq < 16; ++jq) {
                        const int q = qb + c0 + jq;
                        const float lse_q = __shfl_sync(0xffffffffu, my_lse, hc * 16 + jq);
                        const float delta_q = __shfl_sync(0xffffffffu, my_delta, hc * 16 + jq);
                        float p = exp2f(__uint_as_float(rs[jq]) * scale_log2 - lse_q);
                        if (!FULL) p = (kv_ok && q < len) ? p : 0.f;
                        float mk = 1.f;
                        if (DROP == 1) {
                            const uint32_t w = __shfl_sync(0xffffffffu, my_bits, hc * 16 + jq);
                            mk = (w >> lane) & 1u ? inv_keep : 0.f;
                        } else if (DROP == 2) {
                            mk = dropout_lane(attn_drop_word(attn_drop_base(drop_key_b, drop_row0 + (uint32_t)q, (uint32_t)kv >> 5),
                                                             ((uint32_t)kv & 31u) >> 1), drop_lane, keep_thresh, inv_keep);
                        }
                        pd[jq] = p * mk;
                        ds[jq] = p * (__uint_as_float(rp[jq]) * mk - delta_q) * scale;
                    }
                };
                using I0 = std::integral_constant<int, 0>; using I1 = std::integral_constant<int, 1>;
                using I2 = std::integral_constant<int, 2>;
                if (full_tile) {
                    if (drop_mode == 1) elems(I1{}, std::true_type{});
                    else if (drop_mode == 0) elems(I0{}, std::true_type{});
                    else elems(I2{}, std::true_type{});
                } else {
                    if (drop_mode == 1) elems(I1{}, std::false_type{});
                    else if (drop_mode == 0) elems(I0{}, std::false_type{});
                    else elems(I2{}, std::false_type{});
                }
                if (PT_TMEM) {
                    uint32_t pk[8];
#pragma unroll
                    for (int q = 0; q < 8; ++q) pk[q] = pack_bf16x2(pd[2 * q], pd[2 * q + 1]);
                    if (hc == 0 && i > 0) {      // the dV product of tile i-1 has read its Pd^T (issued a whole tile ago)
                        mbar_wait(&dv_done, (uint32_t)(i - 1) & 1u);
                        tc_fence_after_sync();
                    }
                    tmem_st_32x8(tm_pt + lane_addr + cg * 16 + hc * 8, pk);
                }
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    if (!PT_TMEM) st_swizzled_chunk(tp, r, (cg & 1) * 4 + hc * 2 + g, pd + g * 8);
                    st_swizzled_chunk(td, r, (cg & 1) * 4 + hc * 2 + g, ds + g * 8);
                }
            }
            if (PT_TMEM) tmem_st_wait();
            fence_async_smem();
            tc_fence_before_sync();
            mbar_arrive(&pds_ready);
            if (i > 0) drain_dq(i - 1);          // overlaps the S^T/dP^T MMAs of the next tile
            my_lse = nx_lse; my_delta = nx_delta; my_bits = nx_bits;
        }
        drain_dq(n_q - 1);
        mbar_wait(&dkv_full, 0);
        tc_fence_after_sync();
        {   // column groups 0-1 write dK (32 columns each) into the K block of dqkv, groups 2-3 write dV
            const int which = cg >> 1, c = cg & 1;
            uint32_t raw[32];
            tmem_ld_32x32((which ? tm_dv : tm_dk) + lane_addr + c * 32, raw);
            tmem_ld_wait();
            if (kv_ok) {
                __nv_bfloat16* dst = dqkv + (size_t)(tok0 + kv) * 3 * H + (which ? 2 * H : H) + h * 64 + c * 32;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    uint4 pk;
                    pk.x = pack_bf16x2(__uint_as_float(raw[g * 8 + 0]), __uint_as_float(raw[g * 8 + 1]));
                    pk.y = pack_bf16x2(__uint_as_float(raw[g * 8 + 2]), __uint_as_float(raw[g * 8 + 3]));
                    pk.z = pack_bf16x2(__uint_as_float(raw[g * 8 + 4]), __uint_as_float(raw[g * 8 + 5]));
                    pk.w = pack_bf16x2(__uint_as_float(raw[g * 8 + 6]), __uint_as_float(raw[g * 8 + 7]));
                    *reinterpret_cast<uint4*>(dst + g * 8) = pk;
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_slot, 512);
}

}  // namespace fedb

using namespace fedb;

static long long* g_attn_trace = nullptr;
static int g_attn_diag = 8;          // bits 0-2: diagnostics of the traced forward (see the kernel); bit 3: early TMA issue (default on,
                                     // FEDB200_ATTN_EARLY_TMA=0 switches it off: 59.9 -> 59.4 us with dropout, 47.9 -> 47.0 us without)
static int g_attn_smem_pad = 0;      // extra dynamic shared memory: > 30 KB leaves room for only one CTA per SM

extern "C" int fedb200_attn_fwd(const void* qkv, void* ctx, float* lse, const int* cu_seqlens, int B, int T, int heads,
                                int max_seqlen, float scale, unsigned long long seed, float dropout_p, void* drop_bits,
                                void* stream) {
    FEDB_REQUIRE(B > 0 && T > 0 && heads > 0 && max_seqlen > 0, "attn_fwd: empty problem");
    const int H = heads * 64;
    CUtensorMap tm;
    int e = make_tmap_bf16_2d(&tm, qkv, (uint64_t)T, (uint64_t)3 * H, (uint64_t)3 * H, 128, 64);
    if (e) return e;
    const size_t smem = 5 * AT_TILE + 1024;
    static bool configured = false;
    static int variant = 4;
    if (!configured) {
        FEDB_CHECK(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        FEDB_CHECK(cudaFuncSetAttribute(attn_fwd4_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        FEDB_CHECK(cudaFuncSetAttribute(attn_fwd4_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem + 64 * 1024));
        const char* v = getenv("FEDB200_ATTN_FWD");
        if (v != nullptr && (v[0] == '2' || v[0] == '4')) variant = v[0] - '0';
        const char* et = getenv("FEDB200_ATTN_EARLY_TMA");
        if (et != nullptr) g_attn_diag = (g_attn_diag & ~8) | (et[0] == '1' ? 8 : 0);
        configured = true;
    }
    dim3 grid((max_seqlen + 127) / 128, heads, B);
    if (variant == 4) {
        auto kern = g_attn_trace != nullptr ? attn_fwd4_kernel<true> : attn_fwd4_kernel<false>;
        FEDB_CHECK(launch_pdl(kern, dim3(grid), dim3(AT4_THREADS), (size_t)(3 * AT_TILE + 1024) + (size_t)g_attn_smem_pad, (cudaStream_t)stream,
            tm, (__nv_bfloat16*)ctx, lse, cu_seqlens, T, H, scale * LOG2E, seed, dropout_thresh16(dropout_p),
            1.0f / (1.0f - dropout_p), dropout_p > 0.f, (uint32_t*)drop_bits, (max_seqlen + 31) / 32, g_attn_trace, g_attn_diag));
        return 0;
    }
    FEDB_CHECK(launch_pdl(attn_fwd_kernel, dim3(grid), dim3(AT_THREADS), smem, (cudaStream_t)stream, 
        tm, (__nv_bfloat16*)ctx, lse, cu_seqlens, T, H, scale * LOG2E, seed, dropout_thresh16(dropout_p),
        1.0f / (1.0f - dropout_p), dropout_p > 0.f, (uint32_t*)drop_bits, (max_seqlen + 31) / 32, g_attn_trace));
    return 0;
}

// Diagnostics: when set, every attention-forward CTA writes 32 clock64() stamps (see the kernel) at trace[cta * 32 ...];
// the buffer must hold ceil(max_seqlen / 128) * heads * B * 32 int64.  NULL switches tracing off.
extern "C" int fedb200_attn_set_trace(void* trace) { g_attn_trace = (long long*)trace; return 0; }
extern "C" int fedb200_attn_set_diag(int diag, int smem_pad_bytes) {
    FEDB_REQUIRE(smem_pad_bytes >= 0 && smem_pad_bytes <= 64 * 1024, "attn_set_diag: pad %d", smem_pad_bytes);
    g_attn_diag = diag;
    g_attn_smem_pad = smem_pad_bytes;
    return 0;
}

extern "C" int fedb200_attn_bwd(const void* qkv, const void* ctx, const void* dctx, const float* lse, float* delta,
                                void* dqkv, float* dq_accum, float* dbias_qkv, const int* cu_seqlens, int B, int T,
                                int heads, int max_seqlen, float scale, unsigned long long seed, float dropout_p,
                                const void* drop_bits, void* stream) {
    FEDB_REQUIRE(B > 0 && T > 0 && heads > 0 && max_seqlen > 0, "attn_bwd: empty problem");
    const int H = heads * 64;
    cudaStream_t s = (cudaStream_t)stream;
    CUtensorMap tm, tmdo;
    int e = make_tmap_bf16_2d(&tm, qkv, (uint64_t)T, (uint64_t)3 * H, (uint64_t)3 * H, 128, 64);
    if (e) return e;
    e = make_tmap_bf16_2d(&tmdo, dctx, (uint64_t)T, (uint64_t)H, (uint64_t)H, 128, 64);
    if (e) return e;
    FEDB_CHECK(launch_pdl(attn_delta_kernel, dim3((T * heads + 255) / 256), dim3(256), 0, s, (const __nv_bfloat16*)ctx, (const __nv_bfloat16*)dctx, delta, T, heads));
    static bool configured = false;
    static int pt_tmem = 1;                 // Pd^T kept in TMEM (A operand of the dV product); FEDB200_ATTN_BWD_PT=smem: through shared memory
    if (!configured) {
        FEDB_CHECK(cudaFuncSetAttribute(attn_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 14 * AT_TILE + 1024));
        FEDB_CHECK(cudaFuncSetAttribute(attn_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 10 * AT_TILE + 1024));
        const char* v = getenv("FEDB200_ATTN_BWD_PT");
        if (v != nullptr) pt_tmem = (v[0] == 't');
        configured = true;
    }
    const size_t smem = (size_t)(pt_tmem ? 10 : 14) * AT_TILE + 1024;
    auto bwd = pt_tmem ? attn_bwd_kernel<true> : attn_bwd_kernel<false>;
    dim3 grid((max_seqlen + 127) / 128, heads, B);
    FEDB_CHECK(launch_pdl(bwd, dim3(grid), dim3(ATB_THREADS), smem, s, tm, tmdo, lse, delta, (__nv_bfloat16*)dqkv, dq_accum, cu_seqlens, T, H,
                                                   scale, scale * LOG2E, seed, dropout_thresh16(dropout_p),
                                                   1.0f / (1.0f - dropout_p), dropout_p > 0.f, (const uint32_t*)drop_bits,
                                                   (max_seqlen + 31) / 32));
    float* part = nullptr;
    if (det_enabled() && dbias_qkv != nullptr && B > 1) {
        part = det_workspace((size_t)B * 3 * H * sizeof(float));
        FEDB_REQUIRE(part != nullptr, "attn_bwd: deterministic workspace too small");
    }
    FEDB_CHECK(launch_pdl(attn_bwd_finish_kernel, dim3((3 * H + 63) / 64, B), dim3(8, 32), 0, s, (const __nv_bfloat16*)dq_accum,
                          (__nv_bfloat16*)dqkv, dbias_qkv, cu_seqlens, T, H, part));
    if (part != nullptr) FEDB_TRY(det_ordered_colsum(part, B, (size_t)3 * H, 3 * H, 1, dbias_qkv, nullptr, nullptr, nullptr, s));
    return 0;
}
