// gemm_tcgen05 (K8): the encoder / pooler-free GEMMs of the local trainer on the 5th-gen tensor cores.
//   D[M, N] = A[M, K] * B[N, K]^T   with bf16 operands, fp32 accumulation in TMEM, fused epilogues.
// Replaces the cuBLAS calls behind nn.Linear in the reference's train loop
// (src/Servercase/server_IID_IMDB.py:108-118; shapes in SURVEY §2.6).
//
// Structure (persistent, warp-specialised, one CTA per SM, 320 threads):
//   warp 0  TMA producer   : cp.async.bulk.tensor tiles of A and B into a 4-deep 128B-swizzled smem ring
//   warp 1  MMA issuer     : one elected thread issues tcgen05.mma (cta_group::1, M=128, N=BN, K=16),
//                            accumulators double-buffered in TMEM (2 x BN columns), tcgen05.commit frees slots
//   warps 2-9 epilogue     : tcgen05.ld 32 rows x 32 columns per warp (two warps per lane quadrant interleave
//                            the column chunks), fused bias / dropout / residual / GELU / GELU' / fp32 split-K
//                            reduction, residual chunks prefetched one ahead, direct 16-byte global stores
// Operand layouts (selected per call, all fed by TMA, no transposes are ever materialised):
//   forward  y = x W^T      A K-major  (x  [M, K]),        B K-major  (W [N, K], the HF Linear layout)
//   dgrad    dx = dy W      A K-major  (dy [M, N']),       B MN-major (W [N', K'] read as [K = N', N = K'])
//   wgrad    dW = dy^T x    A MN-major (dy [tokens, N']),  B MN-major (x [tokens, K']), split-K over tokens,
//                           fp32 vector atomics into the gradient arena
#include "gemm_epilogue.cuh"
#include <stdlib.h>

namespace fedb {

constexpr int GM_BM = 128;
constexpr int GM_BK = 64;
constexpr int GM_STAGES = 4;
constexpr int GM_EPI_WARPS = 8;            // two warps per TMEM lane quadrant; they interleave the 32-column chunks
constexpr int GM_THREADS = 64 + GM_EPI_WARPS * 32;
constexpr int GM_A_BYTES = GM_BM * GM_BK * 2;     // 16 KB

template <int A_MN, int B_MN, int BN>
__global__ void __launch_bounds__(GM_THREADS, 1) gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                     const __grid_constant__ CUtensorMap tmB,
                                                                     const GemmParams p) {
    constexpr int B_BYTES = BN * GM_BK * 2;
    constexpr int STAGE_BYTES = GM_A_BYTES + B_BYTES;
    constexpr uint32_t TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64 ? 64 : (2 * BN <= 128 ? 128 : (2 * BN <= 256 ? 256 : 512)));

    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[GM_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[GM_STAGES];
    __shared__ __align__(8) uint64_t acc_full[2];
    __shared__ __align__(8) uint64_t acc_empty[2];
    __shared__ uint32_t tmem_base_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_tiles = (p.M + GM_BM - 1) / GM_BM, n_tiles = (p.N + BN - 1) / BN;
    const int k_blocks = (p.K + GM_BK - 1) / GM_BK;
    const int kb_per_split = (k_blocks + p.splits - 1) / p.splits;
    const int total = m_tiles * n_tiles * p.splits;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int s = 0; s < GM_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], GM_EPI_WARPS); }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(&tmem_base_slot, TMEM_COLS);
    tc_fence_before_sync();
    __syncthreads();
    pdl_sync();                                             // everything above overlaps the previous kernel's tail
    tc_fence_after_sync();
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // =============================== TMA producer ===============================
        // warp-uniform loop, one elected lane issues (see gemm2_tcgen05.cu for why)
        uint32_t s = 0, ph = 0;
        for (int w = blockIdx.x; w < total; w += gridDim.x) {
            const int split = w % p.splits, tile = w / p.splits;
            const int n_blk = tile % n_tiles, m_blk = tile / n_tiles;
            const int kb0 = split * kb_per_split, kb1 = min(k_blocks, kb0 + kb_per_split);
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(&empty_bar[s], ph ^ 1u);
                if (elect_one()) {
                    mbar_expect_tx(&full_bar[s], STAGE_BYTES);
                    uint8_t* sa = smem + s * STAGE_BYTES;
                    uint8_t* sb = sa + GM_A_BYTES;
                    if (A_MN == 0) {
                        tma_load_2d(sa, &tmA, &full_bar[s], kb * GM_BK, m_blk * GM_BM);
                    } else {
#pragma unroll
                        for (int j = 0; j < GM_BM / 64; ++j)
                            tma_load_2d(sa + j * 8192, &tmA, &full_bar[s], m_blk * GM_BM + j * 64, kb * GM_BK);
                    }
                    if (B_MN == 0) {
                        tma_load_2d(sb, &tmB, &full_bar[s], kb * GM_BK, n_blk * BN);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_2d(sb + j * 8192, &tmB, &full_bar[s], n_blk * BN + j * 64, kb * GM_BK);
                    }
                }
                __syncwarp();
                if (++s == (uint32_t)GM_STAGES) { s = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // =============================== MMA issuer ===============================
        constexpr uint32_t idesc = umma_idesc_bf16(GM_BM, BN, A_MN, B_MN);
        uint32_t s = 0, ph = 0, local = 0;
        for (int w = blockIdx.x; w < total; w += gridDim.x, ++local) {
            const int split = w % p.splits;
            const int kb0 = split * kb_per_split, kb1 = min(k_blocks, kb0 + kb_per_split);
            const uint32_t buf = local & 1u;
            mbar_wait(&acc_empty[buf], ((local >> 1) & 1u) ^ 1u);
            tc_fence_after_sync();
            const uint32_t tmem_d = tmem_base + buf * BN;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(&full_bar[s], ph);
                tc_fence_after_sync();
                const uint32_t sa = smem_u32(smem + s * STAGE_BYTES), sb = sa + GM_A_BYTES;
                const uint64_t da0 = A_MN ? umma_smem_desc(sa, 8192, 1024) : umma_smem_desc(sa, 16, 1024);
                const uint64_t db0 = B_MN ? umma_smem_desc(sb, 8192, 1024) : umma_smem_desc(sb, 16, 1024);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < GM_BK / 16; ++k)
                        umma_bf16(tmem_d, da0 + (uint64_t)(k * ((A_MN ? 2048 : 32) >> 4)),
                                  db0 + (uint64_t)(k * ((B_MN ? 2048 : 32) >> 4)), idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                    umma_commit(&empty_bar[s]);
                }
                __syncwarp();
                if (++s == (uint32_t)GM_STAGES) { s = 0; ph ^= 1u; }
            }
            if (elect_one()) umma_commit(&acc_full[buf]);
            __syncwarp();
        }
    } else {
        // =============================== epilogue warps ===============================
        // Warps w and w+4 read the same TMEM lane quadrant (rows) and take alternate 32-column chunks; with a
        // single warp per SM sub-partition the elementwise math (GELU, dropout hash) was latency-bound.
        const int quad = warp & 3;                       // TMEM lane quadrant this warp may read
        const int half = (warp - 2) >> 2;                // 0: even chunks, 1: odd chunks
        uint32_t local = 0;
        for (int w = blockIdx.x; w < total; w += gridDim.x, ++local) {
            const int tile = w / p.splits;
            const int n_blk = tile % n_tiles, m_blk = tile / n_tiles;
            const uint32_t buf = local & 1u;
            const int split = w % p.splits;
            const bool has_k = split * kb_per_split < k_blocks;
            const int row = m_blk * GM_BM + quad * 32 + lane;
            gemm_epilogue_tile<BN>(p, tmem_base + ((uint32_t)(quad * 32) << 16) + buf * BN, row, n_blk * BN, half, has_k,
                                   &acc_full[buf], (local >> 1) & 1u, split);
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[buf]);
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

template <int A_MN, int B_MN, int BN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int max_ctas, cudaStream_t s) {
    constexpr int STAGE_BYTES = GM_A_BYTES + BN * GM_BK * 2;
    const size_t smem = (size_t)GM_STAGES * STAGE_BYTES + 1024;
    auto kern = gemm_tcgen05_kernel<A_MN, B_MN, BN>;
    static bool configured = false;
    if (!configured) {
        FEDB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = true;
    }
    const int m_tiles = (p.M + GM_BM - 1) / GM_BM, n_tiles = (p.N + BN - 1) / BN;
    int grid = m_tiles * n_tiles * p.splits;
    int cap = sm_count();
    if (max_ctas > 0 && cap > max_ctas) cap = max_ctas;
    if (grid > cap) grid = cap;
    FEDB_CHECK(launch_pdl(kern, dim3(grid), dim3(GM_THREADS), smem, s, ta, tb, p));
    return 0;
}

int gemm2_dispatch(int layout, const void* A, int lda, const void* B, int ldb, const GemmParams& p, int BN, int max_ctas,
                   cudaStream_t s);      // gemm2_tcgen05.cu

}  // namespace fedb

using namespace fedb;

// Diagnostic switch (production = 0), also settable through the FEDB200_GEMM_MODE environment variable (read once); used
// by the A/B benchmarks bench/diag_gemm.py and bench/diag_epi.py:
//   0 auto (CTA pairs whenever N >= 128)      1 force the 1-CTA kernel        3 pair kernel with 128-wide tiles
//   4 pair kernel with direct row-per-thread stores / loads instead of the TMA-staged epilogue
//   5 pair kernel, epilogue stages the tile but skips the global store (isolates store traffic; results are garbage)
static int g_gemm_mode = -1;
extern "C" int fedb200_set_gemm_mode(int mode) {
    g_gemm_mode = mode;
    return 0;
}
static int gemm_mode() {
    if (g_gemm_mode < 0) {
        const char* e = getenv("FEDB200_GEMM_MODE");
        g_gemm_mode = e ? atoi(e) : 0;
    }
    return g_gemm_mode;
}

// layout: 0 = NT (A [M,K] K-major, B [N,K] K-major)           forward
//         1 = NN (A [M,K] K-major, B [K,N] row-major)          dgrad
//         2 = TN (A [K,M] row-major, B [K,N] row-major)        wgrad (use epi = ATOMIC_F32, splits >= 1)
static int gemm_bf16_impl(int layout, const void* A, int lda, const void* B, int ldb, int M, int N, int K,
                          int epi, void* out, void* out2, int ldo, const float* bias, const void* aux,
                          int ld_aux, int act, unsigned long long seed, float dropout_p, int splits,
                          int max_ctas, void* stream, long long split_stride) {
    FEDB_REQUIRE(layout >= 0 && layout <= 2, "gemm: layout %d", layout);
    FEDB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem %dx%dx%d", M, N, K);
    FEDB_REQUIRE(N % 8 == 0 && ldo % 8 == 0 && lda % 8 == 0 && ldb % 8 == 0, "gemm: N, ldo, lda, ldb must be multiples of 8");
    FEDB_REQUIRE(epi >= 0 && epi <= FEDB200_EPI_DISCARD, "gemm: epilogue %d", epi);
    FEDB_REQUIRE(((uintptr_t)out & 15) == 0 && ((uintptr_t)out2 & 15) == 0 && ((uintptr_t)aux & 15) == 0 && ld_aux % 8 == 0,
                 "gemm: out/out2/aux must be 16-byte aligned");
    if (epi == FEDB200_EPI_BIAS_DROP_RES || epi == FEDB200_EPI_MUL_GELU_GRAD) FEDB_REQUIRE(aux != nullptr, "gemm: epilogue needs aux");
    if (epi == FEDB200_EPI_BIAS_GELU) FEDB_REQUIRE(out2 != nullptr, "gemm: GELU epilogue needs out2");
    if (epi != FEDB200_EPI_ATOMIC_F32) splits = 1;
    if (splits < 1) splits = 1;
    const int k_blocks = (K + GM_BK - 1) / GM_BK;
    if (splits > k_blocks) splits = k_blocks;
    // make every split non-empty
    while (splits > 1 && (splits - 1) * ((k_blocks + splits - 1) / splits) >= k_blocks) --splits;

    int BN = N >= 256 ? 256 : (N >= 128 ? 128 : 64);
    const int a_mn = layout == 2, b_mn = layout != 0;
    if (gemm_mode() == 3 && BN == 256) BN = 128;          // experiment: narrower pair tiles
    if (BN >= 128 && gemm_mode() != 1) {
        // CTA-pair kernel: 256 x BN tile per two SMs, each SM loads half of B
        GemmParams pp;
        pp.M = M; pp.N = N; pp.K = K; pp.splits = splits; pp.epi = epi; pp.out = out; pp.out2 = out2; pp.ldo = ldo;
        pp.split_stride = split_stride;
        pp.bias = bias; pp.aux = (const __nv_bfloat16*)aux; pp.ld_aux = ld_aux; pp.act = act; pp.seed = seed;
        pp.keep_thresh = dropout_thresh16(dropout_p);
        pp.inv_keep = 1.0f / (1.0f - dropout_p);
        pp.use_dropout = dropout_p > 0.f;
        pp.tma_store = gemm_mode() == 4 ? 0 : (gemm_mode() == 5 ? 2 : 1);
        return gemm2_dispatch(layout, A, lda, B, ldb, pp, BN, max_ctas, (cudaStream_t)stream);
    }
    CUtensorMap ta, tb;
    int e;
    if (!a_mn) e = make_tmap_bf16_2d(&ta, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, GM_BM, GM_BK);
    else       e = make_tmap_bf16_2d(&ta, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, GM_BK, 64);
    if (e) return e;
    if (!b_mn) e = make_tmap_bf16_2d(&tb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, (uint32_t)BN, GM_BK);
    else       e = make_tmap_bf16_2d(&tb, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, GM_BK, 64);
    if (e) return e;

    GemmParams p;
    p.M = M; p.N = N; p.K = K; p.splits = splits; p.epi = epi; p.out = out; p.out2 = out2; p.ldo = ldo;
    p.split_stride = split_stride;
    p.bias = bias; p.aux = (const __nv_bfloat16*)aux; p.ld_aux = ld_aux; p.act = act; p.seed = seed;
    p.keep_thresh = dropout_thresh16(dropout_p);
    p.inv_keep = 1.0f / (1.0f - dropout_p);
    p.use_dropout = dropout_p > 0.f;
    p.tma_store = 0;
    cudaStream_t s = (cudaStream_t)stream;

#define FEDB_GEMM_CASE(AM, BM, BNV) if (a_mn == AM && b_mn == BM && BN == BNV) return launch_gemm<AM, BM, BNV>(ta, tb, p, max_ctas, s);
    FEDB_GEMM_CASE(0, 0, 256) FEDB_GEMM_CASE(0, 0, 128) FEDB_GEMM_CASE(0, 0, 64)
    FEDB_GEMM_CASE(0, 1, 256) FEDB_GEMM_CASE(0, 1, 128) FEDB_GEMM_CASE(0, 1, 64)
    FEDB_GEMM_CASE(1, 1, 256) FEDB_GEMM_CASE(1, 1, 128) FEDB_GEMM_CASE(1, 1, 64)
#undef FEDB_GEMM_CASE
    FEDB_REQUIRE(false, "gemm: no kernel for layout %d BN %d", layout, BN);
    return 0;
}

extern "C" int fedb200_gemm_bf16(int layout, const void* A, int lda, const void* B, int ldb, int M, int N, int K,
                                 int epi, void* out, void* out2, int ldo, const float* bias, const void* aux,
                                 int ld_aux, int act, unsigned long long seed, float dropout_p, int splits,
                                 int max_ctas, void* stream) {
    if (epi == FEDB200_EPI_ATOMIC_F32 && splits > 1 && det_enabled() && M > 0 && N > 0 && K > 0) {
        // Reproducible mode (determinism.cu): split s accumulates into its own [M, ldo] partial in the workspace (one writer
        // per element, so the atomic add onto the zeroed partial is exact), and the partials are added to `out` in split order.
        const int k_blocks = (K + GM_BK - 1) / GM_BK;
        if (splits > k_blocks) splits = k_blocks;
        const size_t stride = (size_t)M * (size_t)ldo;
        const size_t fit = det_capacity() / (stride * sizeof(float));
        if ((size_t)splits > fit) splits = (int)fit;                       // a smaller split factor is slower, not different in kind
        if (splits > 1) {
            while (splits > 1 && (splits - 1) * ((k_blocks + splits - 1) / splits) >= k_blocks) --splits;   // as the kernel will see it
            float* ws = det_workspace((size_t)splits * stride * sizeof(float));
            FEDB_REQUIRE(ws != nullptr, "gemm: deterministic workspace unavailable");
            FEDB_CHECK(cudaMemsetAsync(ws, 0, (size_t)splits * stride * sizeof(float), (cudaStream_t)stream));
            const int e = gemm_bf16_impl(layout, A, lda, B, ldb, M, N, K, epi, ws, out2, ldo, bias, aux, ld_aux, act, seed, dropout_p,
                                         splits, max_ctas, stream, (long long)stride);
            if (e) return e;
            return det_ordered_split_sum(ws, splits, stride, stride, reinterpret_cast<float*>(out), (cudaStream_t)stream);
        }
    }
    return gemm_bf16_impl(layout, A, lda, B, ldb, M, N, K, epi, out, out2, ldo, bias, aux, ld_aux, act, seed, dropout_p, splits,
                          max_ctas, stream, 0);
}
