// gemm2_tcgen05: the CTA-pair (cta_group::2) variant of the K8 GEMM — the one the trainer uses for every
// shape with N >= 128.
//
// Why a pair: with one CTA per 128 x 256 tile every 64-wide k-block moves 48 KB into shared memory (TMA writes)
// and the MMAs read the same 48 KB back out — 96 KB per 512 tensor cycles = 187 B/cycle against a 128 B/cycle
// shared-memory port, which caps the tensor pipe near 65 % (measured 50-63 %, profiles/).  Two SMs that share one
// 256 x BN tile each load only HALF of the B operand (A: 128 rows each, B: BN/2 rows each), the hardware feeds
// both tensor cores from both shared memories, and the per-SM traffic drops to 64 KB per 512 cycles.
//
// Per CTA (320 threads): warp 0 TMA producer (its half of A and B, completion counted on the LEADER's mbarrier),
// warp 1 = MMA issuer in the leader CTA only (tcgen05.mma.cta_group::2, M = 256; commits are multicast to both
// CTAs' barriers), warps 2-9 epilogue of the CTA's own 128 accumulator rows (same fused epilogues as the 1-CTA
// kernel, gemm_epilogue.cuh).  Accumulators are double-buffered in TMEM (2 x BN columns per CTA); the peer's
// epilogue warps release a buffer by arriving on the leader's barrier through the cluster shared window.
#include "gemm_epilogue.cuh"

namespace fedb {

constexpr int G2_BM = 128;                 // rows per CTA (256 per pair)
constexpr int G2_BK = 64;
constexpr int G2_EPI_WARPS = 8;
constexpr int G2_THREADS = 64 + G2_EPI_WARPS * 32;
constexpr int G2_A_BYTES = G2_BM * G2_BK * 2;
constexpr int G2_EPI_STAGE_BYTES = 8192;   // per epilogue warp: two 128B-swizzled [32 x 64] bf16 slots for the TMA stores

template <int BN> struct G2Cfg {
    static constexpr int B_HALF_ROWS = BN / 2;
    static constexpr int B_BYTES = B_HALF_ROWS * G2_BK * 2;
    static constexpr int STAGE_BYTES = G2_A_BYTES + B_BYTES;
    static constexpr int STAGES = (BN == 256) ? 5 : 6;            // 64 KB of the 227 KB goes to the epilogue's store staging
    static constexpr uint32_t TMEM_COLS = (2 * BN <= 256) ? 256 : 512;
};

template <int A_MN, int B_MN, int BN, int EPI>
__global__ void __launch_bounds__(G2_THREADS, 1) gemm2_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                      const __grid_constant__ CUtensorMap tmB,
                                                                      const __grid_constant__ CUtensorMap tmC,
                                                                      const __grid_constant__ CUtensorMap tmC2,
                                                                      const __grid_constant__ CUtensorMap tmAux,
                                                                      const GemmParams p) {
    using Cfg = G2Cfg<BN>;
    constexpr int STAGES = Cfg::STAGES;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[STAGES];
    __shared__ __align__(8) uint64_t empty_bar[STAGES];
    __shared__ __align__(8) uint64_t acc_full[2];
    __shared__ __align__(8) uint64_t acc_empty[2];
    __shared__ __align__(8) uint64_t aux_bar[G2_EPI_WARPS][2];      // per epilogue warp and staging slot: aux tile landed
    __shared__ uint32_t tmem_base_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();               // 0 = leader
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
    const int m_tiles = (p.M + 2 * G2_BM - 1) / (2 * G2_BM), n_tiles = (p.N + BN - 1) / BN;
    const int k_blocks = (p.K + G2_BK - 1) / G2_BK;
    const int kb_per_split = (k_blocks + p.splits - 1) / p.splits;
    const int total = m_tiles * n_tiles * p.splits;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmC);
        tma_prefetch_desc(&tmC2);
        tma_prefetch_desc(&tmAux);
        for (int w = 0; w < G2_EPI_WARPS; ++w) { mbar_init(&aux_bar[w][0], 1); mbar_init(&aux_bar[w][1], 1); }
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], 2 * G2_EPI_WARPS); }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc_pair(&tmem_base_slot, Cfg::TMEM_COLS);
    tc_fence_before_sync();
    cluster_sync_all();                                     // barriers + TMEM of BOTH CTAs are ready
    pdl_sync();                                             // everything above overlaps the previous kernel's tail
    tc_fence_after_sync();
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // =============================== TMA producer (both CTAs) ===============================
        // The whole warp runs the loop (warp-uniform control flow keeps addresses / coordinates in uniform registers);
        // one elected lane issues.  With a `lane == 0` branch around the loop the compiler has to build every operand
        // in vector registers and move it over with ELECT + R2UR.BROADCAST, which made the issue loops the bottleneck.
        uint32_t s = 0, ph = 0;
        for (int w = pair; w < total; w += n_pairs) {
            const int split = w % p.splits, tile = w / p.splits;
            const int n_blk = tile % n_tiles, m_blk = tile / n_tiles;
            const int kb0 = split * kb_per_split, kb1 = min(k_blocks, kb0 + kb_per_split);
            const int m0 = m_blk * 2 * G2_BM + (int)rank * G2_BM;
            const int n0 = n_blk * BN + (int)rank * Cfg::B_HALF_ROWS;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(&empty_bar[s], ph ^ 1u);
                if (elect_one()) {
                    if (rank == 0) mbar_expect_tx(&full_bar[s], 2 * Cfg::STAGE_BYTES);     // both CTAs' bytes land here
                    uint8_t* sa = smem + s * Cfg::STAGE_BYTES;
                    uint8_t* sb = sa + G2_A_BYTES;
                    if (A_MN == 0) {
                        tma_load_2d_pair(sa, &tmA, &full_bar[s], kb * G2_BK, m0);
                    } else {
#pragma unroll
                        for (int j = 0; j < G2_BM / 64; ++j)
                            tma_load_2d_pair(sa + j * 8192, &tmA, &full_bar[s], m0 + j * 64, kb * G2_BK);
                    }
                    if (B_MN == 0) {
                        tma_load_2d_pair(sb, &tmB, &full_bar[s], kb * G2_BK, n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < Cfg::B_HALF_ROWS / 64; ++j)
                            tma_load_2d_pair(sb + j * 8192, &tmB, &full_bar[s], n0 + j * 64, kb * G2_BK);
                    }
                }
                __syncwarp();
                if (++s == (uint32_t)STAGES) { s = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // =============================== MMA issuer (leader CTA only) ===============================
        if (rank == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(2 * G2_BM, BN, A_MN, B_MN);
            uint32_t s = 0, ph = 0, local = 0;
            for (int w = pair; w < total; w += n_pairs, ++local) {
                const int split = w % p.splits;
                const int kb0 = split * kb_per_split, kb1 = min(k_blocks, kb0 + kb_per_split);
                const uint32_t buf = local & 1u;
                mbar_wait(&acc_empty[buf], ((local >> 1) & 1u) ^ 1u);
                tc_fence_after_sync();
                const uint32_t tmem_d = tmem_base + buf * BN;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[s], ph);
                    tc_fence_after_sync();
                    const uint32_t sa = smem_u32(smem + s * Cfg::STAGE_BYTES), sb = sa + G2_A_BYTES;
                    // descriptors differ between the four K steps only in the 14-bit start-address field (>> 4)
                    const uint64_t da0 = A_MN ? umma_smem_desc(sa, 8192, 1024) : umma_smem_desc(sa, 16, 1024);
                    const uint64_t db0 = B_MN ? umma_smem_desc(sb, 8192, 1024) : umma_smem_desc(sb, 16, 1024);
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < G2_BK / 16; ++k)
                            umma_bf16_pair(tmem_d, da0 + (uint64_t)(k * ((A_MN ? 2048 : 32) >> 4)),
                                           db0 + (uint64_t)(k * ((B_MN ? 2048 : 32) >> 4)), idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                        umma_commit_pair(&empty_bar[s]);          // frees the slot in both CTAs
                    }
                    __syncwarp();
                    if (++s == (uint32_t)STAGES) { s = 0; ph ^= 1u; }
                }
                if (elect_one()) umma_commit_pair(&acc_full[buf]);     // both CTAs' epilogues may start
                __syncwarp();
            }
        }
    } else {
        // =============================== epilogue warps (both CTAs, own 128 rows) ===============================
        const int quad = warp & 3;
        const int half = (warp - 2) >> 2;
        const uint32_t stage = smem_u32(smem + STAGES * Cfg::STAGE_BYTES) + (uint32_t)(warp - 2) * G2_EPI_STAGE_BYTES;
        uint32_t local = 0, slot_it = 0, aux_phase = 0;
        for (int w = pair; w < total; w += n_pairs, ++local) {
            const int tile = w / p.splits;
            const int n_blk = tile % n_tiles, m_blk = tile / n_tiles;
            const uint32_t buf = local & 1u;
            const int split = w % p.splits;
            const bool has_k = split * kb_per_split < k_blocks;
            const int row = m_blk * 2 * G2_BM + (int)rank * G2_BM + quad * 32 + lane;
            gemm_epilogue_tile_tma<BN, EPI>(p, &tmC, &tmC2, &tmAux, stage, slot_it, aux_bar[warp - 2], aux_phase,
                                            tmem_base + ((uint32_t)(quad * 32) << 16) + buf * BN, row, n_blk * BN, half, has_k,
                                            &acc_full[buf], (local >> 1) & 1u, split);
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(&acc_empty[buf], 0);        // the leader's MMA warp owns this barrier
        }
        if (lane == 0) tma_store_wait_read<0>();                // staging must stay valid until the last bulk store read it
    }
    tc_fence_before_sync();
    cluster_sync_all();                                     // nobody frees TMEM / exits while the peer still uses it
    if (warp == 1) tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS);
}

template <int A_MN, int B_MN, int BN, int EPI>
static int launch_gemm2(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& tc2,
                        const CUtensorMap& taux, const GemmParams& p, int max_ctas, cudaStream_t s) {
    using Cfg = G2Cfg<BN>;
    const size_t smem = (size_t)Cfg::STAGES * Cfg::STAGE_BYTES + (size_t)G2_EPI_WARPS * G2_EPI_STAGE_BYTES + 1024;
    auto kern = gemm2_tcgen05_kernel<A_MN, B_MN, BN, EPI>;
    static bool configured = false;
    if (!configured) {
        FEDB_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = true;
    }
    const int m_tiles = (p.M + 2 * G2_BM - 1) / (2 * G2_BM), n_tiles = (p.N + BN - 1) / BN;
    int pairs = m_tiles * n_tiles * p.splits;
    int cap = sm_count() / 2;
    if (max_ctas > 0 && cap > max_ctas / 2) cap = max_ctas / 2 > 0 ? max_ctas / 2 : 1;
    if (pairs > cap) pairs = cap;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * pairs);
    cfg.blockDim = dim3(G2_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    FEDB_CHECK(cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, tc2, taux, p));
    return 0;
}

// Called by fedb200_gemm_bf16 (gemm_tcgen05.cu) for shapes the pair kernel covers.  Tensor maps are encoded here because
// the B box is BN/2 rows (K-major) and the A box 128 rows — different from the 1-CTA kernel.
int gemm2_dispatch(int layout, const void* A, int lda, const void* B, int ldb, const GemmParams& p, int BN, int max_ctas,
                   cudaStream_t s) {
    const int a_mn = layout == 2, b_mn = layout != 0;
    CUtensorMap ta, tb;
    int e;
    if (!a_mn) e = make_tmap_bf16_2d(&ta, A, (uint64_t)p.M, (uint64_t)p.K, (uint64_t)lda, G2_BM, G2_BK);
    else       e = make_tmap_bf16_2d(&ta, A, (uint64_t)p.K, (uint64_t)p.M, (uint64_t)lda, G2_BK, 64);
    if (e) return e;
    if (!b_mn) e = make_tmap_bf16_2d(&tb, B, (uint64_t)p.N, (uint64_t)p.K, (uint64_t)ldb, (uint32_t)(BN / 2), G2_BK);
    else       e = make_tmap_bf16_2d(&tb, B, (uint64_t)p.K, (uint64_t)p.N, (uint64_t)ldb, G2_BK, 64);
    if (e) return e;
    // output maps for the epilogue's TMA stores ([32 rows x 64 cols] boxes, 128B swizzle); unused outputs alias A's map
    CUtensorMap tc = ta, tc2 = ta, taux = ta;
    if (p.epi != FEDB200_EPI_ATOMIC_F32 && p.epi != FEDB200_EPI_DISCARD) {
        if (p.aux != nullptr) {
            e = make_tmap_bf16_2d(&taux, p.aux, (uint64_t)p.M, (uint64_t)p.N, (uint64_t)p.ld_aux, 32, 64);
            if (e) return e;
        }
        if (p.out != nullptr) {
            e = make_tmap_bf16_2d(&tc, p.out, (uint64_t)p.M, (uint64_t)p.N, (uint64_t)p.ldo, 32, 64);
            if (e) return e;
        }
        if (p.out2 != nullptr) {
            e = make_tmap_bf16_2d(&tc2, p.out2, (uint64_t)p.M, (uint64_t)p.N, (uint64_t)p.ldo, 32, 64);
            if (e) return e;
        }
    }
    // Epilogue-specialised kernels for the (layout, epilogue) pairs of the training step; EPI = -1 is the generic kernel
    // that dispatches on p.epi at run time (its per-element branches cost ~20 % on these shapes).
#define FEDB_G2_CASE(AM, BM, BNV, EP)                                                          \
    if (a_mn == AM && b_mn == BM && BN == BNV && (EP < 0 || p.epi == EP))                      \
        return launch_gemm2<AM, BM, BNV, EP>(ta, tb, tc, tc2, taux, p, max_ctas, s);
#define FEDB_G2_LAYOUTS(BNV)                                                                   \
    FEDB_G2_CASE(0, 0, BNV, FEDB200_EPI_BIAS) FEDB_G2_CASE(0, 0, BNV, FEDB200_EPI_BIAS_DROP_RES) \
    FEDB_G2_CASE(0, 0, BNV, FEDB200_EPI_BIAS_GELU) FEDB_G2_CASE(0, 0, BNV, -1)                 \
    FEDB_G2_CASE(0, 1, BNV, FEDB200_EPI_ADD) FEDB_G2_CASE(0, 1, BNV, FEDB200_EPI_MUL_GELU_GRAD) \
    FEDB_G2_CASE(0, 1, BNV, -1)                                                                \
    FEDB_G2_CASE(1, 1, BNV, FEDB200_EPI_ATOMIC_F32) FEDB_G2_CASE(1, 1, BNV, -1)
    FEDB_G2_LAYOUTS(256) FEDB_G2_LAYOUTS(128)
#undef FEDB_G2_LAYOUTS
#undef FEDB_G2_CASE
    FEDB_REQUIRE(false, "gemm2: no pair kernel for layout %d BN %d", layout, BN);
    return 0;
}

}  // namespace fedb
