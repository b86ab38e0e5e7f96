// cosine_gram (K4): G = Delta * Delta^T over the flattened client updates, on the 5th-gen tensor cores.
//
// Delta is [N, D] bf16 with N <= 128 clients and D ~ 1e8 — an extremely skinny GEMM whose cost is reading
// N*D*2 bytes, so the kernel is a streaming reduction over K = D:
//   * the reduction dimension is sharded over GPUs (rank r owns columns [k_begin, k_end)) and, inside a GPU,
//     over persistent CTAs; the all-gathered [N, D] matrix is never materialised — each CTA pulls its K-slice
//     of every rank's rows straight from that rank's (peer-mapped) buffer with TMA into one 128B-swizzled
//     [rows x 64] shared tile;
//   * one tcgen05.mma (M = 128, N = n_pad, K = 16, cta_group::1) per 16 columns uses the SAME tile as the A and
//     the B operand (both K-major), accumulating the whole K-slice into a single [128 x n_pad] fp32 TMEM tile;
//   * epilogue: tcgen05.ld -> fp32 atomics into the N x N result (zeroed by the caller); the cross-rank
//     N x N sum is a few KB.
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer, all four warps drain TMEM at the end.
// There is no reference counterpart (the reference scores a latency graph offline, All_graphs_IMDB_dataset
// .ipynb:168-176); this is the north-star "anomaly scorer" of BASELINE.json.
#include "common.cuh"
#include "peer_flags.cuh"

namespace fedb {

constexpr int GR_STAGES = 8;
constexpr int GR_TILE_BYTES = 128 * 128;          // 128 rows x 64 bf16
constexpr int GR_THREADS = 128;

struct GramMaps {
    CUtensorMap map[FEDB200_MAX_RANKS];
};

// Fused (single-launch) form: cross-rank barriers, zeroing and the N x N cross-rank sum happen inside the kernel.
//   t.flags[p] : rank p's flag words;  t.acc[p] : rank p's partial buffers, fp32 [2][nn] (double-buffered by epoch parity);
//   t.out[rank]: this rank's result, fp32 [nn].  world == 0 selects the plain form (caller zeroes `out`, no barriers).
struct GramFused {
    PeerTable t;
    int rank, world;
    unsigned int epoch;
    long long timeout_cycles;
};

// Segment stacking (S > 1): with few clients a [rows x 64] tile is tiny (8 clients = 1 KB per TMA box) and the kernel
// becomes TMA-issue bound.  Each delta row is therefore cut into segments that are stacked as extra tile rows: the
// buffer is viewed as [rows * nseg_total, seg_len], one box brings S segments of one client ([S x 64], 2 KB at S = 16),
// the tile holds rows_total * S <= 128 rows, and the wanted Gram entry is the sum of the S "diagonal" products
// G[a, b] = sum_s acc[a*S + s, b*S + s]; the off-diagonal products are free riders on a bandwidth-bound kernel.
__global__ void __launch_bounds__(GR_THREADS, 1) gram_tcgen05_kernel(const __grid_constant__ GramMaps maps, int n_ranks,
                                                                     int boxes_per_rank, int box_rows, int n_pad,
                                                                     int kb_begin, int kb_end, float* __restrict__ out,
                                                                     int n_out, int S, int nseg_total, int seg_begin,
                                                                     const GramFused f) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[GR_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[GR_STAGES];
    __shared__ __align__(8) uint64_t done_bar;
    __shared__ uint32_t tmem_base_slot;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rows_total = n_ranks * boxes_per_rank * box_rows;      // tile rows (clients x segments)
    const bool fused = f.world > 0;
    const int nn = n_out * n_out;
    const long long deadline = clock64() + f.timeout_cycles;
    if (fused) {
        // every rank's delta rows are written, and every rank has left the previous call (so nobody still reads the
        // partial buffer of the other parity, which this call clears for the next one)
        start_barrier(f.t, FB_GRAM, f.rank, f.world, f.epoch, deadline, 0, 0.f);
        asm volatile("fence.proxy.async;" ::: "memory");     // the acquire above orders generic accesses; TMA reads are async-proxy
        out = const_cast<float*>(f.t.acc[f.rank]) + (size_t)(f.epoch & 1u) * nn;
        float* nxt = const_cast<float*>(f.t.acc[f.rank]) + (size_t)((f.epoch + 1u) & 1u) * nn;
        for (int i = blockIdx.x * GR_THREADS + threadIdx.x; i < nn; i += gridDim.x * GR_THREADS) nxt[i] = 0.f;
    }
    const uint32_t tmem_cols = n_pad <= 32 ? 32u : (n_pad <= 64 ? 64u : 128u);

    // rows >= rows_total are never written by TMA: clear the ring once so the unused A rows are finite
    for (int i = threadIdx.x; i < GR_STAGES * GR_TILE_BYTES / 16; i += GR_THREADS)
        reinterpret_cast<uint4*>(smem)[i] = make_uint4(0u, 0u, 0u, 0u);
    fence_async_smem();

    if (warp == 0 && lane == 0) {
        for (int r = 0; r < n_ranks; ++r) tma_prefetch_desc(&maps.map[r]);
        for (int s = 0; s < GR_STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&done_bar, 1);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(&tmem_base_slot, tmem_cols);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_d = tmem_base_slot;

    const int n_kb = (kb_end - kb_begin - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // this CTA's k-blocks
    const uint32_t stage_tx = (uint32_t)rows_total * 128u;

    if (warp == 0) {
        // ---------------- TMA producer (warp-uniform loop, one elected lane issues) ----------------
        uint32_t s = 0, ph = 0;
        for (int it = 0; it < n_kb; ++it) {
            mbar_wait(&empty_bar[s], ph ^ 1u);
            if (elect_one()) {
                mbar_expect_tx(&full_bar[s], stage_tx);
                const int kcol = (kb_begin + (int)blockIdx.x + it * (int)gridDim.x) * 64;
                for (int r = 0; r < n_ranks; ++r)
                    for (int pr = 0; pr < boxes_per_rank; ++pr)     // one box = all rows of a rank (S = 1) or S segments of one client
                        tma_load_2d(smem + s * GR_TILE_BYTES + (r * boxes_per_rank + pr) * box_rows * 128, &maps.map[r],
                                    &full_bar[s], kcol, pr * nseg_total + seg_begin);
            }
            __syncwarp();
            if (++s == (uint32_t)GR_STAGES) { s = 0; ph ^= 1u; }
        }
    } else if (warp == 1) {
        // ---------------- MMA issuer ----------------
        const uint32_t idesc = umma_idesc_bf16(128, n_pad, 0, 0);
        uint32_t s = 0, ph = 0;
        for (int it = 0; it < n_kb; ++it) {
            mbar_wait(&full_bar[s], ph);
            tc_fence_after_sync();
            const uint64_t d0 = umma_smem_desc(smem_u32(smem + s * GR_TILE_BYTES), 16, 1024);
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_bf16(tmem_d, d0 + (uint64_t)(k * 2), d0 + (uint64_t)(k * 2), idesc, (it | k) != 0);
                umma_commit(&empty_bar[s]);          // frees the slot once these MMAs have read it
            }
            __syncwarp();
            if (++s == (uint32_t)GR_STAGES) { s = 0; ph ^= 1u; }
        }
        if (elect_one()) umma_commit(&done_bar);
    }
    __syncwarp();

    // ---------------- epilogue: all four warps ----------------
    if (n_kb > 0) {
        mbar_wait(&done_bar, 0);
        tc_fence_after_sync();
        const int row = warp * 32 + lane;                     // tile row = client * S + segment
        const int a = row / S, sg = row % S;
        for (int c0 = 0; c0 < n_pad; c0 += 16) {
            uint32_t v[16];
            tmem_ld_32x16(tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
            tmem_ld_wait();
            if (a < n_out) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const int col = c0 + j;                   // keep only the same-segment ("diagonal") products
                    if (col % S == sg && col / S < n_out) atomicAdd(out + (size_t)a * n_out + col / S, __uint_as_float(v[j]));
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_d, tmem_cols);
    if (fused) {
        // the last CTA of this rank to finish publishes "my partial is complete", waits for every rank's, and sums the
        // partials in rank order (fixed order => every rank computes the bit-identical matrix, so their masks agree)
        unsigned int* my = f.t.flags[f.rank];
        __threadfence();
        __syncthreads();
        __shared__ unsigned int is_last;
        if (threadIdx.x == 0) is_last = atomicAdd(my + FLG_GR_DONE, 1u) == gridDim.x - 1 ? 1u : 0u;
        __syncthreads();
        if (is_last) {
            if (threadIdx.x == 0) my[FLG_GR_DONE] = 0u;
            cross_rank_barrier(f.t, FLG_GR_READY, f.rank, f.world, f.epoch, deadline);
            __syncthreads();
            const size_t off = (size_t)(f.epoch & 1u) * nn;
            for (int i = threadIdx.x; i < nn; i += GR_THREADS) {
                float sum = 0.f;
                for (int p = 0; p < f.world; ++p) sum += __ldcv(f.t.acc[p] + off + i);
                f.t.out[f.rank][i] = sum;
            }
        }
    }
}

}  // namespace fedb

using namespace fedb;

static thread_local GramFused g_fused_arg;
static thread_local bool g_use_fused = false;          // set by the fedb200_gram_fused* wrappers around the plain entry points
static GramFused make_plain() { GramFused f = {}; return f; }
#define g_fused (g_use_fused ? g_fused_arg : make_plain())

// rows_ptrs[r]: base of rank r's [rows_per_rank, ld] bf16 buffer (local or peer-mapped).  Accumulates G[i, j]
// (i, j < n_out = n_ranks * rows_per_rank) over columns [col_begin, col_end) into `out` (fp32 [n_out, n_out], caller
// zeroes it).  col_begin must be a multiple of 64.
extern "C" int fedb200_gram_bf16(const void* const* rows_ptrs, int n_ranks, int rows_per_rank, size_t D, size_t ld,
                                 size_t col_begin, size_t col_end, float* out, int max_ctas, void* stream) {
    FEDB_REQUIRE(n_ranks >= 1 && n_ranks <= FEDB200_MAX_RANKS, "gram: n_ranks=%d outside [1, %d]", n_ranks, FEDB200_MAX_RANKS);
    const int rows_total = n_ranks * rows_per_rank;
    FEDB_REQUIRE(rows_total >= 1 && rows_total <= 128, "gram: %d rows; at most 128 clients per launch", rows_total);
    FEDB_REQUIRE(col_begin % 64 == 0 && col_end <= D && col_begin <= col_end, "gram: bad column range");
    FEDB_REQUIRE(D % 8 == 0 && ld % 8 == 0, "gram: D and ld must be multiples of 8 (16-byte TMA strides)");
    if (col_begin == col_end) return 0;
    GramMaps maps;
    for (int r = 0; r < n_ranks; ++r) {
        int e = make_tmap_bf16_2d(&maps.map[r], rows_ptrs[r], (uint64_t)rows_per_rank, (uint64_t)D, (uint64_t)ld,
                                  (uint32_t)rows_per_rank, 64);
        if (e) return e;
    }
    const int n_pad = (rows_total + 15) / 16 * 16;
    const int kb_begin = (int)(col_begin / 64), kb_end = (int)((col_end + 63) / 64);
    int grid = sm_count();
    if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
    if (grid > kb_end - kb_begin) grid = kb_end - kb_begin;
    const size_t smem = (size_t)GR_STAGES * GR_TILE_BYTES + 1024;
    FEDB_CHECK(cudaFuncSetAttribute(gram_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // S = 1: every client row is one tile row; the single box per rank covers all of the rank's rows (row coordinate 0)
    gram_tcgen05_kernel<<<grid, GR_THREADS, smem, (cudaStream_t)stream>>>(maps, n_ranks, 1, rows_per_rank, n_pad, kb_begin, kb_end, out,
                                                                         rows_total, 1, 0, 0, g_fused);
    FEDB_LAUNCH_CHECK();
    return 0;
}

// Segment-stacked form for few clients.  Every rank's buffer is [rows_per_rank, ld] bf16 with ld = nseg_total * seg_len
// (zero padded past the real row length); this launch reduces segments [seg_begin, seg_begin + S) of every client, i.e.
// columns [seg_begin * seg_len, (seg_begin + S) * seg_len).  n_ranks * rows_per_rank * S <= 128, seg_len % 64 == 0.
extern "C" int fedb200_gram_bf16_seg(const void* const* rows_ptrs, int n_ranks, int rows_per_rank, size_t seg_len,
                                     int nseg_total, int seg_begin, int S, float* out, int max_ctas, void* stream) {
    FEDB_REQUIRE(n_ranks >= 1 && n_ranks <= FEDB200_MAX_RANKS, "gram_seg: n_ranks=%d outside [1, %d]", n_ranks, FEDB200_MAX_RANKS);
    const int clients = n_ranks * rows_per_rank;
    FEDB_REQUIRE(S >= 1 && clients >= 1 && clients * S <= 128, "gram_seg: %d clients x %d segments exceed the 128-row tile", clients, S);
    FEDB_REQUIRE(seg_len % 64 == 0 && seg_begin >= 0 && seg_begin + S <= nseg_total, "gram_seg: bad segment layout");
    GramMaps maps;
    for (int r = 0; r < n_ranks; ++r) {
        int e = make_tmap_bf16_2d(&maps.map[r], rows_ptrs[r], (uint64_t)rows_per_rank * nseg_total, (uint64_t)seg_len,
                                  (uint64_t)seg_len, (uint32_t)S, 64);
        if (e) return e;
    }
    const int n_pad = (clients * S + 15) / 16 * 16;
    const int kb_end = (int)(seg_len / 64);
    int grid = sm_count();
    if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
    if (grid > kb_end) grid = kb_end;
    const size_t smem = (size_t)GR_STAGES * GR_TILE_BYTES + 1024;
    FEDB_CHECK(cudaFuncSetAttribute(gram_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    gram_tcgen05_kernel<<<grid, GR_THREADS, smem, (cudaStream_t)stream>>>(maps, n_ranks, rows_per_rank, S, n_pad, 0, kb_end, out, clients,
                                                                         S, nseg_total, seg_begin, g_fused);
    FEDB_LAUNCH_CHECK();
    return 0;
}

// Single-launch forms: the same kernels with the cross-rank start barrier, the clearing of the partial buffer, and the
// one-shot N x N cross-rank sum inside the launch.  `part_ptrs[p]`: rank p's partial buffer, fp32 [2][n_out * n_out],
// zero-initialised once (the kernel alternates between the halves by epoch parity and clears the idle half itself);
// `flag_ptrs[p]`: rank p's flag words; `result`: this rank's [n_out, n_out] output.  Every rank must call with the same
// epoch (1, 2, 3, ...).  One rank (`world` == 1) works too: the barriers then only involve the rank's own words.
static int set_fused(const float* const* part_ptrs, unsigned int* const* flag_ptrs, float* result, int rank, int world,
                     unsigned int epoch, int timeout_ms) {
    FEDB_REQUIRE(world >= 1 && world <= FEDB200_MAX_RANKS && rank >= 0 && rank < world, "gram_fused: rank %d / world %d", rank, world);
    GramFused& f = g_fused_arg;
    for (int p = 0; p < FEDB200_MAX_RANKS; ++p) {
        f.t.acc[p] = p < world ? part_ptrs[p] : nullptr;
        f.t.flags[p] = p < world ? flag_ptrs[p] : nullptr;
        f.t.out[p] = nullptr;
        f.t.shadow[p] = nullptr;
    }
    f.t.out[rank] = result;
    f.rank = rank; f.world = world; f.epoch = epoch;
    f.timeout_cycles = (long long)(timeout_ms > 0 ? timeout_ms : 20000) * 2000000LL;
    return 0;
}

extern "C" int fedb200_gram_fused(const void* const* rows_ptrs, int n_ranks, int rows_per_rank, size_t D, size_t ld,
                                  size_t col_begin, size_t col_end, const float* const* part_ptrs,
                                  unsigned int* const* flag_ptrs, float* result, int rank, unsigned int epoch,
                                  int timeout_ms, int max_ctas, void* stream) {
    int e = set_fused(part_ptrs, flag_ptrs, result, rank, n_ranks, epoch, timeout_ms);
    if (e) return e;
    g_use_fused = true;
    e = fedb200_gram_bf16(rows_ptrs, n_ranks, rows_per_rank, D, ld, col_begin, col_end, nullptr, max_ctas, stream);
    g_use_fused = false;
    return e;
}

extern "C" int fedb200_gram_fused_seg(const void* const* rows_ptrs, int n_ranks, int rows_per_rank, size_t seg_len,
                                      int nseg_total, int seg_begin, int S, const float* const* part_ptrs,
                                      unsigned int* const* flag_ptrs, float* result, int rank, unsigned int epoch,
                                      int timeout_ms, int max_ctas, void* stream) {
    int e = set_fused(part_ptrs, flag_ptrs, result, rank, n_ranks, epoch, timeout_ms);
    if (e) return e;
    g_use_fused = true;
    e = fedb200_gram_bf16_seg(rows_ptrs, n_ranks, rows_per_rank, seg_len, nseg_total, seg_begin, S, nullptr, max_ctas, stream);
    g_use_fused = false;
    return e;
}
