// gossip_mix (K2): routed neighbour exchange fused with the mixing-matrix combine, ONE launch per rank and round.
//
//   out[r] = sum_j W[r][j] * src_j        for every local peer row r   (w_i <- sum_{j in N(i)} W_ij w_j)
//
// The reference only has the degenerate form — one shared model object handed from client to client and a numpy mean
// (src/Serverlesscase/serverless_IID_IMDB.py:257-270); the routed exchange exists there as a hand calculation of
// "shortest path" transfer times (Medical_Transcriptions_All_graphs.ipynb:984-996: 1 -> 4 goes via 3).  Here:
//   * start barrier (every peer's slot is published), mix, end barrier (nobody overwrites a slot that is still being read)
//     are inside the kernel — round 1 needed a barrier kernel, one mix_n launch per local row, and another barrier kernel;
//   * every distinct source is read ONCE per element and feeds all local rows that use it (rows of one GPU share
//     neighbours), instead of once per row;
//   * routes chosen by the path optimiser are EXECUTED: for a two-hop route u -> r -> v the rank that owns relay peer r
//     first copies u's slot into its relay buffer (phase 1, reading over the u -> r link), all ranks pass a mid barrier,
//     and v's row then reads the relay copy over the r -> v link (phase 2).  Direct routes skip phase 1.
#include "common.cuh"
#include "peer_flags.cuh"

namespace fedb {

constexpr int GM_MAX_SRC = 48;          // distinct sources per launch (own slots + remote / relayed slots)
constexpr int GM_MAX_ROWS = 8;          // local peer rows per launch (peers per GPU)
constexpr int GM_MAX_RELAY = 8;         // relay copies this rank performs per launch
constexpr int FLG_MIX_MID = FLG_MIX_DONE + 1;                    // [FLG_MIX_MID + p]: "my relay copies are complete"
constexpr int FLG_MIX_MID_DONE = FLG_MIX_MID + FEDB200_MAX_RANKS;
static_assert(FLG_MIX_MID_DONE < FEDB200_FLAG_WORDS, "flag buffer too small");

struct GossipArgs {
    const float* src[GM_MAX_SRC];
    float w[GM_MAX_ROWS][GM_MAX_SRC];
    float* out[GM_MAX_ROWS];
    const float* relay_src[GM_MAX_RELAY];
    float* relay_dst[GM_MAX_RELAY];
    unsigned int* flags[FEDB200_MAX_RANKS];
    int n_src, n_rows, n_relay, any_relay;      // any_relay: some rank relays this round => everybody runs the mid barrier
    int rank, world;
    unsigned int epoch;
    long long timeout_cycles;
};

template <int ROWS>
__global__ void __launch_bounds__(256) gossip_mix_kernel(const __grid_constant__ GossipArgs a, size_t n) {
    PeerTable t;
#pragma unroll
    for (int p = 0; p < FEDB200_MAX_RANKS; ++p) { t.flags[p] = a.flags[p]; t.acc[p] = nullptr; t.out[p] = nullptr; t.shadow[p] = nullptr; }
    const long long deadline = clock64() + a.timeout_cycles;
    unsigned int* my = a.flags[a.rank];
    start_barrier(t, FB_MIX, a.rank, a.world, a.epoch, deadline, 0, 0.f);
    const size_t n4 = n / 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;

    if (a.any_relay) {
        // ---- phase 1: execute the first hop of two-hop routes (this rank is the relay) ----
        for (int k = 0; k < a.n_relay; ++k) {
            const float4* s = reinterpret_cast<const float4*>(a.relay_src[k]);
            float4* d = reinterpret_cast<float4*>(a.relay_dst[k]);
            size_t i = tid;
            for (; i + 3 * stride < n4; i += 4 * stride) {
                const float4 v0 = __ldcs(s + i), v1 = __ldcs(s + i + stride), v2 = __ldcs(s + i + 2 * stride), v3 = __ldcs(s + i + 3 * stride);
                d[i] = v0; d[i + stride] = v1; d[i + 2 * stride] = v2; d[i + 3 * stride] = v3;
            }
            for (; i < n4; i += stride) d[i] = __ldcs(s + i);
            for (size_t e = n4 * 4 + tid; e < n; e += stride) a.relay_dst[k][e] = a.relay_src[k][e];
        }
        // mid barrier: the last CTA of this rank announces "relay copies complete"; every CTA waits for every rank's word
        __threadfence_system();
        __syncthreads();
        __shared__ unsigned int last_mid;
        if (threadIdx.x == 0) last_mid = atomicAdd(my + FLG_MIX_MID_DONE, 1u) == gridDim.x - 1 ? 1u : 0u;
        __syncthreads();
        if (last_mid) {
            if (threadIdx.x == 0) my[FLG_MIX_MID_DONE] = 0u;
            if ((int)threadIdx.x < a.world) {
                __threadfence_system();
                st_release_sys(a.flags[threadIdx.x] + FLG_MIX_MID + a.rank, a.epoch);
            }
        }
        if ((int)threadIdx.x < a.world && !spin_until(my + FLG_MIX_MID + threadIdx.x, a.epoch, deadline, true)) my[FLG_ERROR] = 1u;
        __syncthreads();
    }

    // ---- phase 2: every source once per element, all local rows at once ----
    for (size_t i = tid; i < n4; i += stride) {
        float4 acc[ROWS];
#pragma unroll
        for (int r = 0; r < ROWS; ++r) acc[r] = make_float4(0.f, 0.f, 0.f, 0.f);
        int j = 0;
        for (; j + 4 <= a.n_src; j += 4) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = __ldcs(reinterpret_cast<const float4*>(a.src[j + u]) + i);
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int r = 0; r < ROWS; ++r) {
                    const float w = a.w[r][j + u];
                    acc[r].x += w * v[u].x; acc[r].y += w * v[u].y; acc[r].z += w * v[u].z; acc[r].w += w * v[u].w;
                }
        }
        for (; j < a.n_src; ++j) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(a.src[j]) + i);
#pragma unroll
            for (int r = 0; r < ROWS; ++r) {
                const float w = a.w[r][j];
                acc[r].x += w * v.x; acc[r].y += w * v.y; acc[r].z += w * v.z; acc[r].w += w * v.w;
            }
        }
#pragma unroll
        for (int r = 0; r < ROWS; ++r)
            if (r < a.n_rows) reinterpret_cast<float4*>(a.out[r])[i] = acc[r];
    }
    for (size_t e = n4 * 4 + tid; e < n; e += stride)
        for (int r = 0; r < a.n_rows; ++r) {
            float s = 0.f;
            for (int j = 0; j < a.n_src; ++j) s += a.w[r][j] * a.src[j][e];
            a.out[r][e] = s;
        }
    end_barrier(t, FB_MIX, a.rank, a.world, a.epoch, deadline);
}

}  // namespace fedb

using namespace fedb;

// srcs[n_src]: device pointers (local or peer-mapped) of the distinct sources; weights: HOST [n_rows][n_src] row-major;
// outs[n_rows]: local output rows.  relay_srcs / relay_dsts [n_relay]: first-hop copies this rank performs (peer-mapped source
// -> local relay buffer); any_relay != 0 when ANY rank relays this round (all ranks must then agree, they run a mid barrier).
// flag_ptrs[world]: every rank's flag words; every rank calls with the same epoch 1, 2, 3, ...
extern "C" int fedb200_gossip_mix(const float* const* srcs, const float* weights, int n_src, float* const* outs, int n_rows,
                                  const float* const* relay_srcs, float* const* relay_dsts, int n_relay, int any_relay,
                                  unsigned int* const* flag_ptrs, int rank, int world, size_t n, unsigned int epoch,
                                  int timeout_ms, void* stream) {
    FEDB_REQUIRE(n_src >= 1 && n_src <= GM_MAX_SRC, "gossip_mix: %d sources (max %d per launch)", n_src, GM_MAX_SRC);
    FEDB_REQUIRE(n_rows >= 1 && n_rows <= GM_MAX_ROWS, "gossip_mix: %d rows (max %d)", n_rows, GM_MAX_ROWS);
    FEDB_REQUIRE(n_relay >= 0 && n_relay <= GM_MAX_RELAY, "gossip_mix: %d relay copies (max %d)", n_relay, GM_MAX_RELAY);
    FEDB_REQUIRE(world >= 1 && world <= FEDB200_MAX_RANKS && rank >= 0 && rank < world, "gossip_mix: rank %d / world %d", rank, world);
    FEDB_REQUIRE(n_relay == 0 || any_relay, "gossip_mix: relay copies given but any_relay == 0");
    GossipArgs a = {};
    for (int j = 0; j < n_src; ++j) {
        FEDB_REQUIRE(((uintptr_t)srcs[j] & 15) == 0, "gossip_mix: source %d not 16-byte aligned", j);
        a.src[j] = srcs[j];
    }
    for (int r = 0; r < n_rows; ++r) {
        FEDB_REQUIRE(((uintptr_t)outs[r] & 15) == 0, "gossip_mix: output row %d not 16-byte aligned", r);
        a.out[r] = outs[r];
        for (int j = 0; j < n_src; ++j) a.w[r][j] = weights[r * n_src + j];
    }
    for (int k = 0; k < n_relay; ++k) { a.relay_src[k] = relay_srcs[k]; a.relay_dst[k] = relay_dsts[k]; }
    for (int p = 0; p < world; ++p) a.flags[p] = flag_ptrs[p];
    a.n_src = n_src; a.n_rows = n_rows; a.n_relay = n_relay; a.any_relay = any_relay ? 1 : 0;
    a.rank = rank; a.world = world; a.epoch = epoch;
    a.timeout_cycles = (long long)(timeout_ms > 0 ? timeout_ms : 20000) * 2000000LL;
    // the mid barrier makes CTAs of one launch wait for each other's phase 1: keep the grid within one resident wave
    const size_t n4 = n / 4;
    int grid = (int)((n4 + 255) / 256);
    const int cap = sm_count() * (any_relay ? 1 : 4);
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    cudaStream_t s = (cudaStream_t)stream;
    if (n_rows == 1) gossip_mix_kernel<1><<<grid, 256, 0, s>>>(a, n);
    else if (n_rows == 2) gossip_mix_kernel<2><<<grid, 256, 0, s>>>(a, n);
    else if (n_rows <= 4) gossip_mix_kernel<4><<<grid, 256, 0, s>>>(a, n);
    else gossip_mix_kernel<8><<<grid, 256, 0, s>>>(a, n);
    FEDB_LAUNCH_CHECK();
    return 0;
}
