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

// =================================================================================================================
// gossip_mix_async (K3): staleness-weighted mixing applied AS PEER BUFFERS LAND, from OBSERVED versions.
//
//   w_i <- (1 - sum_j a_j) w_i + sum_j a_j w_j(v_j),   a_j = (mix_rate / n_live) / (1 + tau_j)^decay,   tau_j = e_i - v_j
//
// (formula: Medical_Transcriptions_All_graphs.ipynb:979-980, 997-999 give only the max-vs-sum timing model; the staleness
// rule is this framework's, fedb200/p2p/async_engine.py).  Every peer publishes version e into slot e % R of its outbox
// ring: it first bumps its `pub` word to e (publish STARTED), then the publish kernel writes the slot chunk by chunk and
// releases one flag word per chunk (value e).  A consumer never waits for a peer to reach some scheduled epoch: the
// first CTA of the mix kernel reads every source's `pub` word, whatever it is, and fixes the versions for this mix (one
// decision per launch, logged so the host can replay the mix bit for bit); all CTAs then walk the chunks, waiting only
// for the chunk flags of the chosen versions — data that is still in flight over NVLink is consumed the moment it lands.
// Sources that have published nothing yet, or that the on-device anomaly mask eliminates, get weight 0.  After reading a
// chunk its flag is checked again: a changed flag means the producer lapped the consumer by a whole ring (R versions)
// while it was reading; that is counted in log[n_src] ("torn") and must stay 0.
// =================================================================================================================
namespace fedb {

constexpr int AM_MAX_SRC = 8;

struct AsyncMixArgs {
    const float* self_src;                       // this peer's freshly trained weights (local)
    float* out;                                  // this peer's state after the mix (local)
    const float* ring[AM_MAX_SRC];               // source j's outbox ring [R][D] (peer-mapped)
    const unsigned int* chunk_flags[AM_MAX_SRC]; // source j's chunk flags [R][n_chunks] (peer-mapped)
    const unsigned int* pub[AM_MAX_SRC];         // source j's "publish started" version word (peer-mapped)
    int src_peer[AM_MAX_SRC];                    // global peer ids (index into mask)
    const int* mask;                             // device anomaly mask [num_peers], 1 = eliminated (may be null)
    unsigned int* decision;                      // local scratch: [0] claim epoch, [1] ready epoch, [2 + j] chosen version
    unsigned int* log;                           // local: [j] chosen version (0 = not used), [n_src] torn-chunk count
    unsigned int* err;                           // local error word
    int n_src, R, n_chunks;
    unsigned int my_epoch, launch_id;
    float stale_w[32];                           // host-computed mix_rate / (1 + tau)^decay for tau = 0..31 (exact replay)
    size_t D, chunk_elems;
    long long timeout_cycles;
};

__global__ void __launch_bounds__(256) gossip_mix_async_kernel(const __grid_constant__ AsyncMixArgs a) {
    const long long deadline = clock64() + a.timeout_cycles;
    __shared__ unsigned int s_ver[AM_MAX_SRC];
    __shared__ float s_w[AM_MAX_SRC + 1];
    __shared__ unsigned int s_claim;
    // ---- one version decision per launch ------------------------------------------------------------------
    if (threadIdx.x == 0) s_claim = atomicMax(a.decision, a.launch_id) < a.launch_id ? 1u : 0u;
    __syncthreads();
    if (s_claim) {
        if ((int)threadIdx.x < a.n_src) {
            const int j = threadIdx.x;
            unsigned int v = ld_acquire_sys(a.pub[j]);
            if (a.mask != nullptr && a.mask[a.src_peer[j]] != 0) v = 0u;          // eliminated by the anomaly scorer
            a.decision[2 + j] = v;
            a.log[j] = v;
        }
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) st_release_gpu(a.decision + 1, a.launch_id);
    } else if (threadIdx.x == 0) {
        if (!spin_until(a.decision + 1, a.launch_id, deadline, false)) *a.err = 1u;
    }
    __syncthreads();
    if ((int)threadIdx.x < a.n_src) s_ver[threadIdx.x] = ld_acquire_gpu(a.decision + 2 + threadIdx.x);
    __syncthreads();
    if (threadIdx.x == 0) {
        int live = 0;
        for (int j = 0; j < a.n_src; ++j) live += s_ver[j] != 0u;
        float tot = 0.f;
        for (int j = 0; j < a.n_src; ++j) {
            float w = 0.f;
            if (s_ver[j] != 0u) {
                int tau = (int)a.my_epoch - (int)s_ver[j];
                tau = tau < 0 ? 0 : (tau > 31 ? 31 : tau);
                w = a.stale_w[tau] / (float)live;
            }
            s_w[j] = w;
            tot += w;
        }
        s_w[AM_MAX_SRC] = 1.0f - tot;
    }
    __syncthreads();
    const float self_w = s_w[AM_MAX_SRC];
    // ---- chunk by chunk, as the data lands -------------------------------------------------------------------
    for (int c = blockIdx.x; c < a.n_chunks; c += gridDim.x) {
        if ((int)threadIdx.x < a.n_src && s_ver[threadIdx.x] != 0u) {
            const int j = threadIdx.x;
            const unsigned int* f = a.chunk_flags[j] + (size_t)(s_ver[j] % (unsigned)a.R) * a.n_chunks + c;
            if (!spin_until(f, s_ver[j], deadline, true)) *a.err = 1u;
        }
        __syncthreads();
        const size_t lo = (size_t)c * a.chunk_elems, hi = lo + a.chunk_elems < a.D ? lo + a.chunk_elems : a.D;
        const size_t lo4 = lo / 4, hi4 = hi / 4;                   // chunk_elems % 4 == 0, so lo is vector aligned
        for (size_t i = lo4 + threadIdx.x; i < hi4; i += blockDim.x) {
            float4 v[AM_MAX_SRC];
#pragma unroll
            for (int j = 0; j < AM_MAX_SRC; ++j)
                if (j < a.n_src && s_ver[j] != 0u)
                    v[j] = __ldcs(reinterpret_cast<const float4*>(a.ring[j] + (size_t)(s_ver[j] % (unsigned)a.R) * a.D) + i);
            const float4 me = __ldcs(reinterpret_cast<const float4*>(a.self_src) + i);
            float4 acc = make_float4(self_w * me.x, self_w * me.y, self_w * me.z, self_w * me.w);
#pragma unroll
            for (int j = 0; j < AM_MAX_SRC; ++j)
                if (j < a.n_src && s_ver[j] != 0u) {
                    const float w = s_w[j];
                    acc.x += w * v[j].x; acc.y += w * v[j].y; acc.z += w * v[j].z; acc.w += w * v[j].w;
                }
            reinterpret_cast<float4*>(a.out)[i] = acc;
        }
        for (size_t e = hi4 * 4 + threadIdx.x; e < hi; e += blockDim.x) {          // scalar tail of the last chunk
            float acc = self_w * a.self_src[e];
            for (int j = 0; j < a.n_src; ++j)
                if (s_ver[j] != 0u) acc += s_w[j] * a.ring[j][(size_t)(s_ver[j] % (unsigned)a.R) * a.D + e];
            a.out[e] = acc;
        }
        __syncthreads();
        if ((int)threadIdx.x < a.n_src && s_ver[threadIdx.x] != 0u) {              // lapped by the producer while reading?
            const int j = threadIdx.x;
            const unsigned int now = ld_acquire_sys(a.chunk_flags[j] + (size_t)(s_ver[j] % (unsigned)a.R) * a.n_chunks + c);
            if (now != s_ver[j]) atomicAdd(a.log + a.n_src, 1u);
        }
    }
}

// Publish copy with per-chunk landing flags: CTA c copies elements [c * chunk_elems, (c + 1) * chunk_elems) of `src` into `dst`
// (an outbox slot, local or peer-mapped) and then releases flags[c] = value at system scope.
__global__ void __launch_bounds__(256) copy_flagged_kernel(const float* __restrict__ src, float* __restrict__ dst, size_t n,
                                                           size_t chunk_elems, unsigned int* __restrict__ flags,
                                                           unsigned int value, int n_chunks) {
    for (int c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        const size_t lo = (size_t)c * chunk_elems, hi = lo + chunk_elems < n ? lo + chunk_elems : n;
        const size_t lo4 = lo / 4, hi4 = hi / 4;
        for (size_t i = lo4 + threadIdx.x; i < hi4; i += blockDim.x)
            reinterpret_cast<float4*>(dst)[i] = __ldcs(reinterpret_cast<const float4*>(src) + i);
        for (size_t e = hi4 * 4 + threadIdx.x; e < hi; e += blockDim.x) dst[e] = src[e];
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) st_release_sys(flags + c, value);
        __syncthreads();
    }
}

// flags[0 .. n) = value (system-scope release after everything earlier on the stream): used when another kernel wrote the slot
__global__ void __launch_bounds__(256) set_flags_kernel(unsigned int* __restrict__ flags, unsigned int value, int n) {
    __threadfence_system();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) st_release_sys(flags + i, value);
}

}  // namespace fedb

using namespace fedb;

// One asynchronous mix (see gossip_mix_async_kernel).  rings / chunk_flags / pubs: per source, (peer-mapped) pointers to its
// outbox ring [R][D], its chunk flags [R][n_chunks] and its publish-started word; mask: device int32 [num_peers] or NULL;
// decision: >= 2 + n_src zero-initialised local words reused by every launch (launch_id must increase by 1 per launch on
// that buffer); log: n_src + 1 local words written by the launch (chosen versions, torn-chunk count — zero it first);
// stale_w: HOST array of 32 floats, mix_rate / (1 + tau)^decay.  chunk_elems % 4 == 0.
extern "C" int fedb200_gossip_mix_async(const float* self_src, float* out, const float* const* rings,
                                        const unsigned int* const* chunk_flags, const unsigned int* const* pubs,
                                        const int* src_peers, int n_src, const int* mask, unsigned int* decision,
                                        unsigned int* log, unsigned int* err, int R, size_t D, size_t chunk_elems,
                                        unsigned int my_epoch, unsigned int launch_id, const float* stale_w, int max_ctas,
                                        int timeout_ms, void* stream) {
    FEDB_REQUIRE(n_src >= 0 && n_src <= AM_MAX_SRC, "gossip_mix_async: %d sources (max %d: bounded fan-in)", n_src, AM_MAX_SRC);
    FEDB_REQUIRE(R >= 1 && chunk_elems > 0 && chunk_elems % 4 == 0 && D > 0, "gossip_mix_async: bad ring / chunk geometry");
    FEDB_REQUIRE((((uintptr_t)self_src | (uintptr_t)out) & 15) == 0, "gossip_mix_async: 16-byte alignment");
    AsyncMixArgs a = {};
    a.self_src = self_src; a.out = out; a.mask = mask; a.decision = decision; a.log = log; a.err = err;
    for (int j = 0; j < n_src; ++j) {
        a.ring[j] = rings[j]; a.chunk_flags[j] = chunk_flags[j]; a.pub[j] = pubs[j]; a.src_peer[j] = src_peers[j];
        FEDB_REQUIRE(((uintptr_t)rings[j] & 15) == 0 && (D * 4) % 16 == 0, "gossip_mix_async: ring %d / D not 16-byte aligned", j);
    }
    a.n_src = n_src; a.R = R; a.D = D; a.chunk_elems = chunk_elems;
    a.n_chunks = (int)((D + chunk_elems - 1) / chunk_elems);
    a.my_epoch = my_epoch; a.launch_id = launch_id;
    for (int t = 0; t < 32; ++t) a.stale_w[t] = stale_w[t];
    a.timeout_cycles = (long long)(timeout_ms > 0 ? timeout_ms : 20000) * 2000000LL;
    int grid = a.n_chunks;
    const int cap = max_ctas > 0 ? max_ctas : 2 * sm_count();
    if (grid > cap) grid = cap;
    gossip_mix_async_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    FEDB_LAUNCH_CHECK();
    return 0;
}

extern "C" int fedb200_copy_flagged(const float* src, float* dst, size_t n, size_t chunk_elems, unsigned int* flags,
                                    unsigned int value, void* stream) {
    FEDB_REQUIRE(chunk_elems > 0 && chunk_elems % 4 == 0 && n > 0, "copy_flagged: bad geometry");
    FEDB_REQUIRE((((uintptr_t)src | (uintptr_t)dst) & 15) == 0, "copy_flagged: 16-byte alignment");
    const int n_chunks = (int)((n + chunk_elems - 1) / chunk_elems);
    int grid = n_chunks < 2 * sm_count() ? n_chunks : 2 * sm_count();
    copy_flagged_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, dst, n, chunk_elems, flags, value, n_chunks);
    FEDB_LAUNCH_CHECK();
    return 0;
}

extern "C" int fedb200_set_flags(unsigned int* flags, unsigned int value, int n, void* stream) {
    FEDB_REQUIRE(n > 0, "set_flags: n must be positive");
    set_flags_kernel<<<(n + 255) / 256 < 64 ? (n + 255) / 256 : 64, 256, 0, (cudaStream_t)stream>>>(flags, value, n);
    FEDB_LAUNCH_CHECK();
    return 0;
}
