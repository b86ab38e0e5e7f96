// Flag-word protocol of the peer-memory collectives (shared by peer.cu, gram_tcgen05.cu, elementwise.cu).
//   * every rank owns a buffer of FEDB200_FLAG_WORDS uint32 words, mapped into every peer (CUDA IPC);
//   * a rank signals by writing an EPOCH into its word of a peer's buffer with st.release.sys and waits by polling its
//     OWN buffer with ld.acquire.sys; epochs only ever increase, so words are never reset and a late reader cannot miss
//     a signal; every spin has a clock64() deadline and raises FLG_ERROR instead of hanging the GPU.
#pragma once
#include "common.cuh"
#include "../include/fedb200.h"

namespace fedb {

// flag-word layout inside each rank's flag buffer (uint32 words)
constexpr int FLG_START = 0;                       // [FLG_START + p]: start-barrier epoch written by rank p
constexpr int FLG_END = FEDB200_MAX_RANKS;         // [FLG_END + p]  : end-barrier epoch written by rank p
constexpr int FLG_GO = 2 * FEDB200_MAX_RANKS;      // local: block 0 -> other blocks of this rank
constexpr int FLG_DONE = FLG_GO + 1;               // local: finished-block counter
constexpr int FLG_ERROR = FLG_GO + 2;              // local: set to 1 when a spin timed out
constexpr int FLG_BAR = FLG_GO + 8;                // [FLG_BAR + p]: standalone barrier epochs
constexpr int FLG_WEIGHT = FLG_BAR + FEDB200_MAX_RANKS;   // [FLG_WEIGHT + p]: rank p's aggregation weight (float bits)
constexpr int FLG_XCHG = FLG_WEIGHT + FEDB200_MAX_RANKS;  // [FLG_XCHG + p]  : slot-exchange barrier epochs
constexpr int FLG_OS_START = FLG_XCHG + FEDB200_MAX_RANKS;      // one-shot all-reduce: start / end epochs
constexpr int FLG_OS_END = FLG_OS_START + FEDB200_MAX_RANKS;
constexpr int FLG_CLAIM = FLG_OS_END + FEDB200_MAX_RANKS;       // local: epoch claimed by the first CTA to arrive (two-shot)
constexpr int FLG_OS_CLAIM = FLG_CLAIM + 1;                     // local: same for the one-shot family
constexpr int FLG_OS_DONE = FLG_CLAIM + 2;                      // local: finished-block counter of the one-shot family
constexpr int FLG_GR_START = 72;                                // fused Gram: "my delta rows are written" epochs
constexpr int FLG_GR_READY = FLG_GR_START + FEDB200_MAX_RANKS;  // fused Gram: "my N x N partial is complete" epochs
constexpr int FLG_GR_CLAIM = FLG_GR_READY + FEDB200_MAX_RANKS;  // local
constexpr int FLG_GR_DONE = FLG_GR_CLAIM + 1;                   // local: finished-block counter
constexpr int FLG_MIX_START = 96;                               // fused gossip mix: start / end epochs
constexpr int FLG_MIX_END = FLG_MIX_START + FEDB200_MAX_RANKS;
constexpr int FLG_MIX_CLAIM = FLG_MIX_END + FEDB200_MAX_RANKS;  // local
constexpr int FLG_MIX_DONE = FLG_MIX_CLAIM + 1;                 // local
static_assert(FLG_MIX_DONE < FEDB200_FLAG_WORDS, "flag buffer too small");
static_assert(FLG_OS_DONE < FEDB200_FLAG_WORDS, "flag buffer too small");

struct PeerTable {
    const float* acc[FEDB200_MAX_RANKS];
    float* out[FEDB200_MAX_RANKS];
    __nv_bfloat16* shadow[FEDB200_MAX_RANKS];
    unsigned int* flags[FEDB200_MAX_RANKS];
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned int* p, unsigned int v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// epochs are monotone and compared with wrap-around safe arithmetic
__device__ __forceinline__ bool reached(unsigned int seen, unsigned int want) { return (int)(seen - want) >= 0; }

__device__ __forceinline__ bool spin_until(const unsigned int* p, unsigned int want, long long deadline, bool sys) {
    while (true) {
        const unsigned int v = sys ? ld_acquire_sys(p) : ld_acquire_gpu(p);
        if (reached(v, want)) return true;
        if (clock64() > deadline) return false;
        __nanosleep(64);
    }
}

// signal every peer then wait for every peer; executed by `world` threads of one block
__device__ __forceinline__ void cross_rank_barrier(const PeerTable& t, int base, int rank, int world, unsigned int epoch,
                                                   long long deadline, int publish_weight = 0, float weight = 0.f) {
    const int p = threadIdx.x;
    if (p < world) {
        if (publish_weight) t.flags[p][FLG_WEIGHT + rank] = __float_as_uint(weight);   // ordered by the release below
        __threadfence_system();
        st_release_sys(t.flags[p] + base + rank, epoch);
        if (!spin_until(t.flags[rank] + base + p, epoch, deadline, true)) t.flags[rank][FLG_ERROR] = 1u;
    }
}


// Flag bases of one collective family.  Two-shot FedAvg, the one-shot small all-reduce and the slot exchange each own a
// set of words, so their epoch sequences are independent and calls of different families may be interleaved freely.
struct FlagBases { int start, end, claim, done; };
constexpr FlagBases FB_TWOSHOT = {FLG_START, FLG_END, FLG_CLAIM, FLG_DONE};
constexpr FlagBases FB_ONESHOT = {FLG_OS_START, FLG_OS_END, FLG_OS_CLAIM, FLG_OS_DONE};
constexpr FlagBases FB_GRAM = {FLG_GR_START, FLG_GR_READY, FLG_GR_CLAIM, FLG_GR_DONE};
constexpr FlagBases FB_MIX = {FLG_MIX_START, FLG_MIX_END, FLG_MIX_CLAIM, FLG_MIX_DONE};

// Start barrier without an intra-rank dependency: the first CTA of this rank to arrive (atomicMax claim) tells every
// peer "my input is complete" (+ publishes this rank's weight); EVERY CTA then polls this rank's own flag words, which
// the peers write.  No CTA waits for another CTA of the same launch, so the grid need not be co-resident and a small
// grid can run beside compute kernels on a side stream.
__device__ __forceinline__ void start_barrier(const PeerTable& t, const FlagBases fb, int rank, int world, unsigned int epoch,
                                              long long deadline, int publish_weight, float weight) {
    unsigned int* my = t.flags[rank];
    __shared__ unsigned int s_first;
    if (threadIdx.x == 0) s_first = atomicMax(my + fb.claim, epoch) < epoch ? 1u : 0u;
    __syncthreads();
    const int p = threadIdx.x;
    if (p < world) {
        if (s_first) {
            if (publish_weight) t.flags[p][FLG_WEIGHT + rank] = __float_as_uint(weight);   // ordered by the release below
            __threadfence_system();
            st_release_sys(t.flags[p] + fb.start + rank, epoch);
        }
        if (!spin_until(my + fb.start + p, epoch, deadline, true)) my[FLG_ERROR] = 1u;
    }
    __syncthreads();
}
// End barrier: the last CTA of this rank to finish tells every peer "my stores have been issued and fenced" and waits
// until every peer said the same, so the next kernel on this stream sees all peers' pushes.
__device__ __forceinline__ void end_barrier(const PeerTable& t, const FlagBases fb, int rank, int world, unsigned int epoch,
                                            long long deadline) {
    unsigned int* my = t.flags[rank];
    __threadfence_system();
    __syncthreads();
    __shared__ unsigned int is_last;
    if (threadIdx.x == 0) is_last = atomicAdd(my + fb.done, 1u) == gridDim.x - 1 ? 1u : 0u;
    __syncthreads();
    if (is_last) {
        if (threadIdx.x == 0) my[fb.done] = 0u;
        cross_rank_barrier(t, fb.end, rank, world, epoch, deadline);
    }
}


}  // namespace fedb
