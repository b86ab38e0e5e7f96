// Peer-memory runtime over NVLink 5 / NVSwitch and the fused FedAvg reduce (K1).
//
// * symmetric buffers: every rank cudaMalloc's the same set of buffers and exports them with CUDA IPC; after
//   the handle exchange (through torch.distributed on the host) each rank holds a table of peer-mapped
//   pointers, so kernels address any peer's arena with plain ld/st.global over NVLink;
// * signalling: per-rank flag words in peer memory, written with st.release.sys and polled with
//   ld.acquire.sys; epochs only ever increase (no resets), every spin has a clock64() deadline so a protocol
//   bug raises an error word instead of hanging the GPU;
// * fedavg_reduce: ONE launch = cross-rank start barrier -> two-shot reduction (rank r sums slice r of every
//   rank's accumulator, scales by 1 / sum(n_k), casts) -> push of the result slice into every rank's global
//   model (+ bf16 shadow) -> cross-rank end barrier.  Replaces Flower's numpy aggregate + pickled broadcast
//   (src/Servercase/server_IID_IMDB.py:205-218) and `sum(p)/len(p)` + load_state_dict
//   (src/Serverlesscase/serverless_IID_IMDB.py:269-270).
#include "common.cuh"
#include "../include/fedb200.h"
#include <string.h>

namespace fedb {

// flag-word layout inside each rank's flag buffer (uint32 words)
constexpr int FLG_START = 0;                       // [FLG_START + p]: start-barrier epoch written by rank p
constexpr int FLG_END = FEDB200_MAX_RANKS;         // [FLG_END + p]  : end-barrier epoch written by rank p
constexpr int FLG_GO = 2 * FEDB200_MAX_RANKS;      // local: block 0 -> other blocks of this rank
constexpr int FLG_DONE = FLG_GO + 1;               // local: finished-block counter
constexpr int FLG_ERROR = FLG_GO + 2;              // local: set to 1 when a spin timed out
constexpr int FLG_BAR = FLG_GO + 8;                // [FLG_BAR + p]: standalone barrier epochs
constexpr int FLG_WEIGHT = FLG_BAR + FEDB200_MAX_RANKS;   // [FLG_WEIGHT + p]: rank p's aggregation weight (float bits)

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

__global__ void __launch_bounds__(512) fedavg_reduce_kernel(const PeerTable t, int rank, int world, size_t n,
                                                            float my_weight, float inv_w_arg, unsigned int epoch,
                                                            long long timeout_cycles) {
    unsigned int* my = t.flags[rank];
    const long long deadline = clock64() + timeout_cycles;
    // ---- start barrier: every rank's accumulator is complete ------------------------------------------
    if (blockIdx.x == 0) {
        cross_rank_barrier(t, FLG_START, rank, world, epoch, deadline, 1, my_weight);
        __syncthreads();
        if (threadIdx.x == 0) st_release_gpu(my + FLG_GO, epoch);
    } else {
        if (threadIdx.x == 0 && !spin_until(my + FLG_GO, epoch, deadline, false)) my[FLG_ERROR] = 1u;
        __syncthreads();
    }
    // sample-count weighting: every rank published its sum of n_k; the scale is 1 / sum over ranks
    float inv_w = inv_w_arg;
    if (inv_w <= 0.f) {
        float tot = 0.f;
        for (int p = 0; p < world; ++p) tot += __uint_as_float(ld_acquire_gpu(my + FLG_WEIGHT + p));
        inv_w = tot > 0.f ? 1.f / tot : 0.f;
    }
    // ---- two-shot: reduce my slice from all peers, push it to all peers ---------------------------------
    const size_t n4 = n / 4;
    const size_t per = (n4 + world - 1) / world;
    const size_t lo = per * rank, hi = lo + per < n4 ? lo + per : n4;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += stride) {
        float4 v[FEDB200_MAX_RANKS];
#pragma unroll
        for (int p = 0; p < FEDB200_MAX_RANKS; ++p)
            if (p < world) v[p] = __ldcs(reinterpret_cast<const float4*>(t.acc[p]) + i);
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int p = 0; p < FEDB200_MAX_RANKS; ++p)
            if (p < world) { s.x += v[p].x; s.y += v[p].y; s.z += v[p].z; s.w += v[p].w; }
        s.x *= inv_w; s.y *= inv_w; s.z *= inv_w; s.w *= inv_w;
        uint2 sh;
        sh.x = pack_bf16x2(s.x, s.y);
        sh.y = pack_bf16x2(s.z, s.w);
#pragma unroll
        for (int p = 0; p < FEDB200_MAX_RANKS; ++p) {
            if (p < world) {
                reinterpret_cast<float4*>(t.out[p])[i] = s;
                if (t.shadow[p] != nullptr) reinterpret_cast<uint2*>(t.shadow[p])[i] = sh;
            }
        }
    }
    if (rank == world - 1) {       // scalar tail (n not a multiple of 4) is owned by the last rank
        for (size_t i = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
            float s = 0.f;
            for (int p = 0; p < world; ++p) s += t.acc[p][i];
            s *= inv_w;
            for (int p = 0; p < world; ++p) {
                t.out[p][i] = s;
                if (t.shadow[p] != nullptr) t.shadow[p][i] = __float2bfloat16_rn(s);
            }
        }
    }
    // ---- end barrier: all pushes have landed everywhere before any rank's next kernel reads `out` --------
    __threadfence_system();
    __syncthreads();
    __shared__ unsigned int is_last;
    if (threadIdx.x == 0) is_last = atomicAdd(my + FLG_DONE, 1u) == gridDim.x - 1 ? 1u : 0u;
    __syncthreads();
    if (is_last) {
        if (threadIdx.x == 0) my[FLG_DONE] = 0u;
        cross_rank_barrier(t, FLG_END, rank, world, epoch, deadline);
    }
}

__global__ void __launch_bounds__(32) peer_barrier_kernel(const PeerTable t, int rank, int world, unsigned int epoch,
                                                          long long timeout_cycles) {
    cross_rank_barrier(t, FLG_BAR, rank, world, epoch, clock64() + timeout_cycles);
}

// Single-word progress flags for the asynchronous gossip executor (K3): `pub[p]` / `done[p]` epochs are written by
// the owning rank with a system-scope release and polled by consumers on other GPUs with acquire loads.
__global__ void flag_set_kernel(unsigned int* flag, unsigned int value) {
    __threadfence_system();
    st_release_sys(flag, value);
}
__global__ void flag_wait_kernel(const unsigned int* flag, unsigned int value, long long timeout_cycles, unsigned int* err) {
    if (!spin_until(flag, value, clock64() + timeout_cycles, true) && err != nullptr) *err = 1u;
}

}  // namespace fedb

using namespace fedb;

extern "C" int fedb200_peer_alloc(size_t bytes, void** dev_ptr, void* ipc_handle_out64) {
    FEDB_REQUIRE(bytes > 0 && dev_ptr != nullptr && ipc_handle_out64 != nullptr, "peer_alloc: bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is expected to be 64 bytes");
    void* p = nullptr;
    FEDB_CHECK(cudaMalloc(&p, bytes));
    FEDB_CHECK(cudaMemset(p, 0, bytes));
    cudaIpcMemHandle_t h;
    FEDB_CHECK(cudaIpcGetMemHandle(&h, p));
    memcpy(ipc_handle_out64, &h, 64);
    *dev_ptr = p;
    return 0;
}

extern "C" int fedb200_peer_open(const void* ipc_handle64, void** mapped_ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle64, 64);
    FEDB_CHECK(cudaIpcOpenMemHandle(mapped_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}

extern "C" int fedb200_peer_close(void* mapped_ptr) {
    FEDB_CHECK(cudaIpcCloseMemHandle(mapped_ptr));
    return 0;
}

extern "C" int fedb200_peer_free(void* dev_ptr) {
    FEDB_CHECK(cudaFree(dev_ptr));
    return 0;
}

static long long timeout_to_cycles(int timeout_ms) {
    if (timeout_ms <= 0) timeout_ms = 20000;
    return (long long)timeout_ms * 2000000LL;      // ~2 GHz SM clock
}

extern "C" int fedb200_fedavg_reduce(const float* const* acc_ptrs, float* const* out_ptrs, void* const* shadow_ptrs,
                                     unsigned int* const* flag_ptrs, int rank, int world, size_t n, float my_weight,
                                     float inv_total_weight, unsigned int epoch, int timeout_ms, void* stream) {
    FEDB_REQUIRE(world >= 1 && world <= FEDB200_MAX_RANKS && rank >= 0 && rank < world, "fedavg_reduce: rank %d / world %d", rank, world);
    PeerTable t;
    for (int p = 0; p < FEDB200_MAX_RANKS; ++p) {
        t.acc[p] = p < world ? acc_ptrs[p] : nullptr;
        t.out[p] = p < world ? out_ptrs[p] : nullptr;
        t.shadow[p] = (p < world && shadow_ptrs) ? (__nv_bfloat16*)shadow_ptrs[p] : nullptr;
        t.flags[p] = p < world ? flag_ptrs[p] : nullptr;
        if (p < world) FEDB_REQUIRE(((uintptr_t)t.acc[p] & 15) == 0 && ((uintptr_t)t.out[p] & 15) == 0, "fedavg_reduce: 16-byte alignment");
    }
    // every block must be co-resident (blocks of one rank wait on block 0): one 512-thread block per SM
    const int grid = sm_count();
    fedavg_reduce_kernel<<<grid, 512, 0, (cudaStream_t)stream>>>(t, rank, world, n, my_weight, inv_total_weight,
                                                                 epoch, timeout_to_cycles(timeout_ms));
    FEDB_LAUNCH_CHECK();
    return 0;
}

extern "C" int fedb200_peer_barrier(unsigned int* const* flag_ptrs, int rank, int world, unsigned int epoch,
                                    int timeout_ms, void* stream) {
    FEDB_REQUIRE(world >= 1 && world <= FEDB200_MAX_RANKS && rank >= 0 && rank < world, "peer_barrier: rank %d / world %d", rank, world);
    PeerTable t;
    for (int p = 0; p < FEDB200_MAX_RANKS; ++p) {
        t.acc[p] = nullptr; t.out[p] = nullptr; t.shadow[p] = nullptr;
        t.flags[p] = p < world ? flag_ptrs[p] : nullptr;
    }
    peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(t, rank, world, epoch, timeout_to_cycles(timeout_ms));
    FEDB_LAUNCH_CHECK();
    return 0;
}

// Publish / await one monotone progress word (device-side, stream-ordered; no host involvement).
extern "C" int fedb200_flag_set(unsigned int* flag, unsigned int value, void* stream) {
    flag_set_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(flag, value);
    FEDB_LAUNCH_CHECK();
    return 0;
}

extern "C" int fedb200_flag_wait(const unsigned int* flag, unsigned int value, int timeout_ms, unsigned int* err,
                                 void* stream) {
    flag_wait_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(flag, value, timeout_to_cycles(timeout_ms), err);
    FEDB_LAUNCH_CHECK();
    return 0;
}
