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
#include "peer_flags.cuh"
#include <string.h>
#include <stdlib.h>

namespace fedb {

__device__ __forceinline__ float4 ld_stream4(const float* base, size_t i) { return __ldcs(reinterpret_cast<const float4*>(base) + i); }

// ONESHOT = false: two-shot (rank r reduces slice r and pushes it to every rank's `out`) — bandwidth-optimal for whole models.
// ONESHOT = true : every rank reduces the WHOLE range and writes only its own `out` — one pass, for payloads of a few KB
//                  (Gram partials, scalars) where the second hop of the two-shot form costs more than the extra reads.
template <bool ONESHOT>
__global__ void __launch_bounds__(512) fedavg_reduce_kernel(const PeerTable t, int rank, int world, size_t n,
                                                            float my_weight, float inv_w_arg, unsigned int epoch,
                                                            long long timeout_cycles) {
    const FlagBases fb = ONESHOT ? FB_ONESHOT : FB_TWOSHOT;
    unsigned int* my = t.flags[rank];
    const long long deadline = clock64() + timeout_cycles;
    start_barrier(t, fb, rank, world, epoch, deadline, 1, my_weight);
    // sample-count weighting: every rank published its sum of n_k; the scale is 1 / sum over ranks
    float inv_w = inv_w_arg;
    bool skip = false;
    if (inv_w <= 0.f) {
        float tot = 0.f;
        for (int p = 0; p < world; ++p) tot += __uint_as_float(ld_acquire_gpu(my + FLG_WEIGHT + p));
        skip = !(tot > 0.f);             // every contributor masked out: keep the previous global model on every rank
        inv_w = skip ? 0.f : 1.f / tot;
    }
    // a start barrier that timed out means some rank's input is missing: do not push a partial sum over the old model
    if (ld_acquire_gpu(my + FLG_ERROR) != 0u) skip = true;
    if (!skip) {
        const size_t n4 = n / 4;
        const size_t per = ONESHOT ? n4 : (n4 + world - 1) / world;
        const size_t lo = ONESHOT ? 0 : per * rank, hi = lo + per < n4 ? lo + per : n4;
        const size_t stride = (size_t)gridDim.x * blockDim.x;
        auto reduce_store = [&](size_t i, const float4 (&v)[FEDB200_MAX_RANKS]) {
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int p = 0; p < FEDB200_MAX_RANKS; ++p)
                if (p < world) { s.x += v[p].x; s.y += v[p].y; s.z += v[p].z; s.w += v[p].w; }
            s.x *= inv_w; s.y *= inv_w; s.z *= inv_w; s.w *= inv_w;
            uint2 sh;
            sh.x = pack_bf16x2(s.x, s.y);
            sh.y = pack_bf16x2(s.z, s.w);
            if (ONESHOT) {
                reinterpret_cast<float4*>(t.out[rank])[i] = s;
                if (t.shadow[rank] != nullptr) reinterpret_cast<uint2*>(t.shadow[rank])[i] = sh;
            } else {
#pragma unroll
                for (int p = 0; p < FEDB200_MAX_RANKS; ++p) {
                    if (p < world) {
                        reinterpret_cast<float4*>(t.out[p])[i] = s;
                        if (t.shadow[p] != nullptr) reinterpret_cast<uint2*>(t.shadow[p])[i] = sh;
                    }
                }
            }
        };
        // two positions per thread and iteration: 2 x world 16-byte peer loads in flight per thread, which is what lets a
        // 32-CTA grid cover the ~2 us NVLink round trip at full bandwidth
        size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + stride < hi; i += 2 * stride) {
            float4 v0[FEDB200_MAX_RANKS], v1[FEDB200_MAX_RANKS];
#pragma unroll
            for (int p = 0; p < FEDB200_MAX_RANKS; ++p)
                if (p < world) { v0[p] = ld_stream4(t.acc[p], i); v1[p] = ld_stream4(t.acc[p], i + stride); }
            reduce_store(i, v0);
            reduce_store(i + stride, v1);
        }
        if (i < hi) {
            float4 v0[FEDB200_MAX_RANKS];
#pragma unroll
            for (int p = 0; p < FEDB200_MAX_RANKS; ++p)
                if (p < world) v0[p] = ld_stream4(t.acc[p], i);
            reduce_store(i, v0);
        }
        if (ONESHOT || rank == world - 1) {       // scalar tail (n not a multiple of 4): last rank (two-shot) / everyone
            for (size_t k = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
                float s = 0.f;
                for (int p = 0; p < world; ++p) s += t.acc[p][k];
                s *= inv_w;
                for (int p = 0; p < world; ++p) {
                    if (ONESHOT && p != rank) continue;
                    t.out[p][k] = s;
                    if (t.shadow[p] != nullptr) t.shadow[p][k] = __float2bfloat16_rn(s);
                }
            }
        }
    }
    // all pushes have landed everywhere (two-shot) / nobody's source is overwritten while a peer still reads it (one-shot)
    end_barrier(t, fb, rank, world, epoch, deadline);
}

// NVLS form of the two-shot reduce: the buffers are bound to a multicast object (cuMulticastCreate / BindMem — set up on the
// host by the symmetric-memory allocator), so ONE `multimem.ld_reduce` on the multicast address returns the sum over every
// GPU's replica, added inside the NVSwitch, and ONE `multimem.st` writes the result into every GPU's copy.  Each GPU then
// receives D / world reduced elements instead of (world - 1) x D / world operands and sends its result once instead of
// world - 1 times: the link carries about half the bytes of the P2P form.  Barriers, weights and the masked / timed-out
// behaviour are those of fedavg_reduce_kernel (same flag words, same epoch sequence).
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(float* mc, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

template <int U>
__global__ void __launch_bounds__(512) fedavg_reduce_nvls_kernel(const PeerTable t, const float* __restrict__ mc_acc,
                                                                 float* __restrict__ mc_out, int rank, int world, size_t n,
                                                                 float my_weight, float inv_w_arg, unsigned int epoch,
                                                                 long long timeout_cycles) {
    unsigned int* my = t.flags[rank];
    const long long deadline = clock64() + timeout_cycles;
    start_barrier(t, FB_TWOSHOT, rank, world, epoch, deadline, 1, my_weight);
    float inv_w = inv_w_arg;
    bool skip = false;
    if (inv_w <= 0.f) {
        float tot = 0.f;
        for (int p = 0; p < world; ++p) tot += __uint_as_float(ld_acquire_gpu(my + FLG_WEIGHT + p));
        skip = !(tot > 0.f);
        inv_w = skip ? 0.f : 1.f / tot;
    }
    if (ld_acquire_gpu(my + FLG_ERROR) != 0u) skip = true;
    if (!skip) {
        const size_t n4 = n / 4;
        const size_t per = (n4 + world - 1) / world;
        const size_t lo = per * rank, hi = lo + per < n4 ? lo + per : n4;
        const size_t stride = (size_t)gridDim.x * blockDim.x;
        size_t i = lo + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + (U - 1) * stride < hi; i += U * stride) {           // U switch reductions in flight per thread
            float4 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) v[u] = multimem_ld_reduce_add(mc_acc + 4 * (i + u * stride));
#pragma unroll
            for (int u = 0; u < U; ++u) {
                v[u].x *= inv_w; v[u].y *= inv_w; v[u].z *= inv_w; v[u].w *= inv_w;
                multimem_st(mc_out + 4 * (i + u * stride), v[u]);
            }
        }
        for (; i < hi; i += stride) {
            float4 v = multimem_ld_reduce_add(mc_acc + 4 * i);
            v.x *= inv_w; v.y *= inv_w; v.z *= inv_w; v.w *= inv_w;
            multimem_st(mc_out + 4 * i, v);
        }
        if (rank == world - 1) {                                       // scalar tail through the unicast (peer) pointers
            for (size_t k = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
                float s = 0.f;
                for (int p = 0; p < world; ++p) s += t.acc[p][k];
                s *= inv_w;
                for (int p = 0; p < world; ++p) t.out[p][k] = s;
            }
        }
    }
    end_barrier(t, FB_TWOSHOT, rank, world, epoch, deadline);
}

// Slot exchange: every rank owns slot `rank` of a [world, slot_words] table that exists on every rank (symmetric buffer).
// One small kernel pushes the own slot into every peer's copy and runs a cross-rank barrier, after which every rank holds
// the whole table locally — the device-side replacement of a host all_gather of per-round metadata (Merkle roots, sample
// counts, evaluation sums), read back with the round's other results.
__global__ void __launch_bounds__(256) peer_exchange_kernel(const PeerTable t, int rank, int world, int slot_words,
                                                            unsigned int epoch, long long timeout_cycles) {
    const long long deadline = clock64() + timeout_cycles;
    const unsigned int* mine = reinterpret_cast<const unsigned int*>(t.acc[rank]) + (size_t)rank * slot_words;
    for (int p = 0; p < world; ++p) {
        if (p == rank) continue;
        unsigned int* dst = reinterpret_cast<unsigned int*>(t.out[p]) + (size_t)rank * slot_words;
        for (int i = threadIdx.x; i < slot_words; i += blockDim.x) dst[i] = mine[i];
    }
    __threadfence_system();
    __syncthreads();
    cross_rank_barrier(t, FLG_XCHG, rank, world, epoch, deadline);
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

static int fill_table(PeerTable& t, const void* const* acc_ptrs, void* const* out_ptrs, void* const* shadow_ptrs,
                      unsigned int* const* flag_ptrs, int world) {
    for (int p = 0; p < FEDB200_MAX_RANKS; ++p) {
        t.acc[p] = (p < world && acc_ptrs) ? (const float*)acc_ptrs[p] : nullptr;
        t.out[p] = (p < world && out_ptrs) ? (float*)out_ptrs[p] : nullptr;
        t.shadow[p] = (p < world && shadow_ptrs) ? (__nv_bfloat16*)shadow_ptrs[p] : nullptr;
        t.flags[p] = p < world ? flag_ptrs[p] : nullptr;
    }
    return 0;
}

extern "C" int fedb200_fedavg_reduce_ex(const float* const* acc_ptrs, float* const* out_ptrs, void* const* shadow_ptrs,
                                        unsigned int* const* flag_ptrs, int rank, int world, size_t n, float my_weight,
                                        float inv_total_weight, unsigned int epoch, int timeout_ms, int oneshot,
                                        int max_ctas, void* stream) {
    FEDB_REQUIRE(world >= 1 && world <= FEDB200_MAX_RANKS && rank >= 0 && rank < world, "fedavg_reduce: rank %d / world %d", rank, world);
    PeerTable t;
    fill_table(t, (const void* const*)acc_ptrs, (void* const*)out_ptrs, shadow_ptrs, flag_ptrs, world);
    for (int p = 0; p < world; ++p)
        FEDB_REQUIRE(((uintptr_t)t.acc[p] & 15) == 0 && ((uintptr_t)t.out[p] & 15) == 0, "fedavg_reduce: 16-byte alignment");
    // No CTA of this kernel waits for another CTA of the same launch (claim-based start barrier), so any grid size is safe.
    // Default: one CTA per SM when the kernel has the GPU to itself; callers that overlap it with compute pass max_ctas.
    const size_t n4 = n / 4;
    const size_t work = oneshot ? n4 : (n4 + world - 1) / world;
    int grid = (int)((work + 2 * 512 - 1) / (2 * 512));
    const int cap = max_ctas > 0 ? max_ctas : sm_count();
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    if (oneshot)
        fedavg_reduce_kernel<true><<<grid, 512, 0, (cudaStream_t)stream>>>(t, rank, world, n, my_weight, inv_total_weight, epoch,
                                                                          timeout_to_cycles(timeout_ms));
    else
        fedavg_reduce_kernel<false><<<grid, 512, 0, (cudaStream_t)stream>>>(t, rank, world, n, my_weight, inv_total_weight, epoch,
                                                                           timeout_to_cycles(timeout_ms));
    FEDB_LAUNCH_CHECK();
    return 0;
}

extern "C" int fedb200_fedavg_reduce(const float* const* acc_ptrs, float* const* out_ptrs, void* const* shadow_ptrs,
                                     unsigned int* const* flag_ptrs, int rank, int world, size_t n, float my_weight,
                                     float inv_total_weight, unsigned int epoch, int timeout_ms, void* stream) {
    return fedb200_fedavg_reduce_ex(acc_ptrs, out_ptrs, shadow_ptrs, flag_ptrs, rank, world, n, my_weight, inv_total_weight,
                                    epoch, timeout_ms, 0, 0, stream);
}

// NVLS variant.  acc_ptrs / out_ptrs / flag_ptrs as for fedb200_fedavg_reduce_ex (the unicast peer pointers of the same
// buffers: used for the scalar tail); mc_acc / mc_out: the MULTICAST addresses of the accumulator and output buffers (every
// rank passes its own mapping of the same multicast object).  Shares the two-shot epoch sequence.
extern "C" int fedb200_fedavg_reduce_nvls(const float* const* acc_ptrs, float* const* out_ptrs, const float* mc_acc, float* mc_out,
                                          unsigned int* const* flag_ptrs, int rank, int world, size_t n, float my_weight,
                                          float inv_total_weight, unsigned int epoch, int timeout_ms, int max_ctas,
                                          void* stream) {
    FEDB_REQUIRE(world >= 1 && world <= FEDB200_MAX_RANKS && rank >= 0 && rank < world, "fedavg_reduce_nvls: rank %d / world %d", rank, world);
    FEDB_REQUIRE(mc_acc != nullptr && mc_out != nullptr, "fedavg_reduce_nvls: buffers are not bound to a multicast object");
    FEDB_REQUIRE((((uintptr_t)mc_acc | (uintptr_t)mc_out) & 15) == 0, "fedavg_reduce_nvls: 16-byte alignment");
    PeerTable t;
    fill_table(t, (const void* const*)acc_ptrs, (void* const*)out_ptrs, nullptr, flag_ptrs, world);
    const size_t work = (n / 4 + world - 1) / world;
    int grid = (int)((work + 4 * 512 - 1) / (4 * 512));
    const int cap = max_ctas > 0 ? max_ctas : sm_count();
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    // measured at 8 x B200, 433 MB: 4 reductions in flight per thread 0.98 ms, 8 in flight 1.03 ms; a 32-CTA grid 0.95 ms
    static const int unroll = [] { const char* e = getenv("FEDB200_NVLS_UNROLL"); return e ? atoi(e) : 4; }();
    if (unroll >= 8)
        fedavg_reduce_nvls_kernel<8><<<grid, 512, 0, (cudaStream_t)stream>>>(t, mc_acc, mc_out, rank, world, n, my_weight,
                                                                            inv_total_weight, epoch, timeout_to_cycles(timeout_ms));
    else
        fedavg_reduce_nvls_kernel<4><<<grid, 512, 0, (cudaStream_t)stream>>>(t, mc_acc, mc_out, rank, world, n, my_weight,
                                                                            inv_total_weight, epoch, timeout_to_cycles(timeout_ms));
    FEDB_LAUNCH_CHECK();
    return 0;
}

extern "C" int fedb200_peer_exchange(void* const* table_ptrs, unsigned int* const* flag_ptrs, int rank, int world,
                                     int slot_bytes, unsigned int epoch, int timeout_ms, void* stream) {
    FEDB_REQUIRE(world >= 1 && world <= FEDB200_MAX_RANKS && rank >= 0 && rank < world, "peer_exchange: rank %d / world %d", rank, world);
    FEDB_REQUIRE(slot_bytes > 0 && slot_bytes % 4 == 0, "peer_exchange: slot_bytes must be a positive multiple of 4");
    PeerTable t;
    fill_table(t, (const void* const*)table_ptrs, table_ptrs, nullptr, flag_ptrs, world);
    peer_exchange_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(t, rank, world, slot_bytes / 4, epoch, timeout_to_cycles(timeout_ms));
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
