// Shared by the 1-CTA and 2-CTA tcgen05 GEMM kernels: parameter block and the fused epilogue of one output tile.
#pragma once
#include "common.cuh"
#include "../include/fedb200.h"

namespace fedb {

struct GemmParams {
    int M, N, K;
    int splits;             // split-K factor (EPI_ATOMIC only)
    int epi;
    void* out;              // bf16 [M, ldo] (fp32 for EPI_ATOMIC)
    void* out2;             // bf16 [M, ldo]: GELU output for EPI_BIAS_GELU
    int ldo;
    const float* bias;      // [N] or null
    const __nv_bfloat16* aux;   // residual / pre-activation, [M, ld_aux]
    int ld_aux;
    int act;                // 0 = erf GELU (BERT), 1 = tanh GELU (ALBERT gelu_new)
    unsigned long long seed;
    uint32_t keep_thresh;
    float inv_keep;
    int use_dropout;
    int tma_store;          // pair kernel: stage bf16 outputs through smem + TMA store (0 = direct row-per-thread stores)
    long long split_stride; // EPI_ATOMIC: split s accumulates into out + s * split_stride (elements); 0 = all splits share `out`
};

__device__ __forceinline__ void store_bf16x32(__nv_bfloat16* dst, const float (&v)[32], int ncols_valid) {
    // dst is 16-byte aligned (column offsets are multiples of 32, ldo multiple of 8)
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        if (q * 8 < ncols_valid) {
            uint4 o;
            o.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]); o.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
            o.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]); o.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
            *reinterpret_cast<uint4*>(dst + q * 8) = o;
        }
    }
}
__device__ __forceinline__ void load_bf16x32(const __nv_bfloat16* src, float (&v)[32], int ncols_valid) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        if (q * 8 < ncols_valid) {
            const uint4 r = *reinterpret_cast<const uint4*>(src + q * 8);
            const float2 a = unpack_bf16x2(r.x), b = unpack_bf16x2(r.y), c = unpack_bf16x2(r.z), d = unpack_bf16x2(r.w);
            v[q * 8 + 0] = a.x; v[q * 8 + 1] = a.y; v[q * 8 + 2] = b.x; v[q * 8 + 3] = b.y;
            v[q * 8 + 4] = c.x; v[q * 8 + 5] = c.y; v[q * 8 + 6] = d.x; v[q * 8 + 7] = d.y;
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[q * 8 + k] = 0.f;
        }
    }
}

// Epilogue of one [128 x BN] accumulator tile for the calling warp: `tmem_acc` addresses this warp's lane quadrant of
// the accumulator buffer, `row` is the thread's global output row, `col_base` the tile's first output column, `half`
// selects even / odd 32-column chunks (two warps share a lane quadrant).  Waits on `acc_full` itself so the first
// residual chunk can be prefetched while the MMAs are still running.
template <int BN, int EPI = -1>
__device__ __forceinline__ void gemm_epilogue_tile(const GemmParams& p, uint32_t tmem_acc, int row, int col_base, int half,
                                                   bool has_k, uint64_t* acc_full, uint32_t parity, int split = 0) {
    constexpr int N_CHUNKS = BN / 32;
    const int epi = EPI >= 0 ? EPI : p.epi;      // compile-time in the specialised kernels: the dispatch below folds away
    const bool needs_aux = p.aux != nullptr && (epi == FEDB200_EPI_BIAS_DROP_RES || epi == FEDB200_EPI_MUL_GELU_GRAD ||
                                                epi == FEDB200_EPI_ADD);
    const bool row_ok = row < p.M;
    // prefetch the first aux chunk (and the chunk's bias: one coalesced 128-byte load per warp, later broadcast with
    // shuffles) while the accumulator is still being produced
    const int lane_id = threadIdx.x & 31;
    uint4 aux_raw[4];
    float bias_lane = 0.f;
    auto fetch_aux = [&](int c) {
        const int col0 = col_base + c * 32;
        bias_lane = (p.bias != nullptr && c < N_CHUNKS && col0 + lane_id < p.N) ? __ldg(p.bias + col0 + lane_id) : 0.f;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            aux_raw[q] = make_uint4(0u, 0u, 0u, 0u);
            if (needs_aux && row_ok && c < N_CHUNKS && col0 + q * 8 < p.N)
                aux_raw[q] = __ldg(reinterpret_cast<const uint4*>(p.aux + (size_t)row * p.ld_aux + col0) + q);
        }
    };
    fetch_aux(half);
    mbar_wait(acc_full, parity);
    tc_fence_after_sync();
#pragma unroll 1
    for (int c = half; c < N_CHUNKS; c += 2) {
        const int col0 = col_base + c * 32;
        if (col0 >= p.N) break;                  // warp-uniform
        uint32_t raw[32];
        tmem_ld_32x32(tmem_acc + c * 32, raw);
        float r[32];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float2 a0 = unpack_bf16x2(aux_raw[q].x), a1 = unpack_bf16x2(aux_raw[q].y);
            const float2 a2 = unpack_bf16x2(aux_raw[q].z), a3 = unpack_bf16x2(aux_raw[q].w);
            r[q * 8 + 0] = a0.x; r[q * 8 + 1] = a0.y; r[q * 8 + 2] = a1.x; r[q * 8 + 3] = a1.y;
            r[q * 8 + 4] = a2.x; r[q * 8 + 5] = a2.y; r[q * 8 + 6] = a3.x; r[q * 8 + 7] = a3.y;
        }
        const float bias_cur = bias_lane;
        fetch_aux(c + 2);                        // next chunk's residual / pre-activation is in flight during the math
        tmem_ld_wait();
        float bv[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) bv[j] = p.bias != nullptr ? __shfl_sync(0xffffffffu, bias_cur, j) : 0.f;
        const int nvalid = min(32, p.N - col0);
        if (row_ok) {
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = (has_k ? __uint_as_float(raw[j]) : 0.f) + bv[j];
            if (epi == FEDB200_EPI_DISCARD) {
                if (v[0] == 123456.789f) reinterpret_cast<float*>(p.out)[0] = v[1];      // never true: keeps the loads alive
            } else if (epi == FEDB200_EPI_ATOMIC_F32) {
                // reproducible mode: every split owns a private [M, ldo] partial (one writer per element), summed afterwards
                float* o = reinterpret_cast<float*>(p.out) + (size_t)split * (size_t)p.split_stride + (size_t)row * p.ldo + col0;
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (q * 4 < nvalid)
                        atomicAdd(reinterpret_cast<float4*>(o + q * 4),
                                  make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]));
            } else {
                __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)row * p.ldo + col0;
                if (epi == FEDB200_EPI_BIAS) {
                    store_bf16x32(o, v, nvalid);
                } else if (epi == FEDB200_EPI_BIAS_DROP_RES) {
                    if (p.use_dropout) {
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            dropout_apply8(v + q * 8, p.seed, (uint32_t)row, (uint32_t)(col0 + q * 8), p.keep_thresh, p.inv_keep);
                    }
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] += r[j];
                    store_bf16x32(o, v, nvalid);
                } else if (epi == FEDB200_EPI_BIAS_GELU) {
                    if (p.out != nullptr) store_bf16x32(o, v, nvalid);      // inference keeps only the activated output
                    // activate the bf16-rounded pre-activation so backward (which re-reads it) is consistent
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __bfloat162float(__float2bfloat16_rn(v[j]));
                    // the activation choice is hoisted out of the element loop: a per-element branch serialises the 32
                    // independent dependency chains (it cost ~40 us per FFN GEMM)
                    if (p.act) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = gelu_tanh(v[j]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
                    }
                    store_bf16x32(reinterpret_cast<__nv_bfloat16*>(p.out2) + (size_t)row * p.ldo + col0, v, nvalid);
                } else if (epi == FEDB200_EPI_MUL_GELU_GRAD) {
                    if (p.act) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] *= gelu_tanh_grad(r[j]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] *= gelu_erf_grad(r[j]);
                    }
                    store_bf16x32(o, v, nvalid);
                } else {    // FEDB200_EPI_ADD
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] += r[j];       // r is zero when there is no aux
                    store_bf16x32(o, v, nvalid);
                }
            }
        }
        __syncwarp();
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// Same epilogue with the bf16 outputs staged through shared memory and written by TMA (pair kernel).
// A row-per-thread 16-byte store touches 32 different cache lines per warp instruction; two outputs of a 128 x 256 tile
// cost ~8k LSU wavefronts, more than the tile's MMA time, so the GELU GEMMs were epilogue-bound.  Here each warp packs a
// [32 rows x 64 cols] bf16 block into a 128B-swizzled 4 KB slot (conflict-free 16-byte st.shared) and one lane issues a
// cp.async.bulk.tensor store; rows >= M and columns >= N are clipped by the tensor map.  `stage` is the warp's private
// 8 KB (two slots, 1024-byte aligned): two-output epilogues use one slot per output, single-output ones alternate.
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void stage_bf16x32(uint32_t slot, int lane, int sub, const float (&v)[32]) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const uint32_t j = (uint32_t)(sub * 4 + q);
        const uint32_t addr = slot + (uint32_t)lane * 128u + ((j ^ ((uint32_t)lane & 7u)) << 4);
        st_shared_v4(addr, pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]), pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]),
                     pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]), pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]));
    }
}

template <int BN, int EPI = -1>
__device__ __forceinline__ void gemm_epilogue_tile_tma(const GemmParams& p, const CUtensorMap* tm_out,
                                                       const CUtensorMap* tm_out2, const CUtensorMap* tm_aux, uint32_t stage,
                                                       uint32_t& slot_it, uint64_t* aux_bar, uint32_t& aux_phase,
                                                       uint32_t tmem_acc, int row, int col_base, int half, bool has_k,
                                                       uint64_t* acc_full, uint32_t parity, int split = 0) {
    const int epi = EPI >= 0 ? EPI : p.epi;      // compile-time in the specialised kernels: the dispatch below folds away
    if (epi == FEDB200_EPI_ATOMIC_F32 || epi == FEDB200_EPI_DISCARD || !p.tma_store) {
        gemm_epilogue_tile<BN, EPI>(p, tmem_acc, row, col_base, half, has_k, acc_full, parity, split);
        return;
    }
    constexpr int N_CHUNKS = BN / 32;
    const bool needs_aux = p.aux != nullptr && (epi == FEDB200_EPI_BIAS_DROP_RES || epi == FEDB200_EPI_MUL_GELU_GRAD ||
                                                epi == FEDB200_EPI_ADD);
    const bool row_ok = row < p.M;
    const int lane_id = threadIdx.x & 31;
    const bool two_out = epi == FEDB200_EPI_BIAS_GELU && p.out != nullptr;
    uint4 aux_raw[4];
    float bias_lane = 0.f;
    // Residual / pre-activation tiles arrive by TMA into the SAME swizzled slots the result is later stored from (each
    // thread reads its row's 16-byte chunks and overwrites them in place), issued for the whole tile before the
    // accumulator is even complete; row-per-thread global loads were the epilogue's main stall (long scoreboard).
    const bool aux_tma = needs_aux;
    auto fetch_aux = [&](int c) {
        const int col0 = col_base + c * 32;
        bias_lane = (p.bias != nullptr && c < N_CHUNKS && col0 + lane_id < p.N) ? __ldg(p.bias + col0 + lane_id) : 0.f;
#pragma unroll
        for (int q = 0; q < 4; ++q) aux_raw[q] = make_uint4(0u, 0u, 0u, 0u);
    };
    if (aux_tma) {
        if (lane_id == 0) {
            tma_store_wait_read<0>();            // both slots: last tile's stores have been read out
            uint32_t k = 0;
            for (int g = half; g < N_CHUNKS / 2; g += 2, ++k) {
                const int gcol0 = col_base + g * 64;
                if (gcol0 >= p.N) break;
                const uint32_t sl = (slot_it + k) & 1u;
                mbar_expect_tx(&aux_bar[sl], 4096u);
                tma_load_smem_2d(stage + sl * 4096u, tm_aux, &aux_bar[sl], gcol0, row);
            }
        }
        __syncwarp();
    }
    fetch_aux(2 * half);
    mbar_wait(acc_full, parity);
    tc_fence_after_sync();
#pragma unroll 1
    for (int g = half; g < N_CHUNKS / 2; g += 2) {
        const int gcol0 = col_base + g * 64;
        if (gcol0 >= p.N) break;                 // warp-uniform
        // claim the staging slot(s): the bulk stores that last read them must have finished reading
        uint32_t slot_a, slot_b = 0;
        if (two_out) {
            if (lane_id == 0) tma_store_wait_read<0>();
            slot_a = stage; slot_b = stage + 4096u;
        } else {
            if (!aux_tma && lane_id == 0) tma_store_wait_read<1>();
            const uint32_t sl = slot_it & 1u;
            slot_a = stage + sl * 4096u;
            ++slot_it;
            if (aux_tma) {
                mbar_wait(&aux_bar[sl], (aux_phase >> sl) & 1u);
                aux_phase ^= 1u << sl;
            }
        }
        __syncwarp();
#pragma unroll
        for (int sub = 0; sub < 2; ++sub) {
            const int c = 2 * g + sub;
            const int col0 = col_base + c * 32;
            if (col0 >= p.N) break;              // warp-uniform; the tensor map clips the untouched half of the slot
            uint32_t raw[32];
            tmem_ld_32x32(tmem_acc + c * 32, raw);
            float r[32];
            if (epi == FEDB200_EPI_BIAS_DROP_RES || epi == FEDB200_EPI_MUL_GELU_GRAD || epi == FEDB200_EPI_ADD) {
                if (aux_tma) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint32_t j = (uint32_t)(sub * 4 + q);
                        aux_raw[q] = ld_shared_v4(slot_a + (uint32_t)lane_id * 128u + ((j ^ ((uint32_t)lane_id & 7u)) << 4));
                    }
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float2 a0 = unpack_bf16x2(aux_raw[q].x), a1 = unpack_bf16x2(aux_raw[q].y);
                    const float2 a2 = unpack_bf16x2(aux_raw[q].z), a3 = unpack_bf16x2(aux_raw[q].w);
                    r[q * 8 + 0] = a0.x; r[q * 8 + 1] = a0.y; r[q * 8 + 2] = a1.x; r[q * 8 + 3] = a1.y;
                    r[q * 8 + 4] = a2.x; r[q * 8 + 5] = a2.y; r[q * 8 + 6] = a3.x; r[q * 8 + 7] = a3.y;
                }
            }
            const float bias_cur = bias_lane;
            fetch_aux(sub == 0 ? c + 1 : c + 3);     // next chunk this warp will process
            tmem_ld_wait();
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(raw[j]);
            if (!has_k) {                            // degenerate split: nothing was accumulated (warp-uniform)
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = 0.f;
            }
            if (p.bias != nullptr) {                 // warp-uniform
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] += __shfl_sync(0xffffffffu, bias_cur, j);
            }
            if (epi == FEDB200_EPI_BIAS) {
                stage_bf16x32(slot_a, lane_id, sub, v);
            } else if (epi == FEDB200_EPI_BIAS_DROP_RES) {
                if (p.use_dropout) {
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        dropout_apply8(v + q * 8, p.seed, (uint32_t)row, (uint32_t)(col0 + q * 8), p.keep_thresh, p.inv_keep);
                }
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] += r[j];
                stage_bf16x32(slot_a, lane_id, sub, v);
            } else if (epi == FEDB200_EPI_BIAS_GELU) {
                if (two_out) stage_bf16x32(slot_a, lane_id, sub, v);
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __bfloat162float(__float2bfloat16_rn(v[j]));
                if (p.act) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = gelu_tanh(v[j]);
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
                }
                stage_bf16x32(two_out ? slot_b : slot_a, lane_id, sub, v);
            } else if (epi == FEDB200_EPI_MUL_GELU_GRAD) {
                if (p.act) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] *= gelu_tanh_grad(r[j]);
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] *= gelu_erf_grad(r[j]);
                }
                stage_bf16x32(slot_a, lane_id, sub, v);
            } else {    // FEDB200_EPI_ADD
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] += r[j];
                stage_bf16x32(slot_a, lane_id, sub, v);
            }
        }
        fence_async_smem();
        __syncwarp();
        if (lane_id == 0 && row < p.M && p.tma_store != 2) {   // lane 0 holds the warp's first row; later rows are clipped by the map (2 = profiling: stage only)
            const int row0 = row;
            if (epi == FEDB200_EPI_BIAS_GELU) {
                if (two_out) tma_store_smem_2d(tm_out, slot_a, gcol0, row0);
                tma_store_smem_2d(tm_out2, two_out ? slot_b : slot_a, gcol0, row0);
            } else {
                tma_store_smem_2d(tm_out, slot_a, gcol0, row0);
            }
            tma_store_commit();
        }
    }
}

}  // namespace fedb
