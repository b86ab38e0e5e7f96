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
template <int BN>
__device__ __forceinline__ void gemm_epilogue_tile(const GemmParams& p, uint32_t tmem_acc, int row, int col_base, int half,
                                                   bool has_k, uint64_t* acc_full, uint32_t parity) {
    constexpr int N_CHUNKS = BN / 32;
    const bool needs_aux = p.aux != nullptr && (p.epi == FEDB200_EPI_BIAS_DROP_RES || p.epi == FEDB200_EPI_MUL_GELU_GRAD ||
                                                p.epi == FEDB200_EPI_ADD);
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
            if (p.epi == FEDB200_EPI_DISCARD) {
                if (v[0] == 123456.789f) reinterpret_cast<float*>(p.out)[0] = v[1];      // never true: keeps the loads alive
            } else if (p.epi == FEDB200_EPI_ATOMIC_F32) {
                float* o = reinterpret_cast<float*>(p.out) + (size_t)row * p.ldo + col0;
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (q * 4 < nvalid)
                        atomicAdd(reinterpret_cast<float4*>(o + q * 4),
                                  make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]));
            } else {
                __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + (size_t)row * p.ldo + col0;
                if (p.epi == FEDB200_EPI_BIAS) {
                    store_bf16x32(o, v, nvalid);
                } else if (p.epi == FEDB200_EPI_BIAS_DROP_RES) {
                    if (p.use_dropout) {
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            dropout_apply8(v + q * 8, p.seed, (uint32_t)row, (uint32_t)(col0 + q * 8), p.keep_thresh, p.inv_keep);
                    }
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] += r[j];
                    store_bf16x32(o, v, nvalid);
                } else if (p.epi == FEDB200_EPI_BIAS_GELU) {
                    if (p.out != nullptr) store_bf16x32(o, v, nvalid);      // inference keeps only the activated output
                    // activate the bf16-rounded pre-activation so backward (which re-reads it) is consistent
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float u = __bfloat162float(__float2bfloat16_rn(v[j]));
                        v[j] = p.act ? gelu_tanh(u) : gelu_erf(u);
                    }
                    store_bf16x32(reinterpret_cast<__nv_bfloat16*>(p.out2) + (size_t)row * p.ldo + col0, v, nvalid);
                } else if (p.epi == FEDB200_EPI_MUL_GELU_GRAD) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] *= p.act ? gelu_tanh_grad(r[j]) : gelu_erf_grad(r[j]);
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

}  // namespace fedb
