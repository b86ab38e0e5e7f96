// LayerNorm forward / backward over bf16 activations (K10) with the reference's eps 1e-12
// (BertLayerNorm, serverless_cancer_classification_with_BioBERT.ipynb:526-570).  One warp per row, rows held
// in registers as 16-byte vectors (H % 8 == 0, H <= 1024).  The backward also emits, for free, everything
// that is a column reduction of its own traffic: dgamma, dbeta and the bias gradient of the dense layer that
// produced the pre-LN tensor (optionally through that layer's dropout mask, regenerated from the RNG counter).
#include "common.cuh"
#include "../include/fedb200.h"

namespace fedb {

constexpr int LN_MAXV = 4;          // vectors of 8 bf16 per lane -> H <= 1024 (kernels are templated on the count actually needed)
constexpr int LN_WARPS = 8;

template <int NV>
struct RowVec {
    float x[NV][8];
};

template <int NV>
__device__ __forceinline__ void load_row(const __nv_bfloat16* row, int nvec, int lane, RowVec<NV>& r) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const int v = lane + 32 * i;
        if (v < nvec) {
            const uint4 raw = *reinterpret_cast<const uint4*>(row + v * 8);
            const float2 a = unpack_bf16x2(raw.x), b = unpack_bf16x2(raw.y), c = unpack_bf16x2(raw.z), d = unpack_bf16x2(raw.w);
            r.x[i][0] = a.x; r.x[i][1] = a.y; r.x[i][2] = b.x; r.x[i][3] = b.y;
            r.x[i][4] = c.x; r.x[i][5] = c.y; r.x[i][6] = d.x; r.x[i][7] = d.y;
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) r.x[i][k] = 0.f;
        }
    }
}
template <int NV>
struct RawRow {
    uint4 v[NV];
};
template <int NV>
__device__ __forceinline__ void load_raw(const __nv_bfloat16* row, int nvec, int lane, RawRow<NV>& r) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const int v = lane + 32 * i;
        r.v[i] = v < nvec ? *reinterpret_cast<const uint4*>(row + v * 8) : make_uint4(0u, 0u, 0u, 0u);
    }
}
template <int NV>
__device__ __forceinline__ void unpack_row(const RawRow<NV>& raw, RowVec<NV>& r) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const float2 a = unpack_bf16x2(raw.v[i].x), b = unpack_bf16x2(raw.v[i].y), c = unpack_bf16x2(raw.v[i].z), d = unpack_bf16x2(raw.v[i].w);
        r.x[i][0] = a.x; r.x[i][1] = a.y; r.x[i][2] = b.x; r.x[i][3] = b.y;
        r.x[i][4] = c.x; r.x[i][5] = c.y; r.x[i][6] = d.x; r.x[i][7] = d.y;
    }
}
__device__ __forceinline__ void store_vec(__nv_bfloat16* row, int v, const float (&y)[8]) {
    uint4 o;
    o.x = pack_bf16x2(y[0], y[1]); o.y = pack_bf16x2(y[2], y[3]);
    o.z = pack_bf16x2(y[4], y[5]); o.w = pack_bf16x2(y[6], y[7]);
    *reinterpret_cast<uint4*>(row + v * 8) = o;
}
template <int NV>
__device__ __forceinline__ void row_stats(const RowVec<NV>& r, int nvec, int lane, int H, float eps, float& mean, float& rstd) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k) s += r.x[i][k];
    mean = warp_sum(s) / (float)H;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        if (lane + 32 * i < nvec) {
#pragma unroll
            for (int k = 0; k < 8; ++k) { const float d = r.x[i][k] - mean; q += d * d; }
        }
    }
    rstd = rsqrtf(warp_sum(q) / (float)H + eps);
}

template <int NV>
__global__ void __launch_bounds__(LN_WARPS * 32) ln_fwd_kernel(const __nv_bfloat16* __restrict__ z,
                                                               const float* __restrict__ gamma,
                                                               const float* __restrict__ beta,
                                                               __nv_bfloat16* __restrict__ y, float* __restrict__ mean_out,
                                                               float* __restrict__ rstd_out, int T, int H, float eps) {
    pdl_sync();
    const int lane = threadIdx.x & 31, nvec = H / 8;
    const int warp = blockIdx.x * LN_WARPS + (threadIdx.x >> 5), nwarps = gridDim.x * LN_WARPS;
    for (int row = warp; row < T; row += nwarps) {
        RowVec<NV> r;
        load_row(z + (size_t)row * H, nvec, lane, r);
        float mean, rstd;
        row_stats(r, nvec, lane, H, eps, mean, rstd);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int v = lane + 32 * i;
            if (v < nvec) {
                float o[8];
                const float4 g0 = *reinterpret_cast<const float4*>(gamma + v * 8), g1 = *reinterpret_cast<const float4*>(gamma + v * 8 + 4);
                const float4 b0 = *reinterpret_cast<const float4*>(beta + v * 8), b1 = *reinterpret_cast<const float4*>(beta + v * 8 + 4);
                const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
                const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
                for (int k = 0; k < 8; ++k) o[k] = (r.x[i][k] - mean) * rstd * gg[k] + bb[k];
                store_vec(y + (size_t)row * H, v, o);
            }
        }
        if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
    }
}

// dz = LN'(dy);   df = dz * dropmask (optional);   dgamma += sum dy*xhat;  dbeta += sum dy;  dbias += sum df
template <int NV>
__global__ void __launch_bounds__(LN_WARPS * 32, (NV <= 2) ? 2 : 1) ln_bwd_kernel(
    const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ z, const float* __restrict__ mean_in,
    const float* __restrict__ rstd_in, const float* __restrict__ gamma, __nv_bfloat16* __restrict__ dz,
    __nv_bfloat16* __restrict__ df, float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dbias,
    int T, int H, uint64_t seed, uint32_t keep_thresh, float inv_keep, float* __restrict__ part) {
    pdl_sync();
    extern __shared__ float red[];        // [LN_WARPS][H]
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, nvec = H / 8;
    const int warp = blockIdx.x * LN_WARPS + wib, nwarps = gridDim.x * LN_WARPS;
    float ag[NV][8], ab[NV][8], as[NV][8];
#pragma unroll
    for (int i = 0; i < NV; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k) ag[i][k] = ab[i][k] = as[i][k] = 0.f;
    float gg[NV][8];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const int v = lane + 32 * i;
#pragma unroll
        for (int k = 0; k < 8; ++k) gg[i][k] = v < nvec ? gamma[v * 8 + k] : 0.f;
    }
    // one row ahead: with ~200 registers per thread only eight warps fit on an SM, so the next row's loads are issued
    // before this row's math to keep enough bytes in flight
    RawRow<NV> xr, gr;
    float mean_n = 0.f, rstd_n = 0.f;
    if (warp < T) {
        load_raw(z + (size_t)warp * H, nvec, lane, xr);
        load_raw(dy + (size_t)warp * H, nvec, lane, gr);
        mean_n = mean_in[warp]; rstd_n = rstd_in[warp];
    }
    for (int row = warp; row < T; row += nwarps) {
        RowVec<NV> x, g;
        unpack_row(xr, x);
        unpack_row(gr, g);
        const float mean = mean_n, rstd = rstd_n;
        const int nrow = row + nwarps;
        if (nrow < T) {
            load_raw(z + (size_t)nrow * H, nvec, lane, xr);
            load_raw(dy + (size_t)nrow * H, nvec, lane, gr);
            mean_n = mean_in[nrow]; rstd_n = rstd_in[nrow];
        }
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            if (lane + 32 * i < nvec) {
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float xh = (x.x[i][k] - mean) * rstd;
                    const float dxh = g.x[i][k] * gg[i][k];
                    x.x[i][k] = xh;
                    ag[i][k] += g.x[i][k] * xh;
                    ab[i][k] += g.x[i][k];
                    g.x[i][k] = dxh;
                    s1 += dxh; s2 += dxh * xh;
                }
            }
        }
        const float c1 = warp_sum(s1) / (float)H, c2 = warp_sum(s2) / (float)H;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int v = lane + 32 * i;
            if (v < nvec) {
                float o[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) o[k] = rstd * (g.x[i][k] - c1 - x.x[i][k] * c2);
                store_vec(dz + (size_t)row * H, v, o);
                if (df != nullptr) {
                    dropout_apply8(o, seed, (uint32_t)row, (uint32_t)(v * 8), keep_thresh, inv_keep);
                    store_vec(df + (size_t)row * H, v, o);
                }
                if (dbias != nullptr) {
                    // accumulate the bf16-rounded value the downstream GEMMs will actually consume
#pragma unroll
                    for (int k = 0; k < 8; ++k) as[i][k] += __bfloat162float(__float2bfloat16_rn(o[k]));
                }
            }
        }
    }
    // block reduction of the three column accumulators, one quantity at a time
    float* outs[3] = {dgamma, dbeta, dbias};
    for (int q = 0; q < 3; ++q) {
        if (outs[q] == nullptr) continue;
        __syncthreads();
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int v = lane + 32 * i;
            if (v < nvec) {
#pragma unroll
                for (int k = 0; k < 8; ++k) red[wib * H + v * 8 + k] = q == 0 ? ag[i][k] : (q == 1 ? ab[i][k] : as[i][k]);
            }
        }
        __syncthreads();
        for (int c = threadIdx.x; c < H; c += LN_WARPS * 32) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < LN_WARPS; ++w) s += red[w * H + c];
            if (part != nullptr) part[((size_t)blockIdx.x * 3 + q) * H + c] = s;      // reproducible mode: summed in block order afterwards
            else atomicAdd(outs[q] + c, s);
        }
    }
}

}  // namespace fedb

using namespace fedb;

static int ln_grid(int T) {
    int blocks = (T + LN_WARPS - 1) / LN_WARPS;
    const int cap = sm_count() * 4;
    return blocks < 1 ? 1 : (blocks > cap ? cap : blocks);
}

extern "C" int fedb200_layernorm_fwd(const void* z, const float* gamma, const float* beta, void* y, float* mean,
                                     float* rstd, int T, int H, float eps, void* stream) {
    FEDB_REQUIRE(H % 8 == 0 && H <= 8 * 32 * LN_MAXV, "layernorm: H=%d must be a multiple of 8 and <= 1024", H);
    if (T <= 0) return 0;
    const int nv = (H / 8 + 31) / 32;
    cudaError_t le;
#define FEDB_LN_FWD(NVV) le = launch_pdl(ln_fwd_kernel<NVV>, dim3(ln_grid(T)), dim3(LN_WARPS * 32), 0, (cudaStream_t)stream, \
        (const __nv_bfloat16*)z, gamma, beta, (__nv_bfloat16*)y, mean, rstd, T, H, eps)
    if (nv == 1) FEDB_LN_FWD(1); else if (nv == 2) FEDB_LN_FWD(2); else if (nv == 3) FEDB_LN_FWD(3); else FEDB_LN_FWD(4);
#undef FEDB_LN_FWD
    FEDB_CHECK(le);
    return 0;
}

extern "C" int fedb200_layernorm_bwd(const void* dy, const void* z, const float* mean, const float* rstd,
                                     const float* gamma, void* dz, void* df, float* dgamma, float* dbeta,
                                     float* dbias, int T, int H, unsigned long long seed, float dropout_p,
                                     void* stream) {
    FEDB_REQUIRE(H % 8 == 0 && H <= 8 * 32 * LN_MAXV, "layernorm: H=%d must be a multiple of 8 and <= 1024", H);
    if (T <= 0) return 0;
    const float keep = 1.0f - dropout_p;
    const uint32_t thresh = dropout_thresh16(dropout_p);
    if (dropout_p <= 0.f) df = nullptr;
    int blocks = ln_grid(T);
    const int nv = (H / 8 + 31) / 32;
    const int cap = nv <= 2 ? sm_count() * 2 : sm_count();       // resident blocks (register-limited); one wave
    if (blocks > cap) blocks = cap;
    if (nv <= 2 && blocks < ln_grid(T)) blocks = ln_grid(T) < sm_count() * 4 ? ln_grid(T) : sm_count() * 4;   // two blocks per SM fit
    cudaError_t le;
#define FEDB_LN_BWD(NVV) le = launch_pdl(ln_bwd_kernel<NVV>, dim3(blocks), dim3(LN_WARPS * 32), LN_WARPS * H * sizeof(float), \
        (cudaStream_t)stream, (const __nv_bfloat16*)dy, (const __nv_bfloat16*)z, mean, rstd, gamma, (__nv_bfloat16*)dz, \
        (__nv_bfloat16*)df, dgamma, dbeta, dbias, T, H, seed, thresh, 1.0f / keep, part)
    float* part = nullptr;
    if (det_enabled()) {
        part = det_workspace((size_t)blocks * 3 * H * sizeof(float));
        FEDB_REQUIRE(part != nullptr, "layernorm_bwd: deterministic workspace too small");
    }
    if (nv == 1) FEDB_LN_BWD(1); else if (nv == 2) FEDB_LN_BWD(2); else if (nv == 3) FEDB_LN_BWD(3); else FEDB_LN_BWD(4);
#undef FEDB_LN_BWD
    FEDB_CHECK(le);
    if (part != nullptr) FEDB_TRY(det_ordered_colsum(part, blocks, (size_t)3 * H, H, 3, dgamma, dbeta, dbias, nullptr, (cudaStream_t)stream));
    return 0;
}
