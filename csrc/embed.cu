// Embedding forward/backward (K11): three-table gather-add + LayerNorm + dropout in one pass, one warp per
// token.  Reference op: BertEmbeddings (word + position + token_type -> LayerNorm(1e-12) -> Dropout(0.1)),
// serverless_cancer_classification_with_BioBERT.ipynb:528-534.  Tables are read from the fp32 master arena.
// Backward recomputes the pre-LN sum from the tables (nothing but mean/rstd is saved), applies the LN
// backward and scatters into the three table gradients; the token-type rows (hit by every token) are
// reduced per block instead of by global atomics.
#include "common.cuh"
#include "../include/fedb200.h"

namespace fedb {

constexpr int EM_MAXV = 4;
constexpr int EM_WARPS = 8;

__device__ __forceinline__ void gather3(const float* w, const float* p, const float* t, int v, float (&o)[8]) {
    const float4 a0 = *reinterpret_cast<const float4*>(w + v * 8), a1 = *reinterpret_cast<const float4*>(w + v * 8 + 4);
    const float4 b0 = *reinterpret_cast<const float4*>(p + v * 8), b1 = *reinterpret_cast<const float4*>(p + v * 8 + 4);
    const float4 c0 = *reinterpret_cast<const float4*>(t + v * 8), c1 = *reinterpret_cast<const float4*>(t + v * 8 + 4);
    o[0] = a0.x + b0.x + c0.x; o[1] = a0.y + b0.y + c0.y; o[2] = a0.z + b0.z + c0.z; o[3] = a0.w + b0.w + c0.w;
    o[4] = a1.x + b1.x + c1.x; o[5] = a1.y + b1.y + c1.y; o[6] = a1.z + b1.z + c1.z; o[7] = a1.w + b1.w + c1.w;
}

__global__ void __launch_bounds__(EM_WARPS * 32) embed_fwd_kernel(
    const int* __restrict__ ids, const int* __restrict__ pos, const int* __restrict__ tt,
    const float* __restrict__ word, const float* __restrict__ posw, const float* __restrict__ typew,
    const float* __restrict__ gamma, const float* __restrict__ beta, __nv_bfloat16* __restrict__ out,
    float* __restrict__ mean_out, float* __restrict__ rstd_out, int T, int H, float eps, uint64_t seed,
    uint32_t keep_thresh, float inv_keep, int use_dropout) {
    pdl_sync();
    const int lane = threadIdx.x & 31, nvec = H / 8;
    const int warp = blockIdx.x * EM_WARPS + (threadIdx.x >> 5), nwarps = gridDim.x * EM_WARPS;
    for (int row = warp; row < T; row += nwarps) {
        const float* w = word + (size_t)ids[row] * H;
        const float* p = posw + (size_t)pos[row] * H;
        const float* t = typew + (size_t)tt[row] * H;
        float x[EM_MAXV][8];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < EM_MAXV; ++i) {
            const int v = lane + 32 * i;
            if (v < nvec) {
                gather3(w, p, t, v, x[i]);
#pragma unroll
                for (int k = 0; k < 8; ++k) s += x[i][k];
            }
        }
        const float mean = warp_sum(s) / (float)H;
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < EM_MAXV; ++i)
            if (lane + 32 * i < nvec) {
#pragma unroll
                for (int k = 0; k < 8; ++k) { const float d = x[i][k] - mean; q += d * d; }
            }
        const float rstd = rsqrtf(warp_sum(q) / (float)H + eps);
#pragma unroll
        for (int i = 0; i < EM_MAXV; ++i) {
            const int v = lane + 32 * i;
            if (v < nvec) {
                float o[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) o[k] = (x[i][k] - mean) * rstd * gamma[v * 8 + k] + beta[v * 8 + k];
                if (use_dropout) dropout_apply8(o, seed, (uint32_t)row, (uint32_t)(v * 8), keep_thresh, inv_keep);
                uint4 pk;
                pk.x = pack_bf16x2(o[0], o[1]); pk.y = pack_bf16x2(o[2], o[3]);
                pk.z = pack_bf16x2(o[4], o[5]); pk.w = pack_bf16x2(o[6], o[7]);
                *reinterpret_cast<uint4*>(out + (size_t)row * H + v * 8) = pk;
            }
        }
        if (lane == 0) { mean_out[row] = mean; rstd_out[row] = rstd; }
    }
}

__global__ void __launch_bounds__(EM_WARPS * 32) embed_bwd_kernel(
    const __nv_bfloat16* __restrict__ dx, const int* __restrict__ ids, const int* __restrict__ pos,
    const int* __restrict__ tt, const float* __restrict__ word, const float* __restrict__ posw,
    const float* __restrict__ typew, const float* __restrict__ gamma, const float* __restrict__ mean_in,
    const float* __restrict__ rstd_in, float* __restrict__ gword, float* __restrict__ gpos, float* __restrict__ gtype,
    float* __restrict__ dgamma, float* __restrict__ dbeta, int T, int H, uint64_t seed, uint32_t keep_thresh,
    float inv_keep, int use_dropout, float* __restrict__ dz_out, float* __restrict__ part) {
    // reproducible mode (dz_out / part given): the per-token gradient rows go to dz_out [T, H] and are scattered into the
    // tables by an ordered kernel afterwards; the four block-reduced quantities go to part [blocks][4][H]
    pdl_sync();
    extern __shared__ float red[];        // [EM_WARPS][H]
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, nvec = H / 8;
    const int warp = blockIdx.x * EM_WARPS + wib, nwarps = gridDim.x * EM_WARPS;
    float ag[EM_MAXV][8], ab[EM_MAXV][8], t0[EM_MAXV][8], t1[EM_MAXV][8];
#pragma unroll
    for (int i = 0; i < EM_MAXV; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k) ag[i][k] = ab[i][k] = t0[i][k] = t1[i][k] = 0.f;
    for (int row = warp; row < T; row += nwarps) {
        const int id = ids[row], ps = pos[row], ty = tt[row];
        const float* w = word + (size_t)id * H;
        const float* p = posw + (size_t)ps * H;
        const float* t = typew + (size_t)ty * H;
        const float mean = mean_in[row], rstd = rstd_in[row];
        float xh[EM_MAXV][8], dxh[EM_MAXV][8];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int i = 0; i < EM_MAXV; ++i) {
            const int v = lane + 32 * i;
            if (v < nvec) {
                float x[8];
                gather3(w, p, t, v, x);
                const uint4 raw = *reinterpret_cast<const uint4*>(dx + (size_t)row * H + v * 8);
                const float2 a = unpack_bf16x2(raw.x), b = unpack_bf16x2(raw.y), c = unpack_bf16x2(raw.z), d = unpack_bf16x2(raw.w);
                float g[8] = {a.x, a.y, b.x, b.y, c.x, c.y, d.x, d.y};
                if (use_dropout) dropout_apply8(g, seed, (uint32_t)row, (uint32_t)(v * 8), keep_thresh, inv_keep);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    xh[i][k] = (x[k] - mean) * rstd;
                    ag[i][k] += g[k] * xh[i][k];
                    ab[i][k] += g[k];
                    dxh[i][k] = g[k] * gamma[v * 8 + k];
                    s1 += dxh[i][k]; s2 += dxh[i][k] * xh[i][k];
                }
            }
        }
        const float c1 = warp_sum(s1) / (float)H, c2 = warp_sum(s2) / (float)H;
#pragma unroll
        for (int i = 0; i < EM_MAXV; ++i) {
            const int v = lane + 32 * i;
            if (v < nvec) {
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float dzv = rstd * (dxh[i][k] - c1 - xh[i][k] * c2);
                    if (dz_out != nullptr) {
                        dz_out[(size_t)row * H + v * 8 + k] = dzv;
                    } else {
                        atomicAdd(gword + (size_t)id * H + v * 8 + k, dzv);
                        atomicAdd(gpos + (size_t)ps * H + v * 8 + k, dzv);
                    }
                    if (ty == 0) t0[i][k] += dzv;
                    else if (ty == 1) t1[i][k] += dzv;
                    else if (dz_out == nullptr) atomicAdd(gtype + (size_t)ty * H + v * 8 + k, dzv);
                }
            }
        }
    }
    float* outs[4] = {dgamma, dbeta, gtype, gtype + H};
    for (int q = 0; q < 4; ++q) {
        __syncthreads();
#pragma unroll
        for (int i = 0; i < EM_MAXV; ++i) {
            const int v = lane + 32 * i;
            if (v < nvec) {
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    red[wib * H + v * 8 + k] = q == 0 ? ag[i][k] : (q == 1 ? ab[i][k] : (q == 2 ? t0[i][k] : t1[i][k]));
            }
        }
        __syncthreads();
        for (int c = threadIdx.x; c < H; c += EM_WARPS * 32) {
            float s = 0.f;
#pragma unroll
            for (int w2 = 0; w2 < EM_WARPS; ++w2) s += red[w2 * H + c];
            if (part != nullptr) part[((size_t)blockIdx.x * 4 + q) * H + c] = s;
            else if (s != 0.f) atomicAdd(outs[q] + c, s);
        }
    }
}

}  // namespace fedb

using namespace fedb;

static void drop_params(float p, uint32_t& thresh, float& inv_keep) {
    const float keep = 1.0f - p;
    thresh = dropout_thresh16(p);
    inv_keep = 1.0f / keep;
}

extern "C" int fedb200_embed_fwd(const int* ids, const int* pos, const int* tt, const float* word, const float* posw,
                                 const float* typew, const float* gamma, const float* beta, void* out, float* mean,
                                 float* rstd, int T, int H, float eps, unsigned long long seed, float dropout_p,
                                 void* stream) {
    FEDB_REQUIRE(H % 8 == 0 && H <= 1024, "embed: H=%d must be a multiple of 8 and <= 1024", H);
    if (T <= 0) return 0;
    uint32_t th; float ik;
    drop_params(dropout_p, th, ik);
    int blocks = (T + EM_WARPS - 1) / EM_WARPS;
    const int cap = sm_count() * 4;
    if (blocks > cap) blocks = cap;
    FEDB_CHECK(launch_pdl(embed_fwd_kernel, dim3(blocks), dim3(EM_WARPS * 32), 0, (cudaStream_t)stream, 
        ids, pos, tt, word, posw, typew, gamma, beta, (__nv_bfloat16*)out, mean, rstd, T, H, eps, seed, th, ik,
        dropout_p > 0.f));
    return 0;
}

extern "C" int fedb200_embed_bwd(const void* dx, const int* ids, const int* pos, const int* tt, const float* word,
                                 const float* posw, const float* typew, const float* gamma, const float* mean,
                                 const float* rstd, float* gword, float* gpos, float* gtype, float* dgamma,
                                 float* dbeta, int T, int H, int type_vocab, unsigned long long seed, float dropout_p,
                                 void* stream) {
    FEDB_REQUIRE(H % 8 == 0 && H <= 1024, "embed: H=%d must be a multiple of 8 and <= 1024", H);
    FEDB_REQUIRE(type_vocab >= 2, "embed_bwd: type_vocab must be >= 2");
    if (T <= 0) return 0;
    uint32_t th; float ik;
    drop_params(dropout_p, th, ik);
    int blocks = (T + EM_WARPS - 1) / EM_WARPS;
    const int cap = sm_count() * 2;
    if (blocks > cap) blocks = cap;
    float* dz = nullptr;
    float* part = nullptr;
    if (det_enabled()) {
        const size_t dz_bytes = (((size_t)T * H * sizeof(float)) + 255) & ~(size_t)255;
        dz = det_workspace(dz_bytes);
        part = det_workspace((size_t)blocks * 4 * H * sizeof(float), dz_bytes);
        FEDB_REQUIRE(dz != nullptr && part != nullptr, "embed_bwd: deterministic workspace too small");
    }
    FEDB_CHECK(launch_pdl(embed_bwd_kernel, dim3(blocks), dim3(EM_WARPS * 32), EM_WARPS * H * sizeof(float), (cudaStream_t)stream, 
        (const __nv_bfloat16*)dx, ids, pos, tt, word, posw, typew, gamma, mean, rstd, gword, gpos, gtype, dgamma,
        dbeta, T, H, seed, th, ik, dropout_p > 0.f, dz, part));
    if (dz != nullptr) {
        cudaStream_t s = (cudaStream_t)stream;
        FEDB_TRY(det_ordered_colsum(part, blocks, (size_t)4 * H, H, 4, dgamma, dbeta, gtype, gtype + H, s));
        FEDB_TRY(det_scatter_rows(ids, dz, T, H, 0, gword, s));
        FEDB_TRY(det_scatter_rows(pos, dz, T, H, 0, gpos, s));
        FEDB_TRY(det_scatter_rows(tt, dz, T, H, 2, gtype, s));            // token types 0 / 1 went through the block partials
    }
    return 0;
}
