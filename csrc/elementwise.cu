// Memory-bound kernels over the flat parameter arena:
//   adamw_flat   (K13)  one launch over D elements, HF-AdamW semantics, fused bf16 shadow write + zero_grad
//   mix_n        (K1 local stage / K2 / K3 combine) out = sum_j w_j * src_j, sources may be peer-mapped pointers
//   colsum_bf16         bias gradients: fp32 column sums of a bf16 [M, N] matrix, accumulated into the grad arena
//   cast / delta helpers
// Replaces: the python-loop transformers.AdamW (src/Servercase/server_IID_IMDB.py:109,117) and the host
// numpy averaging `sum(param)/len(param)` (src/Serverlesscase/serverless_IID_IMDB.py:269).
#include "common.cuh"
#include <stdlib.h>
#include "../include/fedb200.h"

namespace fedb {

static inline int stream_grid(size_t work_items, int threads, int per_sm = 8) {
    size_t blocks = (work_items + threads - 1) / threads;
    size_t cap = (size_t)sm_count() * per_sm;
    return (int)(blocks < cap ? (blocks ? blocks : 1) : cap);
}

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) adamw_flat_kernel(float* __restrict__ p, float* __restrict__ g,
                                                         float* __restrict__ m, float* __restrict__ v,
                                                         __nv_bfloat16* __restrict__ shadow, size_t n4, size_t n,
                                                         float step_size, float b1, float b2, float omb1, float omb2,
                                                         float eps, float lr_wd, float gscale, int zero_grad) {
    pdl_sync();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 pp = reinterpret_cast<float4*>(p)[i];
        float4 gg = reinterpret_cast<float4*>(g)[i];
        float4 mm = reinterpret_cast<float4*>(m)[i];
        float4 vv = reinterpret_cast<float4*>(v)[i];
        float* pe = &pp.x; float* ge = &gg.x; float* me = &mm.x; float* ve = &vv.x;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float gr = ge[k] * gscale;
            me[k] = b1 * me[k] + omb1 * gr;                        // omb = 1 - beta evaluated in double on the host, as
            ve[k] = b2 * ve[k] + omb2 * gr * gr;                   // HF's `value=1.0 - beta2` is (1.0f - 0.999f is 4.7e-5 off)
            pe[k] -= step_size * me[k] / (sqrtf(ve[k]) + eps);     // eps outside the bias correction (HF AdamW)
            pe[k] -= lr_wd * pe[k];
        }
        reinterpret_cast<float4*>(p)[i] = pp;
        reinterpret_cast<float4*>(m)[i] = mm;
        reinterpret_cast<float4*>(v)[i] = vv;
        if (zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (shadow) {
            uint2 s;
            s.x = pack_bf16x2(pp.x, pp.y);
            s.y = pack_bf16x2(pp.z, pp.w);
            reinterpret_cast<uint2*>(shadow)[i] = s;
        }
    }
    // scalar tail (n not a multiple of 4)
    for (size_t i = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float gr = g[i] * gscale;
        float mm = b1 * m[i] + omb1 * gr;
        float vv = b2 * v[i] + omb2 * gr * gr;
        float pp = p[i] - step_size * mm / (sqrtf(vv) + eps);
        pp -= lr_wd * pp;
        p[i] = pp; m[i] = mm; v[i] = vv;
        if (zero_grad) g[i] = 0.f;
        if (shadow) shadow[i] = __float2bfloat16_rn(pp);
    }
}

// ---------------------------------------------------------------------------------------------
struct MixArgs {
    const float* src[FEDB200_MAX_MIX_SOURCES];
    float w[FEDB200_MAX_MIX_SOURCES];
    int n_src;
};

template <int UNROLL>
__global__ void __launch_bounds__(256) mix_n_kernel(MixArgs a, float* __restrict__ out,
                                                    __nv_bfloat16* __restrict__ shadow, size_t n4, size_t n) {
    pdl_sync();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        int j = 0;
        for (; j + UNROLL <= a.n_src; j += UNROLL) {
            float4 t[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) t[u] = __ldcs(reinterpret_cast<const float4*>(a.src[j + u]) + i);
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const float w = a.w[j + u];
                acc.x += w * t[u].x; acc.y += w * t[u].y; acc.z += w * t[u].z; acc.w += w * t[u].w;
            }
        }
        for (; j < a.n_src; ++j) {
            const float4 t = __ldcs(reinterpret_cast<const float4*>(a.src[j]) + i);
            const float w = a.w[j];
            acc.x += w * t.x; acc.y += w * t.y; acc.z += w * t.z; acc.w += w * t.w;
        }
        reinterpret_cast<float4*>(out)[i] = acc;
        if (shadow) {
            uint2 s;
            s.x = pack_bf16x2(acc.x, acc.y);
            s.y = pack_bf16x2(acc.z, acc.w);
            reinterpret_cast<uint2*>(shadow)[i] = s;
        }
    }
    for (size_t i = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float acc = 0.f;
        for (int j = 0; j < a.n_src; ++j) acc += a.w[j] * a.src[j][i];
        out[i] = acc;
        if (shadow) shadow[i] = __float2bfloat16_rn(acc);
    }
}

// ---------------------------------------------------------------------------------------------
// out[n] += sum_m x[m, n]   (x bf16 row-major with leading dimension ld).  blockDim = (32, 8):
// a warp-row covers 64 contiguous columns (bf16x2 per lane); rows are strided over y and blockIdx.y.
__global__ void __launch_bounds__(256) colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out,
                                                          int M, int N, int ld, int rows_per_block, float* __restrict__ part) {
    pdl_sync();
    __shared__ float2 red[8][32];
    const int col = (blockIdx.x * 32 + threadIdx.x) * 2;
    const int r0 = blockIdx.y * rows_per_block;
    const int r1 = min(M, r0 + rows_per_block);
    float2 acc = make_float2(0.f, 0.f);
    if (col < N) {
        for (int r = r0 + threadIdx.y; r < r1; r += 8) {
            const uint32_t raw = *reinterpret_cast<const uint32_t*>(x + (size_t)r * ld + col);
            const float2 f = unpack_bf16x2(raw);
            acc.x += f.x; acc.y += f.y;
        }
    }
    red[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && col < N) {
#pragma unroll
        for (int y = 1; y < 8; ++y) { acc.x += red[y][threadIdx.x].x; acc.y += red[y][threadIdx.x].y; }
        if (part != nullptr) {                       // reproducible mode: row-block partials, summed in order afterwards
            part[(size_t)blockIdx.y * N + col] = acc.x;
            part[(size_t)blockIdx.y * N + col + 1] = acc.y;
        } else {
            atomicAdd(out + col, acc.x);
            atomicAdd(out + col + 1, acc.y);
        }
    }
}

__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                                            size_t n4, size_t n) {
    pdl_sync();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 t = reinterpret_cast<const float4*>(in)[i];
        uint2 s;
        s.x = pack_bf16x2(t.x, t.y);
        s.y = pack_bf16x2(t.z, t.w);
        reinterpret_cast<uint2*>(out)[i] = s;
    }
    for (size_t i = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = __float2bfloat16_rn(in[i]);
}

// delta_bf16 = a - b (fp32 in), the operand of the cosine-Gram scorer
__global__ void __launch_bounds__(256) delta_bf16_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                         __nv_bfloat16* __restrict__ out, size_t n4, size_t n) {
    pdl_sync();
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 x = __ldcs(reinterpret_cast<const float4*>(a) + i);
        const float4 y = __ldcs(reinterpret_cast<const float4*>(b) + i);
        uint2 s;
        s.x = pack_bf16x2(x.x - y.x, x.y - y.y);
        s.y = pack_bf16x2(x.z - y.z, x.w - y.w);
        reinterpret_cast<uint2*>(out)[i] = s;
    }
    for (size_t i = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = __float2bfloat16_rn(a[i] - b[i]);
}

}  // namespace fedb

using namespace fedb;

// 1 - beta as HF computes it: `value=1.0 - beta2` in Python doubles on the DECIMAL the user wrote (0.999 -> 0.001).  The ABI
// carries beta as a float (0.999f = 0.99900001287...), so the decimal is recovered through its shortest 7-digit form first;
// 1.0f - 0.999f, or 1.0 - (double)0.999f, would both be 1.3e-5 (relative) away from the reference's second-moment weight.
static float one_minus(float beta) {
    char buf[32];
    snprintf(buf, sizeof(buf), "%.7g", (double)beta);
    return (float)(1.0 - strtod(buf, nullptr));
}

extern "C" int fedb200_adamw_flat(float* p, float* g, float* m, float* v, void* shadow_bf16, size_t n, int step,
                                  float lr, float beta1, float beta2, float eps, float weight_decay,
                                  float grad_scale, int zero_grad, void* stream) {
    FEDB_REQUIRE(step >= 1, "adamw_flat: step is 1-based");
    FEDB_REQUIRE((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0 &&
                 ((uintptr_t)shadow_bf16 & 7) == 0, "adamw_flat: buffers must be 16-byte aligned");
    const double bc1 = 1.0 - pow((double)beta1, step), bc2 = 1.0 - pow((double)beta2, step);
    const float step_size = (float)((double)lr * sqrt(bc2) / bc1);
    const size_t n4 = n / 4;
    FEDB_CHECK(launch_pdl(adamw_flat_kernel, dim3(stream_grid(n4 ? n4 : 1, 256, 8)), dim3(256), 0, (cudaStream_t)stream, 
        p, g, m, v, (__nv_bfloat16*)shadow_bf16, n4, n, step_size, beta1, beta2, one_minus(beta1),
        one_minus(beta2), eps, lr * weight_decay, grad_scale,
        zero_grad));
    return 0;
}

extern "C" int fedb200_mix_n(const float* const* srcs, const float* weights, int n_src, float* out,
                             void* shadow_bf16, size_t n, void* stream) {
    FEDB_REQUIRE(n_src >= 1 && n_src <= FEDB200_MAX_MIX_SOURCES, "mix_n: n_src %d outside [1, %d]", n_src,
                 FEDB200_MAX_MIX_SOURCES);
    MixArgs a;
    a.n_src = n_src;
    for (int j = 0; j < n_src; ++j) {
        FEDB_REQUIRE(((uintptr_t)srcs[j] & 15) == 0, "mix_n: source %d not 16-byte aligned", j);
        a.src[j] = srcs[j];
        a.w[j] = weights[j];
    }
    FEDB_REQUIRE(((uintptr_t)out & 15) == 0, "mix_n: out not 16-byte aligned");
    const size_t n4 = n / 4;
    FEDB_CHECK(launch_pdl(mix_n_kernel<4>, dim3(stream_grid(n4 ? n4 : 1, 256, 8)), dim3(256), 0, (cudaStream_t)stream, 
        a, out, (__nv_bfloat16*)shadow_bf16, n4, n));
    return 0;
}

extern "C" int fedb200_colsum_bf16(const void* x, float* out, int M, int N, int ld, void* stream) {
    FEDB_REQUIRE(N % 2 == 0 && ld % 2 == 0, "colsum: N and ld must be even");
    if (M <= 0 || N <= 0) return 0;
    const int col_blocks = (N + 63) / 64;
    int row_blocks = (sm_count() * 4 + col_blocks - 1) / col_blocks;
    if (row_blocks > (M + 63) / 64) row_blocks = (M + 63) / 64;
    if (row_blocks < 1) row_blocks = 1;
    const int rpb = (M + row_blocks - 1) / row_blocks;
    float* part = nullptr;
    if (det_enabled() && row_blocks > 1) {
        part = det_workspace((size_t)row_blocks * N * sizeof(float));
        FEDB_REQUIRE(part != nullptr, "colsum: deterministic workspace too small");
    }
    FEDB_CHECK(launch_pdl(colsum_bf16_kernel, dim3(dim3(col_blocks, row_blocks)), dim3(dim3(32, 8)), 0, (cudaStream_t)stream, 
        (const __nv_bfloat16*)x, out, M, N, ld, rpb, part));
    if (part != nullptr) FEDB_TRY(det_ordered_colsum(part, row_blocks, (size_t)N, N, 1, out, nullptr, nullptr, nullptr, (cudaStream_t)stream));
    return 0;
}

extern "C" int fedb200_cast_f32_bf16(const float* in, void* out, size_t n, void* stream) {
    const size_t n4 = (((uintptr_t)in & 15) == 0 && ((uintptr_t)out & 7) == 0) ? n / 4 : 0;
    FEDB_CHECK(launch_pdl(cast_f32_bf16_kernel, dim3(stream_grid(n4 ? n4 : n, 256, 8)), dim3(256), 0, (cudaStream_t)stream, 
        in, (__nv_bfloat16*)out, n4, n));
    return 0;
}

extern "C" int fedb200_delta_bf16(const float* a, const float* b, void* out, size_t n, void* stream) {
    const size_t n4 = ((((uintptr_t)a | (uintptr_t)b) & 15) == 0 && ((uintptr_t)out & 7) == 0) ? n / 4 : 0;
    FEDB_CHECK(launch_pdl(delta_bf16_kernel, dim3(stream_grid(n4 ? n4 : n, 256, 8)), dim3(256), 0, (cudaStream_t)stream, 
        a, b, (__nv_bfloat16*)out, n4, n));
    return 0;
}
