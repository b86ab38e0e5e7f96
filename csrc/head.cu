// Classification head (K12): pooler (first token -> dense -> tanh), dropout, classifier, softmax
// cross-entropy, and their backward.  B <= a few hundred rows, C in {2, 3, 40, 41}: latency-bound, so plain
// warp-per-output kernels reading the fp32 master weights.  Reference ops: BertPooler + Dropout(0.1) +
// Linear(768 -> C) + CrossEntropyLoss (serverless_cancer_classification_with_BioBERT.ipynb:563-569).
#include "common.cuh"
#include "../include/fedb200.h"

namespace fedb {

// pooled[b, o] = tanh(bp[o] + <Wp[o, :], x[first[b], :]>);  pooled_drop = pooled * dropmask
__global__ void __launch_bounds__(256) head_pool_kernel(const __nv_bfloat16* __restrict__ x, const int* __restrict__ first,
                                                        const float* __restrict__ Wp, const float* __restrict__ bp,
                                                        float* __restrict__ pooled, float* __restrict__ pooled_drop,
                                                        int H, uint64_t seed, uint32_t keep_thresh, float inv_keep,
                                                        int use_dropout) {
    pdl_sync();
    const int b = blockIdx.x, lane = threadIdx.x & 31;
    const int o = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (o >= H) return;
    const __nv_bfloat16* h0 = x + (size_t)first[b] * H;
    const float* w = Wp + (size_t)o * H;
    float s = 0.f;
    for (int i = lane * 2; i < H; i += 64) {
        const float2 hv = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(h0 + i));
        const float2 wv = *reinterpret_cast<const float2*>(w + i);
        s += hv.x * wv.x + hv.y * wv.y;
    }
    s = warp_sum(s);
    if (lane == 0) {
        const float t = tanhf(s + bp[o]);
        pooled[(size_t)b * H + o] = t;
        pooled_drop[(size_t)b * H + o] =
            use_dropout ? t * dropout_scale(seed, (uint32_t)b, (uint32_t)o, keep_thresh, inv_keep) : t;
    }
}

// one block per sample: logits, loss, prediction, dlogits = (softmax - onehot) * grad_scale
__global__ void __launch_bounds__(128) head_cls_kernel(const float* __restrict__ pd, const float* __restrict__ Wc,
                                                       const float* __restrict__ bc, const int* __restrict__ labels,
                                                       float* __restrict__ logits, float* __restrict__ loss,
                                                       int* __restrict__ pred, float* __restrict__ dlogits, int H, int C,
                                                       float grad_scale) {
    pdl_sync();
    extern __shared__ float sl[];      // [C]
    const int b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int c = warp; c < C; c += 4) {
        const float* w = Wc + (size_t)c * H;
        float s = 0.f;
        for (int i = lane; i < H; i += 32) s += w[i] * pd[(size_t)b * H + i];
        s = warp_sum(s);
        if (lane == 0) sl[c] = s + bc[c];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float mx = -INFINITY;
        int arg = 0;
        for (int c = 0; c < C; ++c) if (sl[c] > mx) { mx = sl[c]; arg = c; }
        float z = 0.f;
        for (int c = 0; c < C; ++c) z += expf(sl[c] - mx);
        const int y = labels ? labels[b] : -1;
        const float lse = mx + logf(z);
        for (int c = 0; c < C; ++c) {
            logits[(size_t)b * C + c] = sl[c];
            if (dlogits) dlogits[(size_t)b * C + c] = (expf(sl[c] - lse) - (c == y ? 1.f : 0.f)) * grad_scale;
        }
        if (loss) loss[b] = (y >= 0 && y < C) ? lse - sl[y] : 0.f;
        if (pred) pred[b] = arg;
    }
}

// thread per hidden column h: dWc[:, h], dpre[:, h];  block 0 also does dbc
__global__ void __launch_bounds__(128) head_bwd_cls_kernel(const float* __restrict__ dlogits, const float* __restrict__ pooled,
                                                           const float* __restrict__ pd, const float* __restrict__ Wc,
                                                           float* __restrict__ gWc, float* __restrict__ gbc,
                                                           float* __restrict__ dpre, int B, int H, int C, uint64_t seed,
                                                           uint32_t keep_thresh, float inv_keep, int use_dropout) {
    pdl_sync();
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h < H) {
        for (int c = 0; c < C; ++c) {
            float s = 0.f;
            for (int b = 0; b < B; ++b) s += dlogits[(size_t)b * C + c] * pd[(size_t)b * H + h];
            gWc[(size_t)c * H + h] += s;
        }
        for (int b = 0; b < B; ++b) {
            float s = 0.f;
            for (int c = 0; c < C; ++c) s += dlogits[(size_t)b * C + c] * Wc[(size_t)c * H + h];
            if (use_dropout) s *= dropout_scale(seed, (uint32_t)b, (uint32_t)h, keep_thresh, inv_keep);
            const float t = pooled[(size_t)b * H + h];
            dpre[(size_t)b * H + h] = s * (1.f - t * t);
        }
    }
    if (blockIdx.x == 0) {
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            float s = 0.f;
            for (int b = 0; b < B; ++b) s += dlogits[(size_t)b * C + c];
            gbc[c] += s;
        }
    }
}

// block per pooler output row o: gWp[o, :] += sum_b dpre[b, o] * h0[b, :];  gbp[o] += sum_b dpre[b, o]
__global__ void __launch_bounds__(256) head_bwd_pool_w_kernel(const float* __restrict__ dpre, const __nv_bfloat16* __restrict__ x,
                                                              const int* __restrict__ first, float* __restrict__ gWp,
                                                              float* __restrict__ gbp, int B, int H) {
    pdl_sync();
    extern __shared__ float sd[];      // dpre[:, o]
    const int o = blockIdx.x;
    for (int b = threadIdx.x; b < B; b += blockDim.x) sd[b] = dpre[(size_t)b * H + o];
    __syncthreads();
    for (int i = threadIdx.x; i < H; i += blockDim.x) {
        float s = 0.f;
        for (int b = 0; b < B; ++b) s += sd[b] * __bfloat162float(x[(size_t)first[b] * H + i]);
        gWp[(size_t)o * H + i] += s;
    }
    if (threadIdx.x == 0) {
        float s = 0.f;
        for (int b = 0; b < B; ++b) s += sd[b];
        gbp[o] += s;
    }
}

// dx[first[b], :] = sum_o dpre[b, o] * Wp[o, :].  Grid (sample, 64-column slab); the four 64-thread groups of a block split
// the reduction over o and combine through shared memory (one block per sample was a 250 us latency-bound tail).
__global__ void __launch_bounds__(256) head_bwd_pool_x_kernel(const float* __restrict__ dpre, const float* __restrict__ Wp,
                                                              const int* __restrict__ first, __nv_bfloat16* __restrict__ dx,
                                                              int H) {
    pdl_sync();
    extern __shared__ float sd[];      // dpre[b, :] followed by 4 x 64 partial sums
    float* part = sd + H;
    const int b = blockIdx.x, il = threadIdx.x & 63, grp = threadIdx.x >> 6;
    const int i = blockIdx.y * 64 + il;
    for (int o = threadIdx.x; o < H; o += blockDim.x) sd[o] = dpre[(size_t)b * H + o];
    __syncthreads();
    float s = 0.f;
    if (i < H) {
        const int per = (H + 3) / 4, o0 = grp * per, o1 = min(H, o0 + per);
#pragma unroll 8
        for (int o = o0; o < o1; ++o) s = fmaf(sd[o], __ldg(Wp + (size_t)o * H + i), s);
    }
    part[grp * 64 + il] = s;
    __syncthreads();
    if (grp == 0 && i < H)
        dx[(size_t)first[b] * H + i] = __float2bfloat16_rn(part[il] + part[64 + il] + part[128 + il] + part[192 + il]);
}

}  // namespace fedb

using namespace fedb;

extern "C" int fedb200_head_fwd(const void* x, const int* first_tok, const float* Wp, const float* bp,
                                const float* Wc, const float* bc, const int* labels, float* pooled,
                                float* pooled_drop, float* logits, float* loss, int* pred, float* dlogits, int B,
                                int H, int C, float grad_scale, unsigned long long seed, float dropout_p,
                                void* stream) {
    FEDB_REQUIRE(H % 2 == 0 && C >= 1 && C <= 4096, "head_fwd: bad H=%d / C=%d", H, C);
    if (B <= 0) return 0;
    const float keep = 1.0f - dropout_p;
    const uint32_t th = dropout_thresh16(dropout_p);
    FEDB_CHECK(launch_pdl(head_pool_kernel, dim3(dim3(B, (H + 7) / 8)), dim3(256), 0, (cudaStream_t)stream, 
        (const __nv_bfloat16*)x, first_tok, Wp, bp, pooled, pooled_drop, H, seed, th, 1.0f / keep, dropout_p > 0.f));
    FEDB_CHECK(launch_pdl(head_cls_kernel, dim3(B), dim3(128), C * sizeof(float), (cudaStream_t)stream, pooled_drop, Wc, bc, labels, logits, loss, pred,
                                                                       dlogits, H, C, grad_scale));
    return 0;
}

extern "C" int fedb200_head_bwd(const void* x, const int* first_tok, const float* Wp, const float* Wc,
                                const float* pooled, const float* pooled_drop, const float* dlogits, float* dpre,
                                float* gWp, float* gbp, float* gWc, float* gbc, void* dx, int B, int H, int C,
                                unsigned long long seed, float dropout_p, void* stream) {
    if (B <= 0) return 0;
    FEDB_REQUIRE(B <= 8192 && H <= 8192, "head_bwd: B/H too large for the shared staging");
    const float keep = 1.0f - dropout_p;
    const uint32_t th = dropout_thresh16(dropout_p);
    cudaStream_t s = (cudaStream_t)stream;
    FEDB_CHECK(launch_pdl(head_bwd_cls_kernel, dim3((H + 127) / 128), dim3(128), 0, s, dlogits, pooled, pooled_drop, Wc, gWc, gbc, dpre, B, H, C, seed,
                                                       th, 1.0f / keep, dropout_p > 0.f));
    FEDB_CHECK(launch_pdl(head_bwd_pool_w_kernel, dim3(H), dim3(256), B * sizeof(float), s, dpre, (const __nv_bfloat16*)x, first_tok, gWp, gbp, B, H));
    FEDB_CHECK(launch_pdl(head_bwd_pool_x_kernel, dim3(B, (H + 63) / 64), dim3(256), (H + 256) * sizeof(float), s, dpre, Wp, first_tok,
                          (__nv_bfloat16*)dx, H));
    return 0;
}
