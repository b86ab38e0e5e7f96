// Reproducible-reduction mode (`fedb200_set_deterministic`).
//
// By default every cross-CTA sum of the training step — split-K weight gradients, bias / LayerNorm gradient column sums,
// the embedding-table scatter — is an fp32 atomic add, so the order of the additions (and the last bits of the result)
// depends on CTA scheduling: two runs from the same seed end with different weights and different Merkle roots.  With a
// workspace registered here the same kernels write their per-CTA (or per-split) partials into the workspace with plain
// stores, and the kernels of this file add them up in one fixed order.  The trainer's result then depends only on the
// inputs, the seed and the launch geometry (SM count), which is what an auditor re-deriving a ledger entry needs.
// Cost: one extra small kernel per reduction and the partial round trip through L2; off by default.
#include "common.cuh"
#include "../include/fedb200.h"

namespace fedb {

static float* g_det_ws = nullptr;
static size_t g_det_bytes = 0;

bool det_enabled() { return g_det_ws != nullptr; }
size_t det_capacity() { return g_det_bytes; }
float* det_workspace(size_t need_bytes, size_t offset_bytes) {
    if (g_det_ws == nullptr || offset_bytes + need_bytes > g_det_bytes) return nullptr;
    return reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(g_det_ws) + offset_bytes);
}

// o_q[c] += sum_b part[b * stride + q * seg + c], b ascending: one thread per output column, so the order is fixed.
__global__ void __launch_bounds__(128) ordered_colsum_kernel(const float* __restrict__ part, int nparts, size_t stride, int seg,
                                                             int nseg, float* o0, float* o1, float* o2, float* o3) {
    pdl_sync();
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= seg * nseg) return;
    const int q = c / seg;
    float* o = q == 0 ? o0 : (q == 1 ? o1 : (q == 2 ? o2 : o3));
    if (o == nullptr) return;
    float s = 0.f;
#pragma unroll 8
    for (int b = 0; b < nparts; ++b) s += part[(size_t)b * stride + c];
    o[c - q * seg] += s;
}

// out[i] += sum_s ws[s * stride + i], s ascending (split-K partials of one weight gradient); n and stride multiples of 4
__global__ void __launch_bounds__(256) ordered_split_sum_kernel(const float* __restrict__ ws, int splits, size_t stride, size_t n4,
                                                                float* __restrict__ out) {
    pdl_sync();
    const size_t step = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += step) {
        float4 a = reinterpret_cast<const float4*>(ws)[i];
        for (int s = 1; s < splits; ++s) {
            const float4 b = reinterpret_cast<const float4*>(ws + (size_t)s * stride)[i];
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        float4 o = reinterpret_cast<float4*>(out)[i];
        o.x += a.x; o.y += a.y; o.z += a.z; o.w += a.w;
        reinterpret_cast<float4*>(out)[i] = o;
    }
}

// out[keys[t]] += dz[t] for every row t with keys[t] >= min_key, rows of one key added in ascending t.  One block per row:
// a block whose key already occurs in an earlier row retires; the first row of a key walks the rest of the key array in
// chunks of 256 (ballot -> eight 32-bit masks) and adds the matching rows one at a time, each thread owning H/256 columns.
__global__ void __launch_bounds__(256) scatter_rows_ordered_kernel(const int* __restrict__ keys, const float* __restrict__ dz, int T,
                                                                   int H, int min_key, float* __restrict__ out) {
    pdl_sync();
    __shared__ uint32_t masks[8];
    const int t = blockIdx.x;
    const int key = keys[t];
    if (key < min_key) return;                                    // block-uniform
    int found = 0;
    for (int j = threadIdx.x; j < t; j += 256) found |= (keys[j] == key) ? 1 : 0;
    if (__syncthreads_or(found)) return;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int base = t; base < T; base += 256) {
        const int j = base + (int)threadIdx.x;
        const bool m = j < T && keys[j] == key;
        const uint32_t b = __ballot_sync(0xffffffffu, m);
        if ((threadIdx.x & 31) == 0) masks[threadIdx.x >> 5] = b;
        __syncthreads();
#pragma unroll 1
        for (int w = 0; w < 8; ++w) {
            uint32_t mm = masks[w];
            while (mm) {
                const int bit = __ffs((int)mm) - 1;
                mm &= mm - 1;
                const float* r = dz + (size_t)(base + w * 32 + bit) * H;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int c = (int)threadIdx.x + 256 * i;
                    if (c < H) acc[i] += r[c];
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = (int)threadIdx.x + 256 * i;
        if (c < H) out[(size_t)key * H + c] += acc[i];
    }
}

int det_ordered_colsum(const float* part, int nparts, size_t stride, int seg, int nseg, float* o0, float* o1, float* o2, float* o3,
                       cudaStream_t s) {
    FEDB_REQUIRE(nseg >= 1 && nseg <= 4, "ordered_colsum: %d segments", nseg);
    const int cols = seg * nseg;
    FEDB_CHECK(launch_pdl(ordered_colsum_kernel, dim3((cols + 127) / 128), dim3(128), 0, s, part, nparts, stride, seg, nseg, o0, o1, o2, o3));
    return 0;
}

int det_ordered_split_sum(const float* ws, int splits, size_t stride, size_t n, float* out, cudaStream_t s) {
    FEDB_REQUIRE(n % 4 == 0 && stride % 4 == 0, "ordered_split_sum: n and stride must be multiples of 4");
    const size_t n4 = n / 4;
    size_t blocks = (n4 + 255) / 256;
    const size_t cap = (size_t)sm_count() * 8;
    if (blocks > cap) blocks = cap;
    FEDB_CHECK(launch_pdl(ordered_split_sum_kernel, dim3((unsigned)blocks), dim3(256), 0, s, ws, splits, stride, n4, out));
    return 0;
}

int det_scatter_rows(const int* keys, const float* dz, int T, int H, int min_key, float* out, cudaStream_t s) {
    FEDB_REQUIRE(H <= 1024, "scatter_rows: H=%d > 1024", H);
    if (T <= 0) return 0;
    FEDB_CHECK(launch_pdl(scatter_rows_ordered_kernel, dim3(T), dim3(256), 0, s, keys, dz, T, H, min_key, out));
    return 0;
}

}  // namespace fedb

// Register (or, with ws == NULL, drop) the workspace that switches the training kernels to ordered reductions.  The caller
// owns the memory and keeps it alive; it is used in stream order by whichever stream runs the training step.
extern "C" int fedb200_set_deterministic(void* ws, size_t bytes) {
    FEDB_REQUIRE(ws == nullptr || (((uintptr_t)ws & 255) == 0 && bytes >= (1u << 20)),
                 "set_deterministic: workspace must be 256-byte aligned and at least 1 MiB");
    fedb::g_det_ws = reinterpret_cast<float*>(ws);
    fedb::g_det_bytes = ws != nullptr ? bytes : 0;
    return 0;
}
extern "C" int fedb200_get_deterministic(void) { return fedb::det_enabled() ? 1 : 0; }
