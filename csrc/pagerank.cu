// On-device PageRank power iteration + mean ± k·sigma threshold -> anomaly mask (K5).
// Same algorithm as nx.pagerank (alpha 0.85, L1 tolerance N*tol, <= max_iter sweeps, out-weight row
// normalisation, uniform teleport and dangling redistribution) followed by the reference's thresholding
// (All_graphs_IMDB_dataset.ipynb:168-176).  mode 0: `mat` is the edge-weight matrix w[u][v];
// mode 1: `mat` is a Gram matrix of client updates -> (1 + cos)/2 * norm ratio, no self loops (the on-line
// scorer).  N <= 128 nodes: one CTA, the matrix lives in shared memory, arithmetic in fp64 so the result
// matches the CPU oracle to ~1e-12.  Latency-bound by construction.
#include "common.cuh"
#include "../include/fedb200.h"

namespace fedb {

constexpr int PR_MAX_N = 128;

__global__ void __launch_bounds__(PR_MAX_N) pagerank_kernel(const float* __restrict__ mat, int n, int mode, double alpha,
                                                            double tol, int max_iter, double k_sigma, int low_only,
                                                            double* __restrict__ ranks, int* __restrict__ mask,
                                                            double* __restrict__ info) {
    extern __shared__ double sm[];
    double* P = sm;                       // [n][n]
    double* x = P + n * n;                // [n]
    double* xn = x + n;                   // [n]
    double* red = xn + n;                 // [PR_MAX_N]
    __shared__ int dangling[PR_MAX_N];
    const int u = threadIdx.x;
    if (u < n) {
        double out = 0.0;
        for (int v = 0; v < n; ++v) {
            double w = (double)mat[u * n + v];
            if (mode == 1) {
                // w = (1 + cos) / 2 * (smaller norm / larger norm), no self loops (fedb200/anomaly/cosine.py)
                const double nu = fmax((double)mat[u * n + u], 1e-30), nv = fmax((double)mat[v * n + v], 1e-30);
                const double cs = fmin(fmax(w / sqrt(nu * nv), -1.0), 1.0);
                w = (u == v) ? 0.0 : 0.5 * (1.0 + cs) * sqrt(fmin(nu, nv) / fmax(nu, nv));
            }
            P[u * n + v] = w;
            out += w;
        }
        dangling[u] = out == 0.0;
        if (out > 0.0)
            for (int v = 0; v < n; ++v) P[u * n + v] /= out;
        x[u] = 1.0 / n;
    }
    __syncthreads();
    int it = 0;
    double err = 0.0;
    for (it = 1; it <= max_iter; ++it) {
        double mine = 0.0;
        if (u < n) {
            double s = 0.0, d = 0.0;
            for (int q = 0; q < n; ++q) {
                s += x[q] * P[q * n + u];
                if (dangling[q]) d += x[q];
            }
            xn[u] = alpha * (s + d / n) + (1.0 - alpha) / n;
            mine = fabs(xn[u] - x[u]);
        }
        red[u] = mine;
        __syncthreads();
        if (u == 0) {
            double e = 0.0;
            for (int q = 0; q < n; ++q) e += red[q];
            red[0] = e;
        }
        __syncthreads();
        err = red[0];
        if (u < n) x[u] = xn[u];
        __syncthreads();
        if (err < n * tol) break;
    }
    if (u == 0) {
        double mean = 0.0;
        for (int q = 0; q < n; ++q) mean += x[q];
        mean /= n;
        double var = 0.0;
        for (int q = 0; q < n; ++q) var += (x[q] - mean) * (x[q] - mean);
        const double sd = sqrt(var / n);
        info[0] = (double)(it > max_iter ? max_iter : it);
        info[1] = mean - k_sigma * sd;
        info[2] = mean + k_sigma * sd;
        info[3] = err;
        red[1] = info[1]; red[2] = info[2]; red[3] = sd;
    }
    __syncthreads();
    if (u < n) {
        ranks[u] = x[u];
        const bool bad = red[3] >= 1e-9 && (x[u] < red[1] || (!low_only && x[u] > red[2]));
        mask[u] = bad ? 1 : 0;
    }
}

}  // namespace fedb

using namespace fedb;

extern "C" int fedb200_pagerank(const float* mat, int n, int mode, double alpha, double tol, int max_iter,
                                double k_sigma, int low_only, double* ranks, int* mask, double* info, void* stream) {
    FEDB_REQUIRE(n >= 1 && n <= PR_MAX_N, "pagerank: n=%d outside [1, %d]", n, PR_MAX_N);
    const size_t smem = ((size_t)n * n + 2 * n + PR_MAX_N) * sizeof(double);
    FEDB_CHECK(cudaFuncSetAttribute(pagerank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    pagerank_kernel<<<1, PR_MAX_N, smem, (cudaStream_t)stream>>>(mat, n, mode, alpha, tol, max_iter, k_sigma, low_only,
                                                                 ranks, mask, info);
    FEDB_LAUNCH_CHECK();
    return 0;
}
