// Shared device/host helpers for the fedb200 sm_100a kernels.
//   * error plumbing for the C ABI (every entry point returns 0 or a cudaError_t value)
//   * thin inline-PTX wrappers: mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit / ld)
//   * UMMA shared-memory + instruction descriptor builders (bit layout as documented for sm_100)
//   * counter-based dropout RNG shared by forward and backward kernels
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>

namespace fedb {

// ---------------------------------------------------------------------------------------------
// error handling
// ---------------------------------------------------------------------------------------------
void set_last_error(const char* fmt, ...);

#define FEDB_CHECK(expr)                                                                   \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            fedb::set_last_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return (int)_e;                                                                \
        }                                                                                  \
    } while (0)

#define FEDB_REQUIRE(cond, ...)                                                            \
    do {                                                                                   \
        if (!(cond)) {                                                                     \
            fedb::set_last_error(__VA_ARGS__);                                             \
            return (int)cudaErrorInvalidValue;                                             \
        }                                                                                  \
    } while (0)

#define FEDB_LAUNCH_CHECK() FEDB_CHECK(cudaGetLastError())
#define FEDB_TRY(expr) do { int _rc = (expr); if (_rc) return _rc; } while (0)

int sm_count();   // cached multiprocessor count of the current device

// determinism.cu: reproducible-reduction mode.  When a workspace is registered (fedb200_set_deterministic) the training
// kernels store their per-CTA / per-split partials there and these helpers add them up in a fixed order.
bool det_enabled();
size_t det_capacity();
float* det_workspace(size_t need_bytes, size_t offset_bytes = 0);     // null when off or too small
int det_ordered_colsum(const float* part, int nparts, size_t stride, int seg, int nseg, float* o0, float* o1, float* o2, float* o3,
                       cudaStream_t s);
int det_ordered_split_sum(const float* ws, int splits, size_t stride, size_t n, float* out, cudaStream_t s);
int det_scatter_rows(const int* keys, const float* dz, int T, int H, int min_key, float* out, cudaStream_t s);

// ---------------------------------------------------------------------------------------------
// programmatic dependent launch (PDL)
//   A kernel launched through launch_pdl() may be scheduled while its predecessor on the stream is still
//   draining; it must execute pdl_wait() before its first global-memory access.  pdl_trigger() lets the NEXT kernel's
//   CTAs be scheduled as soon as every CTA of this one has reached it (they then park in their own pdl_wait()).
//   Both instructions are no-ops for a normal launch, so the same kernel works either way.
// ---------------------------------------------------------------------------------------------
int pdl_enabled();   // runtime.cu: FEDB200_PDL env (default on) / fedb200_set_pdl()

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() { pdl_wait(); pdl_trigger(); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                              Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------------------------------------
// small device utilities
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t v) {
    __nv_bfloat162 t = *reinterpret_cast<__nv_bfloat162*>(&v);
    return __bfloat1622float2(t);
}

// Counter-based dropout.  One 32-bit hash of (seed, row, col / 2) yields two 16-bit lanes, one per column of
// the aligned column pair; element (row, col) is kept iff its lane < thresh16 = round((1 - p) * 65536).
// Forward and backward kernels regenerate the same mask from the seed, so no mask tensor is ever stored.  A
// 32-bit mixer (two multiplies) keeps the cost at ~5 instructions per element even in kernels that walk a
// column (attention backward), where only one lane of each word is used.  fedb200/ops/rng.py is the host replica.
__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU;
    x ^= x >> 15; x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}
__device__ __forceinline__ uint32_t dropout_key(uint64_t seed) {
    return (uint32_t)seed ^ ((uint32_t)(seed >> 32) * 0xC2B2AE3DU);
}
__device__ __forceinline__ uint32_t dropout_word(uint32_t key, uint32_t row, uint32_t col2) {
    return hash32(key ^ (row * 0x9E3779B1U) ^ (col2 * 0x85EBCA77U));
}
__device__ __forceinline__ float dropout_lane(uint32_t word, uint32_t lane2, uint32_t thresh16, float inv_keep) {
    return ((word >> (16u * lane2)) & 0xFFFFu) < thresh16 ? inv_keep : 0.0f;
}
// scalar form: multiplier (0 or 1/(1-p)) for element (row, col); p == 0 callers skip the call entirely
__device__ __forceinline__ float dropout_scale(uint64_t seed, uint32_t row, uint32_t col, uint32_t thresh16,
                                               float inv_keep) {
    return dropout_lane(dropout_word(dropout_key(seed), row, col >> 1), col & 1u, thresh16, inv_keep);
}
// scales 8 consecutive values starting at column col0 (col0 % 8 == 0)
__device__ __forceinline__ void dropout_apply8(float* v, uint64_t seed, uint32_t row, uint32_t col0,
                                               uint32_t thresh16, float inv_keep) {
    const uint32_t key = dropout_key(seed);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t w = dropout_word(key, row, (col0 >> 1) + k);
        v[2 * k] *= dropout_lane(w, 0u, thresh16, inv_keep);
        v[2 * k + 1] *= dropout_lane(w, 1u, thresh16, inv_keep);
    }
}
// 32 keep decisions for columns [col0, col0 + 32) of `row` as one bit mask (bit i = column col0 + i is kept)
__device__ __forceinline__ uint32_t dropout_bits32(uint32_t key, uint32_t row, uint32_t col0, uint32_t thresh16) {
    uint32_t bits = 0u;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        const uint32_t w = dropout_word(key, row, (col0 >> 1) + k);
        bits |= ((w & 0xFFFFu) < thresh16 ? 1u : 0u) << (2 * k);
        bits |= ((w >> 16) < thresh16 ? 1u : 0u) << (2 * k + 1);
    }
    return bits;
}
// Attention-probability dropout (fedb200/ops/rng.py: attn_drop_base / attn_drop_word are the host replica).  The forward is
// instruction-issue bound and a two-multiply hash per column pair was a third of its instructions, so here ONE full hash
// covers a (row, 32-column chunk) and each of its 16 column pairs takes a Weyl step + xorshift-multiply-xorshift (measured
// on the host: keep rate 0.9000, |correlation| < 0.003 at every column lag up to 32, per-chunk keep counts binomial).
__device__ __forceinline__ uint32_t attn_drop_base(uint32_t key, uint32_t row, uint32_t chunk) {
    return hash32(key ^ (row * 0x9E3779B1U) ^ (chunk * 0x85EBCA77U));
}
__device__ __forceinline__ uint32_t attn_drop_word(uint32_t base, uint32_t pair) {
    uint32_t x = base + pair * 0x9E3779B9U;
    x ^= x >> 16; x *= 0x846CA68BU;
    x ^= x >> 15;
    return x;
}
__host__ __device__ inline uint32_t dropout_thresh16(float p) {
    const float keep = 1.0f - p;
    return keep >= 1.0f ? 65536u : (uint32_t)(keep * 65536.0f + 0.5f);
}

// erf-form GELU (BERT): gelu(x) = x * Phi(x).  The Gaussian CDF is evaluated as
//   Phi(x) = 0.5 + 0.5 * tanh(x * (c0 + c1 x^2 + c2 x^4)),   |x| clamped to 8,
// a least-squares fit of atanh(erf(x / sqrt 2)) (bench/fit_gelu.py): max |Phi error| 3.7e-5, max |gelu error| 3.0e-5 —
// two orders of magnitude below bf16 resolution, so the result rounds like the exact erf form.  With the hardware
// tanh (MUFU.TANH) this is 9 instructions; erff + expf cost ~3x more and made the FFN GEMM epilogues ALU-bound
// (profiles/).  NOT the "gelu_new" approximation (that one is gelu_tanh below, ALBERT's activation).
__device__ __forceinline__ float tanh_fast(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float gauss_cdf(float x) {
    const float xc = fminf(fmaxf(x, -8.0f), 8.0f);
    const float x2 = xc * xc;
    const float p = fmaf(x2, fmaf(x2, -0.0003438328493771962f, 0.03694645550401846f), 0.7975903260635063f);
    return fmaf(0.5f, tanh_fast(xc * p), 0.5f);
}
__device__ __forceinline__ float gelu_erf(float x) { return x * gauss_cdf(x); }
__device__ __forceinline__ float gelu_erf_grad(float x) {
    // Phi(x) + x * phi(x),  phi(x) = exp(-x^2 / 2) / sqrt(2 pi)
    return fmaf(x * 0.39894228040143268f, exp2f(-0.72134752044448170f * x * x), gauss_cdf(x));
}
__device__ __forceinline__ float gelu_tanh(float x) {
    const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
    return 0.5f * x * (1.0f + tanh_fast(u));          // MUFU.TANH: branch-free, error far below bf16 resolution
}
__device__ __forceinline__ float gelu_tanh_grad(float x) {
    const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
    const float t = tanh_fast(u);
    const float du = 0.7978845608028654f * (1.0f + 3.0f * 0.044715f * x * x);
    return 0.5f * (1.0f + t) + 0.5f * x * (1.0f - t * t) * du;
}

// ---------------------------------------------------------------------------------------------
// PTX: shared-state-space addresses, mbarrier, fences
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// Bounded spin: a pipeline bug shows up as a trapped kernel (launch error) instead of a hung GPU.
#ifndef FEDB_SPIN_LIMIT
#define FEDB_SPIN_LIMIT (1u << 24)
#endif
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > FEDB_SPIN_LIMIT) {
            printf("fedb200: mbarrier wait timed out (block %d thread %d smem 0x%x parity %u)\n", (int)blockIdx.x,
                   (int)threadIdx.x, smem_u32(bar), parity);
            __trap();
        }
    }
}
// generic-proxy writes (st.shared) -> visible to the async proxy (TMA / tcgen05 reads)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

// ---------------------------------------------------------------------------------------------
// PTX: TMA tiled loads (global -> shared, completion on an mbarrier)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
// TMA store (shared::cta -> global, bulk-group completion).  The issuing thread owns the bulk group: it commits and later
// waits (".read" = the shared-memory source has been read and may be overwritten; the global write may still be in flight,
// kernel completion makes it visible).
__device__ __forceinline__ void tma_store_smem_2d(const CUtensorMap* m, uint32_t smem_src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_src), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
// TMA load into a shared-memory address given as a 32-bit shared-window offset (own CTA)
__device__ __forceinline__ void tma_load_smem_2d(uint32_t smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

// ---------------------------------------------------------------------------------------------
// PTX: tcgen05 (TMEM allocation, UMMA issue, commit, TMEM loads)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result, uint32_t ncols) {   // one full warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(smem_result)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {        // same warp as alloc
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]; bf16 inputs, fp32 accumulate.  One thread issues for the CTA.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// same with the A operand in TMEM (lane = row, one 32-bit column = two consecutive k elements): D[tmem] (+)= A[tmem] * B[smem]
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on an mbarrier once every previously issued UMMA of this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                 ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// 32 lanes x 32 consecutive fp32 columns: thread t of the warp receives lane (base_lane + t)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}

__device__ __forceinline__ void tmem_ld_32x8(uint32_t taddr, uint32_t (&r)[8]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
        : "r"(taddr) : "memory");
}
// 32 lanes x 16 consecutive fp32 columns back into TMEM (thread t writes lane base_lane + t); pair with tmem_st_wait()
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
          "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_32x8(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------
// PTX: CTA pairs (cta_group::2) — two SMs of one cluster cooperate on one 256-row UMMA; the leader (cluster
// rank 0) issues the MMAs, both CTAs feed their half of the operands and drain their half of the accumulator.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared-window addresses carry the CTA rank in bit 24; clearing it names the same offset in the leader CTA
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint64_t* leader_bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(leader_bar) & kPeerBitMask), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_result, uint32_t ncols) {   // one full warp in EACH CTA
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(smem_result)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on the same-offset mbarrier of BOTH CTAs once the pair's previously issued UMMAs have completed
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    const uint16_t mask = 3;
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
// arrive on the mbarrier at the same offset in cluster CTA `cta`
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t cta) {
    asm volatile(
        "{\n\t.reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}"
        ::"r"(smem_u32(bar)), "r"(cta) : "memory");
}

// ---------------------------------------------------------------------------------------------
// UMMA descriptors (sm_100).  Shared-memory matrix descriptor, 64 bit:
//   [0,14)  start address >> 4         [16,30) leading-dim byte offset >> 4
//   [32,46) stride-dim byte offset >> 4  [46,48) version = 1      [49,52) base offset (0: 1024B-aligned atoms)
//   [61,64) layout: 0 none, 2 = 128B swizzle, 4 = 64B, 6 = 32B
// 128B-swizzle canonical layouts (T = 8 bf16 per 16 B):
//   K-major  : rows of 64 elements (128 B); 8-row groups are 1024 B apart  -> SBO = 1024, LBO unused
//   MN-major : 64 contiguous MN elements (128 B) x 8 k-rows per 1024 B atom; SBO = distance between
//              consecutive 8-k groups, LBO = distance between consecutive 64-element MN chunks
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
    d |= (uint64_t)1 << 46;          // descriptor version for Blackwell
    d |= (uint64_t)2 << 61;          // SWIZZLE_128B
    return d;
}
// Instruction descriptor for kind::f16 with bf16 A/B and fp32 D:
//   [4,6) D format (1 = f32)  [7,10) A format (1 = bf16)  [10,13) B format (1 = bf16)
//   [15] A major (0 K, 1 MN)  [16] B major  [17,23) N >> 3  [24,29) M >> 4
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int m, int n, int a_mn_major, int b_mn_major) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// ---------------------------------------------------------------------------------------------
// host: TMA tensor-map encoding through the driver entry point (no link-time libcuda dependency)
// ---------------------------------------------------------------------------------------------
// 2-D bf16 tensor [rows, cols] with `ld_elems` elements between rows; box = [box_rows, box_cols], 128B swizzle
int make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                      uint32_t box_rows, uint32_t box_cols);
// 3-D bf16 tensor [d2, d1, d0] (d0 contiguous) with element strides s1, s2; box [b2, b1, b0], 128B swizzle
int make_tmap_bf16_3d(CUtensorMap* out, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1_elems,
                      uint64_t s2_elems, uint32_t b0, uint32_t b1, uint32_t b2);

}  // namespace fedb
