// Host runtime shared by all kernels: last-error string, device properties cache, TMA tensor-map
// encoding via cudaGetDriverEntryPoint (so the library links against cudart only).
#include "common.cuh"
#include "../include/fedb200.h"
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>
#include <mutex>

namespace fedb {

static thread_local char g_err[512] = "";

void set_last_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

static int g_pdl = -1;
int pdl_enabled() {
    if (g_pdl < 0) {
        const char* e = getenv("FEDB200_PDL");
        g_pdl = (e && e[0] == '0') ? 0 : 1;
    }
    return g_pdl;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

static int encode_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                               uint32_t box_rows, uint32_t box_cols);

// A tensor map is a pure function of (base, rows, cols, ld, box): the trainer issues ~120 GEMMs per step on a fixed set of
// buffers and only a handful of token counts, so the ~600 driver encodes per step are memoised.  Direct-mapped, 1024
// entries per host thread; a colliding key simply re-encodes.
struct TmapKey { const void* base; uint64_t rows, cols, ld; uint32_t br, bc; };
struct TmapEntry { TmapKey key; CUtensorMap map; bool valid; };
static thread_local TmapEntry g_tmap_cache[1024];
static unsigned long long g_tmap_hits = 0, g_tmap_misses = 0;

static inline size_t tmap_slot(const TmapKey& k) {
    uint64_t h = (uint64_t)(uintptr_t)k.base * 0x9E3779B97F4A7C15ull;
    h ^= (k.rows + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full;
    h ^= (k.cols << 20) ^ (k.ld << 40) ^ ((uint64_t)k.br << 8) ^ k.bc;
    h ^= h >> 29;
    return (size_t)(h * 0xBF58476D1CE4E5B9ull >> 54);          // top 10 bits
}

int make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                      uint32_t box_rows, uint32_t box_cols) {
    static const bool cache_on = [] { const char* e = getenv("FEDB200_TMAP_CACHE"); return !(e && e[0] == '0'); }();
    if (!cache_on) { ++g_tmap_misses; return encode_tmap_bf16_2d(out, base, rows, cols, ld_elems, box_rows, box_cols); }
    const TmapKey key = {base, rows, cols, ld_elems, box_rows, box_cols};
    TmapEntry& ent = g_tmap_cache[tmap_slot(key)];
    if (ent.valid && ent.key.base == base && ent.key.rows == rows && ent.key.cols == cols && ent.key.ld == ld_elems &&
        ent.key.br == box_rows && ent.key.bc == box_cols) {
        *out = ent.map;
        ++g_tmap_hits;
        return 0;
    }
    ++g_tmap_misses;
    int rc = encode_tmap_bf16_2d(out, base, rows, cols, ld_elems, box_rows, box_cols);
    if (rc == 0) { ent.key = key; ent.map = *out; ent.valid = true; }
    return rc;
}

static int encode_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                               uint32_t box_rows, uint32_t box_cols) {
    EncodeTiledFn fn = encode_fn();
    FEDB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
    FEDB_REQUIRE(((uintptr_t)base & 15) == 0 && (ld_elems * 2) % 16 == 0, "TMA needs 16-byte aligned base/stride");
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld_elems * 2};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FEDB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(2d) failed with CUresult %d (rows=%llu cols=%llu ld=%llu box=%ux%u)",
                 (int)r, (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld_elems, box_rows, box_cols);
    return 0;
}

int make_tmap_bf16_3d(CUtensorMap* out, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1_elems,
                      uint64_t s2_elems, uint32_t b0, uint32_t b1, uint32_t b2) {
    EncodeTiledFn fn = encode_fn();
    FEDB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
    cuuint64_t dims[3] = {d0, d1, d2};
    cuuint64_t strides[2] = {s1_elems * 2, s2_elems * 2};
    cuuint32_t box[3] = {b0, b1, b2};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FEDB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(3d) failed with CUresult %d", (int)r);
    return 0;
}

unsigned long long tmap_cache_hits() { return g_tmap_hits; }
unsigned long long tmap_cache_misses() { return g_tmap_misses; }

}  // namespace fedb

extern "C" const char* fedb200_last_error(void) { return fedb::g_err; }
extern "C" int fedb200_abi_version(void) { return FEDB200_ABI_VERSION; }
extern "C" int fedb200_sm_count(void) { return fedb::sm_count(); }
extern "C" int fedb200_set_pdl(int on) { fedb::g_pdl = on ? 1 : 0; return 0; }
extern "C" int fedb200_tmap_cache_stats(unsigned long long* hits, unsigned long long* misses) {
    if (hits) *hits = fedb::tmap_cache_hits();
    if (misses) *misses = fedb::tmap_cache_misses();
    return 0;
}
