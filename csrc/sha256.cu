// SHA-256 on the device (K6 `sha256_copy`, Merkle reduction, K7 `block_validate`).
// The reference has no ledger code at all (README.md:10 only names it); this is the device half of
// fedb200/ledger: every weight update is hashed while it is copied into the peer-visible buffer, so the
// ledger costs no extra pass over the 433 MB model.
//   * leaves: the span is cut into `chunk_bytes` chunks (multiple of 64 B); leaf_i = SHA256(chunk_i || 0x00), one thread
//     per chunk, 128 chunks per CTA staged through shared memory so global traffic stays coalesced
//     (4 lanes fetch one 64 B message block; the same registers feed the copy to `dst`).
//   * tree : SHA256(left || right || 0x01) per level (leaves are SHA256(chunk || 0x00): domain separation), unpaired node
//     promoted — same rule as ledger/merkle.py.
// Bit-exact against hashlib (tests/test_gpu_fed.py).
#include "common.cuh"
#include "../include/fedb200.h"

namespace fedb {

__constant__ uint32_t SHA_K[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98, 0x12835b01,
    0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc,
    0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147,
    0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08,
    0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208,
    0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};

__device__ __forceinline__ uint32_t rotr(uint32_t x, int n) { return __funnelshift_r(x, x, n); }
__device__ __forceinline__ uint32_t bswap(uint32_t x) { return __byte_perm(x, 0, 0x0123); }

__device__ __forceinline__ void sha_init(uint32_t (&s)[8]) {
    s[0] = 0x6a09e667; s[1] = 0xbb67ae85; s[2] = 0x3c6ef372; s[3] = 0xa54ff53a;
    s[4] = 0x510e527f; s[5] = 0x9b05688c; s[6] = 0x1f83d9ab; s[7] = 0x5be0cd19;
}

// one compression; w holds the 16 big-endian message words and is clobbered (rolling schedule)
__device__ __forceinline__ void sha_compress(uint32_t (&s)[8], uint32_t (&w)[16]) {
    uint32_t a = s[0], b = s[1], c = s[2], d = s[3], e = s[4], f = s[5], g = s[6], h = s[7];
#pragma unroll
    for (int i = 0; i < 64; ++i) {
        if (i >= 16) {
            const uint32_t w15 = w[(i + 1) & 15], w2 = w[(i + 14) & 15];
            const uint32_t s0 = rotr(w15, 7) ^ rotr(w15, 18) ^ (w15 >> 3);
            const uint32_t s1 = rotr(w2, 17) ^ rotr(w2, 19) ^ (w2 >> 10);
            w[i & 15] = w[i & 15] + s0 + w[(i + 9) & 15] + s1;
        }
        const uint32_t S1 = rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25);
        const uint32_t ch = (e & f) ^ (~e & g);
        const uint32_t t1 = h + S1 + ch + SHA_K[i] + w[i & 15];
        const uint32_t S0 = rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22);
        const uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
        const uint32_t t2 = S0 + mj;
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    s[0] += a; s[1] += b; s[2] += c; s[3] += d; s[4] += e; s[5] += f; s[6] += g; s[7] += h;
}

// final padding for a message of `len_bytes` whose last (len % 64) / 4 words are in tail[]
// Domain separation of the Merkle tree: a leaf is SHA256(chunk || 0x00), an interior node SHA256(left || right || 0x01), so a
// 64-byte interior pair can never be passed off as a 64-byte leaf (second-preimage splicing).  The tag is a SUFFIX byte:
// the data words keep their 64-byte block alignment, which is what lets the leaf hash run inside the copy kernel.
constexpr uint32_t PAD_LEAF = 0x00800000u;    // tag byte 0x00, then the 0x80 padding byte
constexpr uint32_t PAD_NODE = 0x01800000u;    // tag byte 0x01, then 0x80

__device__ __forceinline__ void sha_finish(uint32_t (&s)[8], uint32_t (&tail)[16], uint32_t tail_words, uint64_t len_bytes) {
    uint32_t w[16];
    len_bytes += 1;                            // the leaf tag byte
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = (uint32_t)i < tail_words ? tail[i] : ((uint32_t)i == tail_words ? PAD_LEAF : 0u);
    if (tail_words >= 14) {
        sha_compress(s, w);
#pragma unroll
        for (int i = 0; i < 16; ++i) w[i] = 0u;
    }
    const uint64_t bits = len_bytes * 8ull;
    w[14] = (uint32_t)(bits >> 32);
    w[15] = (uint32_t)bits;
    sha_compress(s, w);
}

constexpr int SHA_CHUNKS_PER_CTA = 128;

// BLK = 64-byte message blocks fetched per chunk and step.  BLK = 2 (chunk_bytes % 128 == 0) makes 8 neighbouring lanes
// load — and, for the fused send, store — 128 contiguous bytes: full cache lines over NVLink instead of 64-byte fragments,
// whose packet overhead capped the fused send well below the link rate.  BLK = 1 serves the other chunk sizes.
template <int BLK>
__global__ void __launch_bounds__(SHA_CHUNKS_PER_CTA) sha256_leaves_kernel(const uint32_t* __restrict__ src,
                                                                           uint32_t* __restrict__ dst, size_t nbytes,
                                                                           uint32_t chunk_bytes, uint32_t n_chunks,
                                                                           uint32_t* __restrict__ leaves) {
    constexpr int ROW = 16 * BLK + 1;                    // words per chunk row in the tile (+1: conflict-free column reads)
    constexpr int NV = 4 * BLK;                          // 16-byte vectors per thread and step
    constexpr int LPC = 4 * BLK;                         // lanes that cover one chunk's step bytes
    __shared__ uint32_t tile[2][SHA_CHUNKS_PER_CTA][ROW];
    const uint32_t t = threadIdx.x;
    const size_t chunk0 = (size_t)blockIdx.x * SHA_CHUNKS_PER_CTA;
    const size_t my_chunk = chunk0 + t;
    uint64_t my_len = 0;
    if (my_chunk < n_chunks) {
        const size_t off = my_chunk * (size_t)chunk_bytes;
        my_len = nbytes - off < chunk_bytes ? nbytes - off : chunk_bytes;
    }
    const uint32_t full_blocks = (uint32_t)(my_len / 64), rem_words = (uint32_t)((my_len % 64) / 4);
    const uint32_t steps = chunk_bytes / (64 * BLK);
    const size_t nwords = nbytes / 4;

    uint32_t st[8];
    sha_init(st);
    uint32_t tail[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) tail[i] = 0u;

    // Loads of step + 1 are issued before the compression rounds of step and consumed after them; the copy to `dst`
    // (peer memory over NVLink for the fused hash + send) is issued one iteration later, from the registers that are
    // already there.  Round 1 stored right behind each load, which made every thread wait for its load before hashing
    // (push and hash serialised: 1.22 ms = 0.60 hash + 0.56 push); now the stores are fire-and-forget work between
    // two compressions.
    uint4 pre[NV];
    auto word_of = [&](uint32_t step, int i) -> size_t {
        const uint32_t p = t + SHA_CHUNKS_PER_CTA * i;
        return ((chunk0 + p / LPC) * (size_t)chunk_bytes + (size_t)step * (64 * BLK) + (p % LPC) * 16) / 4;
    };
    auto fetch = [&](uint32_t step) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const size_t word = word_of(step, i);
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (word + 4 <= nwords) {
                v = __ldcs(reinterpret_cast<const uint4*>(src + word));
            } else if (word < nwords) {
                uint32_t* pv = &v.x;
                for (int k = 0; k < 4; ++k)
                    if (word + k < nwords) pv[k] = src[word + k];
            }
            pre[i] = v;
        }
    };
    auto push = [&](uint32_t step) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const size_t word = word_of(step, i);
            if (word + 4 <= nwords) {
                __stcs(reinterpret_cast<uint4*>(dst + word), pre[i]);
            } else if (word < nwords) {
                const uint32_t* pv = &pre[i].x;
                for (int k = 0; k < 4; ++k)
                    if (word + k < nwords) dst[word + k] = pv[k];
            }
        }
    };
    fetch(0);
    for (uint32_t step = 0; step < steps; ++step) {
        const int buf = step & 1;
        if (dst) push(step);
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const uint32_t p = t + SHA_CHUNKS_PER_CTA * i;
            uint32_t* row = &tile[buf][p / LPC][(p % LPC) * 4];
            row[0] = pre[i].x; row[1] = pre[i].y; row[2] = pre[i].z; row[3] = pre[i].w;
        }
        __syncthreads();
        if (step + 1 < steps) fetch(step + 1);
#pragma unroll
        for (int b = 0; b < BLK; ++b) {
            const uint32_t blk = step * BLK + b;
            if (blk < full_blocks) {
                uint32_t w[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) w[i] = bswap(tile[buf][t][b * 16 + i]);
                sha_compress(st, w);
            } else if (blk == full_blocks && rem_words != 0) {
#pragma unroll
                for (int i = 0; i < 16; ++i) tail[i] = bswap(tile[buf][t][b * 16 + i]);
            }
        }
    }
    if (my_chunk < n_chunks) {
        sha_finish(st, tail, rem_words, my_len);
#pragma unroll
        for (int i = 0; i < 8; ++i) leaves[my_chunk * 8 + i] = bswap(st[i]);
    }
}

// out[i] = SHA256(in[2i] || in[2i+1]);  an unpaired last node is promoted unchanged
__global__ void __launch_bounds__(128) merkle_level_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                           uint32_t n_in) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n_out = (n_in + 1) / 2;
    if (i >= n_out) return;
    if (2 * i + 1 >= n_in) {
#pragma unroll
        for (int k = 0; k < 8; ++k) out[i * 8 + k] = in[2 * i * 8 + k];
        return;
    }
    uint32_t w[16], st[8];
#pragma unroll
    for (int k = 0; k < 16; ++k) w[k] = bswap(in[2 * i * 8 + k]);
    sha_init(st);
    sha_compress(st, w);
#pragma unroll
    for (int k = 0; k < 16; ++k) w[k] = 0u;
    w[0] = PAD_NODE;
    w[15] = 520u;                              // (64 + 1 tag byte) * 8 bits
    sha_compress(st, w);
#pragma unroll
    for (int k = 0; k < 8; ++k) out[i * 8 + k] = bswap(st[k]);
}

// hash every 128-byte block header
__global__ void __launch_bounds__(128) header_hash_kernel(const uint32_t* __restrict__ headers, uint32_t* __restrict__ hashes,
                                                          uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t w[16], st[8];
    sha_init(st);
    for (int blk = 0; blk < 2; ++blk) {
#pragma unroll
        for (int k = 0; k < 16; ++k) w[k] = bswap(headers[(size_t)i * 32 + blk * 16 + k]);
        sha_compress(st, w);
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) w[k] = 0u;
    w[0] = 0x80000000u;
    w[15] = 1024u;
    sha_compress(st, w);
#pragma unroll
    for (int k = 0; k < 8; ++k) hashes[(size_t)i * 8 + k] = bswap(st[k]);
}

// chain rules: header.index == i, header.prev_hash == hash[i-1] (zeros for i == 0), and, when the caller
// supplies the recorded hashes, hash[i] == recorded[i].  first_bad = min offending index (n if none).
__global__ void __launch_bounds__(128) chain_check_kernel(const uint32_t* __restrict__ headers, const uint32_t* __restrict__ hashes,
                                                          const uint32_t* __restrict__ recorded, uint32_t n,
                                                          unsigned int* __restrict__ first_bad) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    bool ok = headers[(size_t)i * 32] == i;
    for (int k = 0; k < 8 && ok; ++k) {
        const uint32_t prev = i == 0 ? 0u : hashes[(size_t)(i - 1) * 8 + k];
        ok = headers[(size_t)i * 32 + 10 + k] == prev;           // prev_hash lives at byte offset 40
        if (ok && recorded) ok = recorded[(size_t)i * 8 + k] == hashes[(size_t)i * 8 + k];
    }
    if (!ok) atomicMin(first_bad, i);
}

}  // namespace fedb

using namespace fedb;

extern "C" int fedb200_sha256_leaves(const void* src, void* dst, size_t nbytes, unsigned int chunk_bytes,
                                     void* leaves, void* stream) {
    FEDB_REQUIRE(chunk_bytes >= 64 && chunk_bytes % 64 == 0, "sha256_leaves: chunk_bytes must be a multiple of 64");
    FEDB_REQUIRE(nbytes % 4 == 0 && nbytes > 0, "sha256_leaves: nbytes must be a positive multiple of 4");
    FEDB_REQUIRE(((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 15) == 0, "sha256_leaves: src/dst must be 16-byte aligned");
    const size_t n_chunks = (nbytes + chunk_bytes - 1) / chunk_bytes;
    FEDB_REQUIRE(n_chunks < (1ull << 31), "sha256_leaves: too many chunks");
    const unsigned grid = (unsigned)((n_chunks + SHA_CHUNKS_PER_CTA - 1) / SHA_CHUNKS_PER_CTA);
    if (chunk_bytes % 128 == 0)
        sha256_leaves_kernel<2><<<grid, SHA_CHUNKS_PER_CTA, 0, (cudaStream_t)stream>>>(
            (const uint32_t*)src, (uint32_t*)dst, nbytes, chunk_bytes, (uint32_t)n_chunks, (uint32_t*)leaves);
    else
        sha256_leaves_kernel<1><<<grid, SHA_CHUNKS_PER_CTA, 0, (cudaStream_t)stream>>>(
            (const uint32_t*)src, (uint32_t*)dst, nbytes, chunk_bytes, (uint32_t)n_chunks, (uint32_t*)leaves);
    FEDB_LAUNCH_CHECK();
    return 0;
}

// Reduces `n` 32-byte nodes in `nodes` to the root, ping-ponging with `scratch` (>= ceil(n/2) * 32 bytes);
// the 32-byte root is written to `root_out` (device memory).
extern "C" int fedb200_merkle_root(void* nodes, void* scratch, unsigned int n, void* root_out, void* stream) {
    FEDB_REQUIRE(n >= 1, "merkle_root: need at least one leaf");
    uint32_t* a = (uint32_t*)nodes;
    uint32_t* b = (uint32_t*)scratch;
    cudaStream_t s = (cudaStream_t)stream;
    while (n > 1) {
        const unsigned n_out = (n + 1) / 2;
        merkle_level_kernel<<<(n_out + 127) / 128, 128, 0, s>>>(a, b, n);
        FEDB_LAUNCH_CHECK();
        uint32_t* t = a; a = b; b = t;
        n = n_out;
    }
    FEDB_CHECK(cudaMemcpyAsync(root_out, a, 32, cudaMemcpyDeviceToDevice, s));
    return 0;
}

extern "C" int fedb200_block_validate(const void* headers, const void* recorded_hashes, unsigned int n,
                                      void* hashes_out, unsigned int* first_bad, void* stream) {
    if (n == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    FEDB_CHECK(cudaMemsetAsync(first_bad, 0xFF, sizeof(unsigned int), s));
    header_hash_kernel<<<(n + 127) / 128, 128, 0, s>>>((const uint32_t*)headers, (uint32_t*)hashes_out, n);
    FEDB_LAUNCH_CHECK();
    chain_check_kernel<<<(n + 127) / 128, 128, 0, s>>>((const uint32_t*)headers, (const uint32_t*)hashes_out,
                                                       (const uint32_t*)recorded_hashes, n, first_bad);
    FEDB_LAUNCH_CHECK();
    return 0;
}

// -------------------------------------------------------------------------------------------------
// Allocation-free Merkle root.
//
// fedb200_sha256_leaves + fedb200_merkle_root need two caller-supplied workspaces (the leaf digests and the
// ping-pong level buffer).  Some users of this library have no way to provide them: the asynchronous gossip
// receive path verifies an incoming update against the ledger while other peers' kernels are in flight, and
// a cudaMalloc/cudaFree there would synchronise the whole device and stall every peer; an auditing process
// that links the library only to validate a ledger (e.g. next to an inference server) is typically handed a
// fixed memory budget and no allocator; and a caller that sized its buffers before it knew the model size has
// nothing left to carve a workspace from.  For them the workspace lives in the library image itself: a
// file-scope __device__ array that each GPU holds from the moment the library is loaded until the process
// exits (FEDB200_STATIC_MERKLE_MAX_LEAVES * 48 bytes = 6 MiB).
//
// The array's name is only a host-side handle here; the device address is fetched once per device with
// cudaGetSymbolAddress (with that device current) and that pointer is what the kernels receive.
// -------------------------------------------------------------------------------------------------
namespace fedb {

constexpr size_t STATIC_LEAF_BYTES = (size_t)FEDB200_STATIC_MERKLE_MAX_LEAVES * 32;
constexpr size_t STATIC_LEVEL_BYTES = (size_t)(FEDB200_STATIC_MERKLE_MAX_LEAVES / 2 + 1) * 32;
constexpr size_t STATIC_SCRATCH_BYTES = STATIC_LEAF_BYTES + ((STATIC_LEVEL_BYTES + 255) / 256) * 256;

static __device__ __align__(256) unsigned char g_static_merkle_scratch[STATIC_SCRATCH_BYTES];

constexpr int STATIC_MAX_DEVICES = 64;
static unsigned char* g_static_scratch_addr[STATIC_MAX_DEVICES] = {nullptr};

// Device address of the scratch array on the current device (looked up on the first call per device).  The host-side
// table is filled without a lock: two host threads that race on a device's first call both obtain the same address from
// cudaGetSymbolAddress and store the same pointer value, so the race is benign (an aligned pointer store is not torn).
static int static_scratch_for_current_device(unsigned char** out) {
    int dev = -1;
    FEDB_CHECK(cudaGetDevice(&dev));
    FEDB_REQUIRE(dev >= 0 && dev < STATIC_MAX_DEVICES, "sha256_root_static: device ordinal %d not supported", dev);
    if (g_static_scratch_addr[dev] == nullptr) {
        void* p = nullptr;
        FEDB_CHECK(cudaGetSymbolAddress(&p, g_static_merkle_scratch));
        FEDB_REQUIRE(p != nullptr, "sha256_root_static: cudaGetSymbolAddress returned a null address");
        g_static_scratch_addr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_static_scratch_addr[dev];
    return 0;
}

}  // namespace fedb

// Merkle root (same chunking and tree rule as fedb200_sha256_leaves + fedb200_merkle_root) of `nbytes` at `src`,
// written as 32 bytes to `root_out`; both are device pointers owned by the caller.  Allocates nothing: the
// leaf digests and tree levels go to the library's static device array, so at most
// FEDB200_STATIC_MERKLE_MAX_LEAVES chunks are accepted (512 MiB at the default 4 KiB chunk — a whole fp32
// BERT-base arena); larger inputs are refused with an error, not truncated.
//   * The first call on each device resolves the array's device address (cudaGetSymbolAddress); make that
//     first call outside any stream capture.
//   * All calls on a device share the one array and it keeps whatever the previous call left there.  Nothing
//     in it has to start at zero (every digest that is read was written earlier in the same call), but two
//     calls that overlap on the GPU — different streams, or different host threads — race on it: serialise
//     them on one stream or with events.
extern "C" int fedb200_sha256_root_static(const void* src, size_t nbytes, unsigned int chunk_bytes, void* root_out,
                                          void* stream) {
    FEDB_REQUIRE(src != nullptr && root_out != nullptr, "sha256_root_static: null src/root_out");
    FEDB_REQUIRE(chunk_bytes >= 64 && chunk_bytes % 64 == 0, "sha256_root_static: chunk_bytes must be a multiple of 64");
    FEDB_REQUIRE(nbytes % 4 == 0 && nbytes > 0, "sha256_root_static: nbytes must be a positive multiple of 4");
    FEDB_REQUIRE(((uintptr_t)src & 15) == 0, "sha256_root_static: src must be 16-byte aligned");
    const size_t n_chunks = (nbytes + chunk_bytes - 1) / chunk_bytes;
    FEDB_REQUIRE(n_chunks <= (size_t)FEDB200_STATIC_MERKLE_MAX_LEAVES,
                 "sha256_root_static: %zu chunks of %u bytes exceed the static workspace (%d leaves max = %zu bytes of "
                 "input at this chunk size); use fedb200_sha256_leaves + fedb200_merkle_root with your own buffers",
                 n_chunks, chunk_bytes, FEDB200_STATIC_MERKLE_MAX_LEAVES,
                 (size_t)FEDB200_STATIC_MERKLE_MAX_LEAVES * chunk_bytes);
    unsigned char* scratch = nullptr;
    int e = fedb::static_scratch_for_current_device(&scratch);
    if (e) return e;
    uint32_t* a = reinterpret_cast<uint32_t*>(scratch);                               // leaves, then even levels
    uint32_t* b = reinterpret_cast<uint32_t*>(scratch + fedb::STATIC_LEAF_BYTES);     // odd levels
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned grid = (unsigned)((n_chunks + SHA_CHUNKS_PER_CTA - 1) / SHA_CHUNKS_PER_CTA);
    sha256_leaves_kernel<1><<<grid, SHA_CHUNKS_PER_CTA, 0, s>>>((const uint32_t*)src, nullptr, nbytes, chunk_bytes,
                                                            (uint32_t)n_chunks, a);
    FEDB_LAUNCH_CHECK();
    unsigned n = (unsigned)n_chunks;
    while (n > 1) {
        const unsigned n_out = (n + 1) / 2;
        merkle_level_kernel<<<(n_out + 127) / 128, 128, 0, s>>>(a, b, n);
        FEDB_LAUNCH_CHECK();
        uint32_t* t = a; a = b; b = t;
        n = n_out;
    }
    FEDB_CHECK(cudaMemcpyAsync(root_out, a, 32, cudaMemcpyDeviceToDevice, s));
    return 0;
}
