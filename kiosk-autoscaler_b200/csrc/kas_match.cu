// kas_match.cu -- K1: prefix match and per-queue count over packed key heads,
// plus the host-side prefix-table builder.
//
// Takes over autoscaler.py:69-71 (one SCAN MATCH 'processing-<q>:*' and a
// Python-side count PER QUEUE) with ONE streaming pass over the snapshot for
// all queues.  HBM-bound integer/byte work: each key's rec_bytes are read
// exactly once with 128/256-bit loads, the queue table lives in shared memory,
// counts are aggregated warp -> block (shared) -> grid (one 64-bit atomic per
// queue per block).
#include "kas_common.cuh"
#include "kas_table.h"

#include <string.h>
#include <string>
#include <vector>

namespace kas {

constexpr int kMatchThreads = 256;
constexpr uint32_t kNoQueue = 0xFFFFFFFFu;

// ':' detector: 0x80 in every byte of w that equals 0x3A (exact, no carries).
__device__ __forceinline__ uint32_t colon_flags(uint32_t w) {
    uint32_t v = w ^ 0x3A3A3A3Au;
    uint32_t t = ((v & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | v;
    return ~t & 0x80808080u;
}
// 0x80-per-byte flags -> 4-bit mask (bit i <=> byte i).
__device__ __forceinline__ uint32_t flags_to_nibble(uint32_t f) {
    return (((f >> 7) * 0x00204081u) >> 21) & 0xFu;
}

template <int NCHUNK>
struct MatchCfg {
    static constexpr int kRecWords = 4 * NCHUNK;
    static constexpr int kRegionWords = (16 * NCHUNK - (int)kLiteralLen + 3) / 4;
    static constexpr int kKeysPerThread = NCHUNK <= 2 ? 4 : 2;
};

// Examine one record held in registers.  Returns the id of the first matching
// queue (kNoQueue if none); further matches of the same key (nested names
// such as "predict" and "predict:zstack") are added to the block histogram
// directly, they are rare.
template <int NCHUNK>
__device__ __forceinline__ uint32_t match_one(const uint32_t (&x)[4 * NCHUNK],
                                              const uint32_t *__restrict__ s_words,
                                              const uint32_t *__restrict__ s_qid,
                                              uint32_t *__restrict__ s_hist,
                                              uint32_t n_slots, uint32_t log2_slots,
                                              uint64_t len_mask) {
    constexpr int XW = MatchCfg<NCHUNK>::kRecWords;
    constexpr int RW = MatchCfg<NCHUNK>::kRegionWords;
    // "processing-" = 70 72 6f 63 | 65 73 73 69 | 6e 67 2d
    const bool literal = (x[0] == 0x636F7270u) & (x[1] == 0x69737365u) &
                         ((x[2] & 0x00FFFFFFu) == 0x002D676Eu);
    uint32_t first = kNoQueue;
    if (literal) {
        uint32_t r[RW];
#pragma unroll
        for (int k = 0; k < RW; ++k)
            r[k] = (3 + k < XW) ? __byte_perm(x[2 + k], x[3 + k], 0x6543) : (x[2 + k] >> 24);
        uint64_t cand = 0;
#pragma unroll
        for (int k = 0; k < RW; ++k)
            cand |= (uint64_t)flags_to_nibble(colon_flags(r[k])) << (4 * k);
        cand &= len_mask;
        while (cand) {
            const int L = __ffsll((long long)cand) - 1;
            cand &= cand - 1;
            const int nbytes = L + 1;
            uint32_t m[RW];
            uint32_t h = hash_init((uint32_t)nbytes);
#pragma unroll
            for (int k = 0; k < RW; ++k) {
                m[k] = r[k] & word_mask(k, nbytes);
                h = hash_step(h, m[k]);
            }
            uint32_t slot = hash_slot(h, log2_slots);
            for (;;) {
                const uint32_t w0 = s_words[slot];
                if (w0 == 0u) break;
                if (w0 == m[0]) {
                    bool eq = true;
#pragma unroll
                    for (int k = 1; k < RW; ++k) eq &= (s_words[k * n_slots + slot] == m[k]);
                    if (eq) {
                        const uint32_t q = s_qid[slot];
                        if (first == kNoQueue) first = q;
                        else atomicAdd(&s_hist[q], 1u);
                        break;
                    }
                }
                slot = (slot + 1) & (n_slots - 1);
            }
        }
    }
    return first;
}

template <int NCHUNK, bool WIDE>
__global__ void __launch_bounds__(kMatchThreads)
match_count_kernel(const uint8_t *__restrict__ heads, uint64_t n_keys,
                   const uint32_t *__restrict__ tbl, uint32_t n_prefix, uint32_t n_slots,
                   uint32_t log2_slots, unsigned long long *__restrict__ counts) {
    constexpr int XW = MatchCfg<NCHUNK>::kRecWords;
    constexpr int RW = MatchCfg<NCHUNK>::kRegionWords;
    constexpr int U = MatchCfg<NCHUNK>::kKeysPerThread;
    extern __shared__ uint32_t smem[];

    // Sizes come from the launch (they size the shared memory); a table image that
    // disagrees with them is reported by poisoning the counts, never by reading past it.
    if (tbl[0] != kTableMagic || tbl[1] != 16u * NCHUNK || tbl[2] != n_prefix || tbl[3] != n_slots) {
        if (blockIdx.x == 0)
            for (uint32_t i = threadIdx.x; i < n_prefix; i += blockDim.x) counts[i] = ~0ull;
        return;
    }
    const uint64_t len_mask = (uint64_t)tbl[4] | ((uint64_t)tbl[5] << 32);
    uint32_t *s_words = smem;                      // [RW][n_slots]
    uint32_t *s_qid = s_words + RW * n_slots;      // [n_slots]
    uint32_t *s_hist = s_qid + n_slots;            // [n_prefix]
    for (uint32_t i = threadIdx.x; i < (RW + 1) * n_slots; i += blockDim.x)
        smem[i] = tbl[kTableHeaderWords + i];
    for (uint32_t i = threadIdx.x; i < n_prefix; i += blockDim.x) s_hist[i] = 0u;
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t tile = (uint64_t)blockDim.x * U;
    for (uint64_t base = (uint64_t)blockIdx.x * tile; base < n_keys; base += (uint64_t)gridDim.x * tile) {
        uint32_t x[U][XW];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t idx = base + (uint64_t)u * blockDim.x + threadIdx.x;
            if (idx < n_keys) {
                const uint8_t *p = heads + idx * (uint64_t)(16 * NCHUNK);
                if constexpr (WIDE) {
                    static_assert(!WIDE || NCHUNK % 2 == 0, "256-bit loads need 32-byte records");
#pragma unroll
                    for (int c = 0; c < NCHUNK / 2; ++c) {
                        uint32_t v[8];
                        ldg_stream_256(p + 32 * c, v);
#pragma unroll
                        for (int j = 0; j < 8; ++j) x[u][8 * c + j] = v[j];
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < NCHUNK; ++c) {
                        // two 16-byte halves of one sector are fetched back to back by the
                        // same thread: let the first allocate in L1 so the second hits.
                        const uint4 v = (NCHUNK == 1) ? ldg_stream_128(p) : ldg_nc_128(p + 16 * c);
                        x[u][4 * c + 0] = v.x; x[u][4 * c + 1] = v.y;
                        x[u][4 * c + 2] = v.z; x[u][4 * c + 3] = v.w;
                    }
                }
            } else {
#pragma unroll
                for (int j = 0; j < XW; ++j) x[u][j] = 0u;
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t q = match_one<NCHUNK>(x[u], s_words, s_qid, s_hist, n_slots, log2_slots, len_mask);
            __syncwarp();
            if (__any_sync(0xFFFFFFFFu, q != kNoQueue)) {
                // warp histogram: lanes holding the same queue elect one adder
                const uint32_t peers = __match_any_sync(0xFFFFFFFFu, q);
                if (q != kNoQueue && lane == (uint32_t)(__ffs(peers) - 1))
                    atomicAdd(&s_hist[q], (uint32_t)__popc(peers));
            }
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n_prefix; i += blockDim.x) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&counts[i], (unsigned long long)c);
    }
}

static size_t match_smem_bytes(uint32_t n_prefix, uint32_t n_slots, uint32_t rec_bytes) {
    return ((size_t)(region_words(rec_bytes) + 1) * n_slots + n_prefix) * sizeof(uint32_t);
}

static uint32_t slots_for(uint32_t n_prefix) {
    uint32_t want = n_prefix <= 256 ? 4 * n_prefix : 2 * n_prefix;
    uint32_t s = 4;
    while (s < want) s <<= 1;
    return s;
}

template <int NCHUNK, bool WIDE>
static int launch_match(const void *d_heads, uint64_t n_keys, const void *d_tbl, uint32_t n_prefix,
                        uint64_t *d_counts, cudaStream_t stream) {
    const uint32_t rec_bytes = 16 * NCHUNK;
    const uint32_t n_slots = slots_for(n_prefix);
    uint32_t log2_slots = 0;
    while ((1u << log2_slots) < n_slots) ++log2_slots;
    const size_t smem = match_smem_bytes(n_prefix, n_slots, rec_bytes);
    auto kern = match_count_kernel<NCHUNK, WIDE>;
    if (smem > 48 * 1024)
        KAS_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    KAS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kMatchThreads, smem));
    if (per_sm < 1) return KAS_ERR_UNSUPPORTED;
    const int sms = sm_count();
    if (sms <= 0) return KAS_ERR_CUDA;
    const uint64_t tile = (uint64_t)kMatchThreads * MatchCfg<NCHUNK>::kKeysPerThread;
    uint64_t want = (n_keys + tile - 1) / tile;
    uint64_t grid = (uint64_t)sms * per_sm;
    if (want < grid) grid = want;
    if (grid == 0) grid = 1;
    kern<<<(unsigned)grid, kMatchThreads, smem, stream>>>(
        static_cast<const uint8_t *>(d_heads), n_keys, static_cast<const uint32_t *>(d_tbl),
        n_prefix, n_slots, log2_slots, reinterpret_cast<unsigned long long *>(d_counts));
    KAS_CUDA_TRY(cudaGetLastError());
    return KAS_OK;
}

static int match_count_impl(const void *d_heads, uint64_t n_keys, uint32_t rec_bytes,
                            const void *d_tbl, uint32_t n_prefix, uint64_t *d_counts,
                            void *stream_v, bool zero_first) {
    if (!d_tbl || !d_counts || (n_keys && !d_heads)) return KAS_ERR_INVALID;
    if (n_prefix == 0 || n_prefix > KAS_MAX_PREFIXES) return KAS_ERR_INVALID;
    if (rec_bytes == 0 || rec_bytes % 16 || rec_bytes > KAS_MAX_REC_BYTES) return KAS_ERR_UNSUPPORTED;
    if (!aligned(d_heads, 16) || !aligned(d_tbl, 4) || !aligned(d_counts, 8)) return KAS_ERR_ALIGNMENT;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (zero_first) KAS_CUDA_TRY(cudaMemsetAsync(d_counts, 0, sizeof(uint64_t) * n_prefix, stream));
    if (n_keys == 0) return KAS_OK;
    const bool wide = aligned(d_heads, 32);
    switch (rec_bytes / 16) {
        case 1: return launch_match<1, false>(d_heads, n_keys, d_tbl, n_prefix, d_counts, stream);
        case 2: return wide ? launch_match<2, true>(d_heads, n_keys, d_tbl, n_prefix, d_counts, stream)
                            : launch_match<2, false>(d_heads, n_keys, d_tbl, n_prefix, d_counts, stream);
        case 3: return launch_match<3, false>(d_heads, n_keys, d_tbl, n_prefix, d_counts, stream);
        case 4: return wide ? launch_match<4, true>(d_heads, n_keys, d_tbl, n_prefix, d_counts, stream)
                            : launch_match<4, false>(d_heads, n_keys, d_tbl, n_prefix, d_counts, stream);
    }
    return KAS_ERR_UNSUPPORTED;
}

}  // namespace kas

using namespace kas;

extern "C" {

uint32_t kas_rec_bytes_for(uint32_t max_name_len) {
    const uint32_t need = kLiteralLen + max_name_len + 1;
    return 16 * ((need + 15) / 16);
}

size_t kas_prefix_table_bytes(uint32_t n_prefix, uint32_t rec_bytes) {
    if (n_prefix == 0 || n_prefix > KAS_MAX_PREFIXES) return 0;
    if (rec_bytes == 0 || rec_bytes % 16 || rec_bytes > KAS_MAX_REC_BYTES) return 0;
    return (kTableHeaderWords + (size_t)(region_words(rec_bytes) + 1) * slots_for(n_prefix)) * sizeof(uint32_t);
}

int kas_prefix_table_build(const uint8_t *names, const uint32_t *name_off, uint32_t n_prefix,
                           uint32_t rec_bytes, void *host_tbl, size_t tbl_bytes) {
    if (!name_off || !host_tbl || (!names && n_prefix && name_off[n_prefix] != 0)) return KAS_ERR_INVALID;
    const size_t need = kas_prefix_table_bytes(n_prefix, rec_bytes);
    if (need == 0) return n_prefix == 0 || n_prefix > KAS_MAX_PREFIXES ? KAS_ERR_INVALID : KAS_ERR_UNSUPPORTED;
    if (tbl_bytes < need) return KAS_ERR_TOO_SMALL;
    const uint32_t RW = region_words(rec_bytes);
    const uint32_t n_slots = slots_for(n_prefix);
    uint32_t log2_slots = 0;
    while ((1u << log2_slots) < n_slots) ++log2_slots;
    uint32_t *t = static_cast<uint32_t *>(host_tbl);
    memset(t, 0, need);
    uint64_t len_mask = 0;
    uint32_t *words = t + kTableHeaderWords;
    uint32_t *qid = words + (size_t)RW * n_slots;
    for (uint32_t i = 0; i < n_slots; ++i) qid[i] = kNoQueue;
    std::vector<std::string> seen;
    for (uint32_t q = 0; q < n_prefix; ++q) {
        if (name_off[q + 1] < name_off[q]) return KAS_ERR_INVALID;
        const uint32_t len = name_off[q + 1] - name_off[q];
        if (len + kLiteralLen + 1 > rec_bytes) return KAS_ERR_UNSUPPORTED;
        const uint8_t *nm = names + name_off[q];
        for (uint32_t i = 0; i < len; ++i) {
            const uint8_t c = nm[i];
            if (c == '*' || c == '?' || c == '[' || c == '\\') return KAS_ERR_GLOB_META;
            if (c == 0) return KAS_ERR_UNSUPPORTED;
        }
        std::string s(reinterpret_cast<const char *>(nm), len);
        for (const auto &o : seen)
            if (o == s) return KAS_ERR_INVALID;  // redis_keys is a dict: names are unique
        seen.push_back(s);
        uint8_t buf[KAS_MAX_REC_BYTES + 4] = {0};
        memcpy(buf, nm, len);
        buf[len] = ':';
        const uint32_t nbytes = len + 1;
        uint32_t m[(KAS_MAX_REC_BYTES + 3) / 4];
        uint32_t h = hash_init(nbytes);
        for (uint32_t k = 0; k < RW; ++k) {
            m[k] = (uint32_t)buf[4 * k] | ((uint32_t)buf[4 * k + 1] << 8) |
                   ((uint32_t)buf[4 * k + 2] << 16) | ((uint32_t)buf[4 * k + 3] << 24);
            m[k] &= word_mask((int)k, (int)nbytes);
            h = hash_step(h, m[k]);
        }
        uint32_t slot = hash_slot(h, log2_slots);
        while (words[slot] != 0u) slot = (slot + 1) & (n_slots - 1);
        for (uint32_t k = 0; k < RW; ++k) words[(size_t)k * n_slots + slot] = m[k];
        qid[slot] = q;
        len_mask |= 1ull << len;
    }
    t[0] = kTableMagic;
    t[1] = rec_bytes;
    t[2] = n_prefix;
    t[3] = n_slots;
    t[4] = (uint32_t)len_mask;
    t[5] = (uint32_t)(len_mask >> 32);
    t[6] = RW;
    t[7] = log2_slots;
    return KAS_OK;
}

int kas_match_count(const void *d_heads, uint64_t n_keys, uint32_t rec_bytes,
                    const void *d_prefix_tbl, uint32_t n_prefix, uint64_t *d_counts, void *stream) {
    return match_count_impl(d_heads, n_keys, rec_bytes, d_prefix_tbl, n_prefix, d_counts, stream, true);
}

int kas_match_count_accumulate(const void *d_heads, uint64_t n_keys, uint32_t rec_bytes,
                               const void *d_prefix_tbl, uint32_t n_prefix, uint64_t *d_counts,
                               void *stream) {
    return match_count_impl(d_heads, n_keys, rec_bytes, d_prefix_tbl, n_prefix, d_counts, stream, false);
}

}  // extern "C"
