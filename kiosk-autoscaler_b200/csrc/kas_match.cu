// kas_match.cu -- K1: prefix match and per-queue count over packed key heads,
// plus the host-side prefix-table builder.
//
// Takes over autoscaler.py:69-71 (one SCAN MATCH 'processing-<q>:*' and a
// Python-side count PER QUEUE) with ONE streaming pass over the snapshot for
// all queues.  HBM-bound integer/byte work: each key's rec_bytes are read
// exactly once with 128/256-bit loads, the queue table lives in shared memory,
// counts are aggregated lane -> block (shared-memory atomics) -> grid (one 64-bit
// atomic per queue per block).
//
// Per key: test the 11-byte literal, find the ':' bytes of the name region
// with byte-parallel arithmetic, and look the text up to the first ':' up in
// the table (hash of the masked region words, multiply-add chain).  Only when
// that entry says a longer queue name continues it does the walk go on to the
// next ':' (see kas_table.h).
#include "kas_common.cuh"
#include "kas_table.h"

#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <map>
#include <type_traits>
#include <string>
#include <vector>

namespace kas {

#ifndef KAS_MATCH_U
#define KAS_MATCH_U 2      // keys per thread per step (32-byte records); two steps are in flight
#endif
#ifndef KAS_MATCH_U1
#define KAS_MATCH_U1 KAS_MATCH_U   // the same for 16-byte records (half the bytes per record in flight)
#endif
#ifndef KAS_MATCH_MINB
#define KAS_MATCH_MINB 1   // resident blocks per SM the register budget is held to (narrow records)
#endif
#ifndef KAS_MATCH_THREADS
#define KAS_MATCH_THREADS 1024   // one big block per SM: the table is staged once per SM (narrow records)
#endif
constexpr uint32_t kNoQueue = 0xFFFFFFFFu;
constexpr size_t kMaxMatchSmem = 226 * 1024;   // of the 227 KB a block can have

// ':' detector, cheap form: 0x80 in every byte of w that equals 0x3A, and possibly
// in bytes ABOVE such a byte of the same word (borrow).  The lowest flag of a word
// is always a real colon, which is all the first step of a walk needs.  The integer
// ALU pipe is this kernel's limiter and the FMA pipe is nearly idle, so the subtract
// is issued there as v * one - 0x01010101 with `one` a kernel argument equal to 1
// (a literal 1 would be folded back into an ALU add).
__device__ __forceinline__ uint32_t colon_flags_lowest_exact(uint32_t w, uint32_t one) {
    const uint32_t v = w ^ 0x3A3A3A3Au;
    const uint32_t d = v * one + 0xFEFEFEFFu;
    return d & ~v & 0x80808080u;
}
// ':' detector: 0x80 in every byte of w that equals 0x3A (exact, no carries).
__device__ __forceinline__ uint32_t colon_flags(uint32_t w) {
    const uint32_t v = w ^ 0x3A3A3A3Au;
    const uint32_t t = ((v & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | v;
    return ~t & 0x80808080u;
}
// 0x80-per-byte flags -> the same four bits in the top nibble (no carries:
// every partial product lands on its own bit).
__device__ __forceinline__ uint32_t flags_to_top_nibble(uint32_t f) { return f * 0x00204081u; }

template <int NCHUNK>
struct MatchCfg {
    static constexpr int kRecWords = 4 * NCHUNK;
    static constexpr int kRW = (16 * NCHUNK - (int)kRegionStart) / 4;       // region words = record words 2..
    static constexpr int kRWP = (kRW + 3) / 4 * 4;                         // padded to 128-bit rows
    static constexpr int kEW = kRW + 2;                                    // entry words
    static constexpr int kRows = 16 * NCHUNK - (int)kRegionStart;           // colon positions inside the region
    static constexpr int kKeysPerThread = NCHUNK == 1 ? KAS_MATCH_U1 : (NCHUNK <= 2 ? KAS_MATCH_U : 1);
    static constexpr int kThreads = NCHUNK <= 2 ? KAS_MATCH_THREADS : 256;
    static constexpr int kMinBlocks = NCHUNK <= 2 ? KAS_MATCH_MINB : 3;    // wide records need the registers
    using mask_t = typename std::conditional<(kRows <= 32), uint32_t, uint64_t>::type;   // one bit per colon position
    // a deferred walk: region words + the position already dealt with, padded to 128-bit rows
    static constexpr int kItemWords = (kRW + 1 + 3) / 4 * 4;
    // Shared-memory copies of every mask row, one per lane of a quarter warp (the unit a 128-bit
    // shared load is served in): lanes that need different rows then never meet in a bank.
    static constexpr int kMaskCopies = NCHUNK <= 2 ? 8 : 1;
    static_assert(kEW % 4 == 0, "entry rows must be 128-bit multiples");
};
// per warp: at most 31 waiting + 32 pushed per record parked in one go.  JOINT: the two
// records of a step park together (one round of bookkeeping, 4 % faster with few queues);
// taken when the larger queue still fits shared memory beside the table.
constexpr int queue_items(bool joint) { return joint ? 32 + 32 * KAS_MATCH_U : 64; }

struct MatchSmem {
    const uint32_t *masks;    // [kRWP / 4][kRows][kMaskCopies] x 128 bits
    const uint32_t *entries;  // [kEW / 4][n_slots][entry copies] x 128 bits
    uint32_t *hist;           // [n_prefix]
    uint32_t *queue;          // this warp's [queue_items(JOINT)][kItemWords]
};

template <typename M> __device__ __forceinline__ int first_bit(M m);
template <> __device__ __forceinline__ int first_bit<uint32_t>(uint32_t m) { return __ffs((int)m) - 1; }
template <> __device__ __forceinline__ int first_bit<uint64_t>(uint64_t m) { return __ffsll((long long)m) - 1; }

template <int NCHUNK>
struct Matcher {
    using Cfg = MatchCfg<NCHUNK>;
    using mask_t = typename Cfg::mask_t;
    static constexpr int XW = Cfg::kRecWords, RW = Cfg::kRW, RWP = Cfg::kRWP, EW = Cfg::kEW,
                         IW = Cfg::kItemWords;
    MatchSmem sm;
    uint32_t seed, slot_shift, lane, lanemask_lt, one;
    // 32-bit shared addresses of this lane's copy of mask row 0 / entry slot 0 (piece 0), the
    // shift that turns a slot into a byte offset, and the byte distance between the pieces
    uint32_t mask_addr, entry_addr, slot_byte_shift, entry_piece_bytes;
    bool two_slot;
    mask_t len_mask;
    uint32_t waiting;   // deferred walks in this warp's queue (warp-uniform)

    // bit P of the result <=> region byte P is ':' and some entry ends there.  EXACT =
    // false uses the cheap detector, whose lowest set bit per word is still exact.
    template <bool EXACT>
    __device__ __forceinline__ mask_t colon_mask(const uint32_t (&r)[RW]) const {
        mask_t colons = 0;
        if constexpr (sizeof(mask_t) == 4) {
#pragma unroll
            for (int k = RW - 1; k >= 0; --k)
                colons = __funnelshift_l(
                    flags_to_top_nibble(EXACT ? colon_flags(r[k]) : colon_flags_lowest_exact(r[k], one)), colons, 4);
        } else {
#pragma unroll
            for (int k = RW - 1; k >= 0; --k)
                colons = (colons << 4) |
                         (mask_t)(flags_to_top_nibble(EXACT ? colon_flags(r[k]) : colon_flags_lowest_exact(r[k], one)) >> 28);
        }
        // Only positions where some entry ends can match.  Dropping the others up front
        // is safe because the table is closed under cutting at inner colons: a candidate
        // that spans a colon at a non-entry position is no entry.
        return colons & len_mask;
    }

    // Region words (record words 2.., i.e. "ng-" + name + ...) and the ':' positions
    // that can end an entry.  Returns false when the record does not start with
    // "processi"; the remaining "ng-" is part of every stored string.
    __device__ __forceinline__ bool prepare(const uint32_t (&x)[XW], uint32_t (&r)[RW], mask_t &colons) const {
        // "processi" = 70 72 6f 63 | 65 73 73 69
        const bool literal = (x[0] == 0x636F7270u) & (x[1] == 0x69737365u);
#pragma unroll
        for (int k = 0; k < RW; ++k) r[k] = x[2 + k];
        // cheap detector: the FIRST set bit is exact (a spurious bit can only sit above a
        // real colon of the same word); walks that go on recompute the exact mask
        colons = colon_mask<false>(r);
        return literal;
    }

    // Is region[0..L] an entry?  One look at its home slot (two for a two-slot
    // table), no probing (kas_table.h).  On a hit: the entry's queue id (kNoQueue for
    // a virtual entry) and whether a longer entry continues it.
    __device__ __forceinline__ bool lookup(const uint32_t (&r)[RW], int L, uint32_t &qid, bool &ext) const {
        uint32_t m[RWP];
        const uint32_t mrow = mask_addr + (uint32_t)L * (Cfg::kMaskCopies * 16);
#pragma unroll
        for (int j = 0; j < RWP / 4; ++j) {
            const uint4 v = lds_128(mrow + j * (Cfg::kRows * Cfg::kMaskCopies * 16));
            m[4 * j] = v.x; m[4 * j + 1] = v.y; m[4 * j + 2] = v.z; m[4 * j + 3] = v.w;
        }
        uint32_t h = 0;
#pragma unroll
        for (int k = 0; k < RW; ++k) {
            m[k] &= r[k];
            h += m[k] * hash_word_constant(k);
        }
        const uint32_t tag = hash_finish(h, (uint32_t)L + 1u, seed);
        uint32_t w[EW];
        auto fetch = [&](uint32_t slot) {
            const uint32_t e = entry_addr + (slot << slot_byte_shift);
#pragma unroll
            for (int j = 0; j < EW / 4; ++j) {
                const uint4 v = lds_128(e + (uint32_t)j * entry_piece_bytes);
                w[4 * j] = v.x; w[4 * j + 1] = v.y; w[4 * j + 2] = v.z; w[4 * j + 3] = v.w;
            }
            uint32_t diff = 0;
#pragma unroll
            for (int k = 0; k < RW; ++k) diff |= w[k] ^ m[k];
            return diff == 0u;
        };
        bool hit = fetch(hash_slot_a(tag, slot_shift));
        if (two_slot && !hit) hit = fetch(hash_slot_b(tag, slot_shift));
        ext = w[RW] != 0u;
        qid = w[RW + 1];
        return hit;
    }

    // Nested queue names ("predict" and "predict:zstack") make a key's walk go on
    // past its first colon.  Few lanes need that, so instead of looping in place
    // with most of the warp idle, they park the walk (region words + the position
    // already dealt with) in the warp's queue; the warp resumes 32 parked walks at
    // a time with every lane busy.
    __device__ __forceinline__ void park(bool more, const uint32_t (&r)[RW], int done_upto) {
        const uint32_t who = __ballot_sync(0xFFFFFFFFu, more);
        if (who == 0u) return;
        if (more) {
            uint32_t *it = sm.queue + (waiting + __popc(who & lanemask_lt)) * IW;
            uint32_t v[IW];
#pragma unroll
            for (int k = 0; k < IW; ++k) v[k] = 0u;
#pragma unroll
            for (int k = 0; k < RW; ++k) v[k] = r[k];
            v[RW] = (uint32_t)done_upto;
#pragma unroll
            for (int j = 0; j < IW / 4; ++j)
                reinterpret_cast<uint4 *>(it)[j] = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
        waiting += __popc(who);
        __syncwarp();
    }

    // park() for the two records of a step at once: one early exit, one update of `waiting`,
    // one warp barrier.  r0 / r1 are the records' region words (x[2..]).
    __device__ __forceinline__ void park_pair(bool more0, const uint32_t *r0, int L0,
                                              bool more1, const uint32_t *r1, int L1) {
        const uint32_t who0 = __ballot_sync(0xFFFFFFFFu, more0);
        const uint32_t who1 = __ballot_sync(0xFFFFFFFFu, more1);
        if ((who0 | who1) == 0u) return;
        const uint32_t n0 = __popc(who0);
        auto put = [&](uint32_t slot, const uint32_t *r, int done_upto) {
            uint32_t *it = sm.queue + slot * IW;
            uint32_t v[IW];
#pragma unroll
            for (int k = 0; k < IW; ++k) v[k] = 0u;
#pragma unroll
            for (int k = 0; k < RW; ++k) v[k] = r[k];
            v[RW] = (uint32_t)done_upto;
#pragma unroll
            for (int j = 0; j < IW / 4; ++j)
                reinterpret_cast<uint4 *>(it)[j] = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        };
        if (more0) put(waiting + __popc(who0 & lanemask_lt), r0, L0);
        if (more1) put(waiting + n0 + __popc(who1 & lanemask_lt), r1, L1);
        waiting += n0 + __popc(who1);
        __syncwarp();
    }

    // first_step() without the parking: what to park is handed back.
    __device__ __forceinline__ uint32_t first_step_only(const uint32_t (&x)[XW], bool &more, int &L) {
        uint32_t r[RW];
        mask_t colons;
        const bool literal = prepare(x, r, colons);
        uint32_t q = kNoQueue;
        more = false;
        L = 0;
        if (literal && colons != 0) {
            L = first_bit<mask_t>(colons);
            colons &= colons - 1;
            uint32_t id;
            bool ext;
            const bool hit = lookup(r, L, id, ext);
            q = hit ? id : kNoQueue;
            more = hit && ext && colons != 0;
        }
        return q;
    }

    // Resume up to 32 parked walks by one colon each.
    __device__ __forceinline__ void resume() {
        const uint32_t n = waiting < 32u ? waiting : 32u;
        const bool mine = lane < n;
        waiting -= n;
        uint32_t r[RW];
        int done_upto = 0;
        if (mine) {
            const uint32_t *it = sm.queue + (waiting + lane) * IW;
            uint32_t v[IW];
#pragma unroll
            for (int j = 0; j < IW / 4; ++j) {
                const uint4 t = reinterpret_cast<const uint4 *>(it)[j];
                v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
            }
#pragma unroll
            for (int k = 0; k < RW; ++k) r[k] = v[k];
            done_upto = (int)v[RW];
        } else {
#pragma unroll
            for (int k = 0; k < RW; ++k) r[k] = 0u;
        }
        __syncwarp();   // every lane has read its item before anyone parks on top of it
        // the exact colon mask, minus what the walk has already dealt with
        mask_t rest = colon_mask<true>(r) & ~((((mask_t)2) << done_upto) - (mask_t)1);
        bool more = false;
        int L = 0;
        if (mine && rest != 0) {
            L = first_bit<mask_t>(rest);
            rest &= rest - 1;
            uint32_t q;
            bool ext;
            if (lookup(r, L, q, ext)) {
                if (q != kNoQueue) atomicAdd(&sm.hist[q], 1u);
                more = ext && rest != 0;
            }
        }
        park(more, r, L);
    }

    // One record: its first colon is looked up here; the id of the queue it counts
    // for (kNoQueue if none) is returned for the warp histogram.
    __device__ __forceinline__ uint32_t first_step(const uint32_t (&x)[XW]) {
        uint32_t r[RW];
        mask_t colons;
        const bool literal = prepare(x, r, colons);
        uint32_t q = kNoQueue;
        bool more = false;
        int L = 0;
        if (literal && colons != 0) {
            L = first_bit<mask_t>(colons);
            colons &= colons - 1;
            uint32_t id;
            bool ext;
            const bool hit = lookup(r, L, id, ext);
            q = hit ? id : kNoQueue;                      // a virtual entry carries kNoQueue itself
            more = hit && ext && colons != 0;             // may over-report: resume() re-checks exactly
        }
        park(more, r, L);
        while (waiting >= 32u) resume();
        return q;
    }
};

template <int NCHUNK, bool WIDE, bool JOINT>
__global__ void __launch_bounds__(MatchCfg<NCHUNK>::kThreads, MatchCfg<NCHUNK>::kMinBlocks)
match_count_kernel(const uint8_t *__restrict__ heads, uint64_t n_keys,
                   const uint32_t *__restrict__ tbl, uint32_t n_prefix, uint32_t n_slots,
                   uint32_t log2_slots, uint32_t ecopy_shift, uint32_t one,
                   unsigned long long *__restrict__ counts) {
    using Cfg = MatchCfg<NCHUNK>;
    constexpr int XW = Cfg::kRecWords;
    constexpr int U = Cfg::kKeysPerThread;
    extern __shared__ uint4 smem4[];
    uint32_t *smem = reinterpret_cast<uint32_t *>(smem4);

    // Sizes come from the launch (they size the shared memory); a table image that
    // disagrees with them is reported by poisoning the counts, never by reading past it.
    if (tbl[0] != kTableMagic || tbl[1] != 16u * NCHUNK || tbl[2] != n_prefix || tbl[3] != n_slots ||
        (tbl[7] & 0xFFu) != log2_slots) {
        if (blockIdx.x == 0)
            for (uint32_t i = threadIdx.x; i < n_prefix; i += blockDim.x) counts[i] = ~0ull;
        return;
    }
    // Stage the table image ([kRows][kRWP] mask rows, [n_slots][kEW] entry rows) in the layout
    // the lookups read: 128-bit pieces, piece-major, every row in several copies.
    const uint32_t mask_words = Cfg::kRows * Cfg::kRWP * Cfg::kMaskCopies;
    const uint32_t entry_words = (n_slots * Cfg::kEW) << ecopy_shift;
    const uint32_t image_words = mask_words + entry_words;
    for (uint32_t i = threadIdx.x; i < Cfg::kRows * Cfg::kRWP; i += blockDim.x) {
        const uint32_t row = i / Cfg::kRWP, k = i % Cfg::kRWP, v = tbl[kTableHeaderWords + i];
        uint32_t *dst = smem + (((k / 4) * Cfg::kRows + row) * Cfg::kMaskCopies) * 4 + (k % 4);
        for (int c = 0; c < Cfg::kMaskCopies; ++c) dst[4 * c] = v;
    }
    for (uint32_t i = threadIdx.x; i < n_slots * Cfg::kEW; i += blockDim.x) {
        const uint32_t slot = i / Cfg::kEW, k = i % Cfg::kEW;
        const uint32_t v = tbl[kTableHeaderWords + Cfg::kRows * Cfg::kRWP + i];
        uint32_t *dst = smem + mask_words + ((((k / 4) * n_slots + slot) << ecopy_shift) * 4) + (k % 4);
        for (uint32_t c = 0; c < (1u << ecopy_shift); ++c) dst[4 * c] = v;
    }
    Matcher<NCHUNK> mt;
    mt.sm.masks = smem;
    mt.sm.entries = smem + mask_words;
    mt.mask_addr = (uint32_t)__cvta_generic_to_shared(smem) + (threadIdx.x & (Cfg::kMaskCopies - 1)) * 16u;
    mt.entry_addr = (uint32_t)__cvta_generic_to_shared(smem + mask_words) +
                    (threadIdx.x & ((1u << ecopy_shift) - 1u)) * 16u;
    mt.slot_byte_shift = ecopy_shift + 4u;
    mt.entry_piece_bytes = (n_slots << ecopy_shift) * 16u;
    mt.sm.hist = smem + image_words;
    mt.sm.queue = smem + image_words + ((n_prefix + 3u) & ~3u) +
                  (threadIdx.x >> 5) * (queue_items(JOINT) * Cfg::kItemWords);
    mt.len_mask = (typename Cfg::mask_t)((uint64_t)tbl[4] | ((uint64_t)tbl[5] << 32));
    mt.slot_shift = 32u - log2_slots;   // n_slots >= 8
    mt.seed = tbl[6];
    mt.two_slot = (tbl[7] >> 8) & 1u;
    mt.lane = threadIdx.x & 31u;
    mt.lanemask_lt = (1u << mt.lane) - 1u;
    mt.waiting = 0u;
    mt.one = one;
    for (uint32_t i = threadIdx.x; i < n_prefix; i += blockDim.x) mt.sm.hist[i] = 0u;
    __syncthreads();

    // U records per thread and step; the next step's records are already in flight
    // while this step's are examined (two register buffers, the loop is unrolled by two
    // so that they never have to be copied).
    auto load_tile = [&](uint64_t base, uint32_t (&x)[U][XW]) {
        // a tile that lies inside the snapshot (all but the last one or two of a block) is
        // loaded without a per-record range test: `base` is the same for the whole block
        const bool whole = base + (uint64_t)blockDim.x * U <= n_keys;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t idx = base + (uint64_t)u * blockDim.x + threadIdx.x;
            if (whole || idx < n_keys) {
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
    };
    // Block histogram: one shared-memory atomic per counting lane.  Measured against a
    // warp-aggregated form (MATCH.ANY + popc, one atomic per distinct queue per warp) this
    // is 6 % faster at 64 queues and no slower at one or two queues: the atomic unit
    // already serialises lanes that hit the same word, and the ten instructions of the
    // aggregation cost more integer-pipe time than they save.
    auto examine = [&](const uint32_t (&x)[U][XW]) {
        if constexpr (JOINT && U == 2) {
            bool more[U];
            int L[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t q = mt.first_step_only(x[u], more[u], L[u]);
                if (q != kNoQueue) atomicAdd(&mt.sm.hist[q], 1u);
            }
            mt.park_pair(more[0], &x[0][2], L[0], more[1], &x[1][2], L[1]);
            while (mt.waiting >= 32u) mt.resume();
        } else {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t q = mt.first_step(x[u]);
                if (q != kNoQueue) atomicAdd(&mt.sm.hist[q], 1u);
            }
        }
    };
    const uint64_t tile = (uint64_t)blockDim.x * U;
    const uint64_t stride = (uint64_t)gridDim.x * tile;
    uint64_t base = (uint64_t)blockIdx.x * tile;
    uint32_t xa[U][XW], xb[U][XW];
    if (base < n_keys) load_tile(base, xa);
    while (base < n_keys) {
        load_tile(base + stride, xb);          // past the end: zero records, no loads
        examine(xa);
        base += stride;
        if (base >= n_keys) break;
        load_tile(base + stride, xa);
        examine(xb);
        base += stride;
    }
    while (mt.waiting) mt.resume();
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n_prefix; i += blockDim.x) {
        const uint32_t c = mt.sm.hist[i];
        if (c) atomicAdd(&counts[i], (unsigned long long)c);
    }
}

static uint32_t slots_for(uint32_t n_prefix) {
    uint32_t want = n_prefix <= 128 ? 8 * n_prefix : (n_prefix <= 256 ? 4 * n_prefix : 2 * n_prefix + 256);
    uint32_t s = 8;
    while (s < want) s <<= 1;
    return s;
}

static size_t image_words(uint32_t n_prefix, uint32_t rec_bytes) {
    return (size_t)length_rows(rec_bytes) * region_words_padded(rec_bytes) +
           (size_t)slots_for(n_prefix) * entry_words(rec_bytes);
}

static uint32_t item_words(uint32_t rec_bytes) {
    return (region_words(rec_bytes) + 1 + 3) / 4 * 4;
}

static size_t match_smem_bytes(uint32_t n_prefix, uint32_t rec_bytes, bool joint = false,
                               uint32_t ecopy_shift = 0) {
    const size_t threads = rec_bytes <= 32 ? KAS_MATCH_THREADS : 256;   // MatchCfg<>::kThreads
    const size_t queues = (threads / 32) * queue_items(joint) * item_words(rec_bytes);
    const size_t mask_copies = rec_bytes <= 32 ? 8 : 1;                  // MatchCfg<>::kMaskCopies
    const size_t masks = (size_t)length_rows(rec_bytes) * region_words_padded(rec_bytes) * mask_copies;
    const size_t entries = ((size_t)slots_for(n_prefix) * entry_words(rec_bytes)) << ecopy_shift;
    return (masks + entries + ((n_prefix + 3u) & ~3u) + queues) * sizeof(uint32_t);
}

template <int NCHUNK, bool WIDE, bool JOINT = false>
static int launch_match(const void *d_heads, uint64_t n_keys, const void *d_tbl, uint32_t n_prefix,
                        uint64_t *d_counts, cudaStream_t stream) {
    const uint32_t rec_bytes = 16 * NCHUNK;
    // Shared memory beside the table goes first to the larger queue of joint parking, then to
    // copies of the entry rows (8, 4 or 2); measured on C2: joint + 4 copies 0.509 ms, joint + 1
    // copy 0.512 ms, 8 copies without joint parking 0.549 ms.
    // (Wide records run three 256-thread blocks per SM: extras only while that stays so.)
    const size_t extras_limit = NCHUNK <= 2 ? kMaxMatchSmem : 72 * 1024;
    if constexpr (!JOINT && NCHUNK <= 2) {   // two records per step: park them together if it fits
        if (match_smem_bytes(n_prefix, rec_bytes, true, 0) <= extras_limit)
            return launch_match<NCHUNK, WIDE, true>(d_heads, n_keys, d_tbl, n_prefix, d_counts, stream);
    }
    uint32_t ecopy_shift = 3;
    while (ecopy_shift > 0 && match_smem_bytes(n_prefix, rec_bytes, JOINT, ecopy_shift) > extras_limit) --ecopy_shift;
    const uint32_t n_slots = slots_for(n_prefix);
    uint32_t log2_slots = 0;
    while ((1u << log2_slots) < n_slots) ++log2_slots;
    const size_t smem = match_smem_bytes(n_prefix, rec_bytes, JOINT, ecopy_shift);
    if (smem > kMaxMatchSmem) return KAS_ERR_UNSUPPORTED;
    auto kern = match_count_kernel<NCHUNK, WIDE, JOINT>;
    if (smem > 48 * 1024)
        KAS_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    KAS_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, MatchCfg<NCHUNK>::kThreads, smem));
    if (per_sm < 1) return KAS_ERR_UNSUPPORTED;
    const int sms = sm_count();
    if (sms <= 0) return KAS_ERR_CUDA;
    const uint64_t tile = (uint64_t)MatchCfg<NCHUNK>::kThreads * MatchCfg<NCHUNK>::kKeysPerThread;
    uint64_t want = (n_keys + tile - 1) / tile;
    // persistent grid over the SMs, minus those the caller set aside (kas_set_reserved_sms) so that
    // a kernel on another stream -- the all-reduce of the previous step -- finds room beside it
    const int free_sms = sms - reserved_sms() > 0 ? sms - reserved_sms() : 1;
    uint64_t grid = (uint64_t)free_sms * per_sm;
    if (want < grid) grid = want;
    if (grid == 0) grid = 1;
    kern<<<(unsigned)grid, MatchCfg<NCHUNK>::kThreads, smem, stream>>>(
        static_cast<const uint8_t *>(d_heads), n_keys, static_cast<const uint32_t *>(d_tbl),
        n_prefix, n_slots, log2_slots, ecopy_shift, 1u, reinterpret_cast<unsigned long long *>(d_counts));
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
    if (match_smem_bytes(n_prefix, rec_bytes) > kMaxMatchSmem) return 0;
    return (kTableHeaderWords + image_words(n_prefix, rec_bytes)) * sizeof(uint32_t);
}

int kas_prefix_table_build(const uint8_t *names, const uint32_t *name_off, uint32_t n_prefix,
                           uint32_t rec_bytes, void *host_tbl, size_t tbl_bytes) {
    if (!name_off || !host_tbl || (!names && n_prefix && name_off[n_prefix] != 0)) return KAS_ERR_INVALID;
    if (n_prefix == 0 || n_prefix > KAS_MAX_PREFIXES) return KAS_ERR_INVALID;
    const size_t need = kas_prefix_table_bytes(n_prefix, rec_bytes);
    if (need == 0) return KAS_ERR_UNSUPPORTED;
    if (tbl_bytes < need) return KAS_ERR_TOO_SMALL;
    const uint32_t RW = region_words(rec_bytes), RWP = region_words_padded(rec_bytes);
    const uint32_t EW = entry_words(rec_bytes), rows = length_rows(rec_bytes);
    const uint32_t n_slots = slots_for(n_prefix);
    uint32_t log2_slots = 0;
    while ((1u << log2_slots) < n_slots) ++log2_slots;

    // entries: every name + ":" and every cut of it after an inner colon
    struct Entry { uint32_t q; bool ext; };
    std::map<std::string, Entry> entries;
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
        std::string s("ng-");                                               // kRegionLiteral bytes
        s.append(reinterpret_cast<const char *>(nm), len);
        s.push_back(':');
        auto it = entries.find(s);
        if (it != entries.end() && it->second.q != kVirtualQueue) return KAS_ERR_INVALID;  // duplicate name
        if (it == entries.end()) entries[s] = Entry{q, false};
        else it->second.q = q;
        for (uint32_t i = 0; i < len; ++i) {
            if (nm[i] != ':') continue;
            std::string cut("ng-");
            cut.append(reinterpret_cast<const char *>(nm), i + 1);        // includes the inner colon
            auto jt = entries.find(cut);
            if (jt == entries.end()) entries[cut] = Entry{kVirtualQueue, true};
            else jt->second.ext = true;
        }
    }
    if (entries.size() * 2 > (size_t)n_slots) return KAS_ERR_UNSUPPORTED;   // too many nested cuts

    uint32_t *t = static_cast<uint32_t *>(host_tbl);
    memset(t, 0, need);
    uint32_t *masks = t + kTableHeaderWords;
    uint32_t *slots = masks + (size_t)rows * RWP;
    for (uint32_t L = 0; L < rows; ++L)
        for (uint32_t k = 0; k < RW; ++k) masks[(size_t)L * RWP + k] = word_mask((int)k, (int)L + 1);

    // masked words, unfinished hash and meta of every entry
    struct Packed { uint32_t m[(KAS_MAX_REC_BYTES + 3) / 4]; uint32_t h, nbytes, q; bool ext; };
    std::vector<Packed> packed;
    uint64_t len_mask = 0;
    for (const auto &kv : entries) {
        const std::string &s = kv.first;
        Packed pk;
        memset(&pk, 0, sizeof(pk));
        pk.nbytes = (uint32_t)s.size();
        uint8_t buf[KAS_MAX_REC_BYTES + 8] = {0};
        memcpy(buf, s.data(), pk.nbytes);
        for (uint32_t k = 0; k < RW; ++k) {
            pk.m[k] = (uint32_t)buf[4 * k] | ((uint32_t)buf[4 * k + 1] << 8) |
                      ((uint32_t)buf[4 * k + 2] << 16) | ((uint32_t)buf[4 * k + 3] << 24);
            pk.h += pk.m[k] * hash_word_constant((int)k);
        }
        pk.q = kv.second.q == kVirtualQueue ? kNoQueue : kv.second.q;
        pk.ext = kv.second.ext;
        len_mask |= 1ull << (pk.nbytes - 1);
        packed.push_back(pk);
    }
    const uint32_t slot_shift = 32u - log2_slots;
    const size_t E = packed.size();
    std::vector<int> owner(n_slots);
    uint32_t seed = 0;
    bool placed = false, two_slot = false;
    // 1) a seed that puts every entry alone in its home slot
    for (uint32_t trial = 0; trial < 20000 && !placed; ++trial) {
        seed = trial * 0x9E3779B9u + 1u;
        std::fill(owner.begin(), owner.end(), -1);
        bool ok = true;
        for (size_t i = 0; i < E && ok; ++i) {
            const uint32_t a = hash_slot_a(hash_finish(packed[i].h, packed[i].nbytes, seed), slot_shift);
            if (owner[a] >= 0) ok = false;
            else owner[a] = (int)i;
        }
        placed = ok;
    }
    // 2) otherwise either of two slots (cuckoo insertion with bounded eviction walks)
    for (uint32_t trial = 0; trial < 64 && !placed; ++trial) {
        seed = trial * 0x7F4A7C15u + 0x51ED27u;
        std::fill(owner.begin(), owner.end(), -1);
        bool ok = true;
        for (size_t i = 0; i < E && ok; ++i) {
            int cur = (int)i;
            uint32_t avoid = 0xFFFFFFFFu;
            bool done = false;
            for (int kick = 0; kick < 512 && !done; ++kick) {
                const uint32_t tag = hash_finish(packed[cur].h, packed[cur].nbytes, seed);
                const uint32_t a = hash_slot_a(tag, slot_shift), b = hash_slot_b(tag, slot_shift);
                if (owner[a] < 0) { owner[a] = cur; done = true; }
                else if (owner[b] < 0) { owner[b] = cur; done = true; }
                else {
                    const uint32_t victim_slot = (a == avoid) ? b : a;
                    const int victim = owner[victim_slot];
                    owner[victim_slot] = cur;
                    cur = victim;
                    avoid = victim_slot;
                }
            }
            ok = done;
        }
        placed = ok;
        two_slot = ok;
    }
    if (!placed) return KAS_ERR_UNSUPPORTED;
    for (uint32_t sl = 0; sl < n_slots; ++sl) {
        if (owner[sl] < 0) continue;
        const Packed &pk = packed[owner[sl]];
        uint32_t *e = slots + (size_t)sl * EW;
        for (uint32_t k = 0; k < RW; ++k) e[k] = pk.m[k];
        e[RW] = pk.ext ? 1u : 0u;
        e[RW + 1] = pk.q;
    }
    t[0] = kTableMagic;
    t[1] = rec_bytes;
    t[2] = n_prefix;
    t[3] = n_slots;
    t[4] = (uint32_t)len_mask;
    t[5] = (uint32_t)(len_mask >> 32);
    t[6] = seed;
    t[7] = log2_slots | (two_slot ? 0x100u : 0u);
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
