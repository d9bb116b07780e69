// kas_table.h -- layout and hash of the queue-prefix table, shared by the host
// builder and the match kernel.
//
// A key matches queue q iff it starts with "processing-" + name_q + ":"
// (autoscaler.py:69 with a literal name).  Bytes 11.. of a record are the
// "region"; a candidate is a ':' found at region offset L, the candidate
// string is region[0..L] INCLUDING the colon, and the table maps such strings
// to queue ids by open addressing.
//
// Device image (uint32 words):
//   [0] magic  [1] rec_bytes  [2] n_prefix  [3] n_slots (power of two)
//   [4] len_mask_lo  [5] len_mask_hi   (bit L set <=> some name has length L)
//   [6] region_words RW  [7] log2(n_slots)
//   then RW arrays of n_slots words: word k of every slot (SoA, so that lanes
//   probing different slots hit different banks), then n_slots queue ids.
//   An empty slot has word 0 == 0 (a stored string always holds its ':').
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define KAS_HD __host__ __device__ __forceinline__
#else
#define KAS_HD inline
#endif

namespace kas {

constexpr uint32_t kTableMagic = 0x4B415331u;  // "KAS1"
constexpr uint32_t kTableHeaderWords = 8;
constexpr uint32_t kLiteralLen = 11;           // strlen("processing-")

KAS_HD constexpr uint32_t region_words(uint32_t rec_bytes) { return (rec_bytes - kLiteralLen + 3) / 4; }

// Mask selecting the bytes of little-endian region word k that lie below nbytes.
KAS_HD uint32_t word_mask(int k, int nbytes) {
    int rem = nbytes - 4 * k;
    return rem >= 4 ? 0xFFFFFFFFu : (rem <= 0 ? 0u : (0xFFFFFFFFu >> (32 - 8 * rem)));
}

KAS_HD uint32_t hash_step(uint32_t h, uint32_t w) {
    h ^= w;
    h *= 0x9E3779B1u;
    return h ^ (h >> 15);
}
KAS_HD uint32_t hash_init(uint32_t nbytes) { return nbytes * 0x85EBCA6Bu + 0x27D4EB2Fu; }
KAS_HD uint32_t hash_slot(uint32_t h, uint32_t log2_slots) {
    h *= 0xC2B2AE35u;
    return log2_slots ? (h >> (32 - log2_slots)) : 0u;
}

}  // namespace kas
