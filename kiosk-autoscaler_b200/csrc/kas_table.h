// kas_table.h -- layout and hash of the queue-prefix table, shared by the host
// builder and the match kernel.
//
// A key matches queue q iff it starts with "processing-" + name_q + ":"
// (autoscaler.py:69 with a literal name).  Bytes 8.. of a record are the
// "region": "ng-" + name + ":" + ...  The kernel checks the first eight bytes
// ("processi") and walks the ':' bytes of the region in order; at a ':' at region
// offset P (= 3 + name length) the candidate string is region[0..P] INCLUDING the
// colon, and the table maps such strings to entries.
//
// The table is closed under "cut at an inner colon": for a name "a:b:c" the
// strings "a:" and "a:b:" are entries too (virtual ones, with no queue id) and
// carry an `ext` flag saying that a longer entry continues them.  So the walk
// can stop at the first colon whose candidate is not an entry, or whose entry
// has no continuation: nothing longer can match.  Nested queue names (a key
// counts for "predict" and for "predict:zstack") fall out of the same walk.
//
// Lookups never loop.  The host searches for a hash seed under which every
// entry sits alone in its home slot ("one-slot" tables: small name sets, the
// usual case); when no such seed is found entries may sit in either of two
// slots (cuckoo placement, "two-slot" tables).  A candidate is an entry iff its
// masked words equal the words stored in (one of) its slot(s); an empty slot
// holds zeros and a candidate always holds its ':'.
//
// Device image (uint32 words):
//   [0] magic  [1] rec_bytes  [2] n_prefix  [3] n_slots (power of two, >= 8)
//   [4] len_mask_lo  [5] len_mask_hi   (bit P set <=> some entry's ':' sits at region offset P)
//   [6] hash seed   [7] log2(n_slots) | two_slot << 8
//   length masks: (rec_bytes-8) rows of RWP words; row P holds, for each region
//       word, the mask of the bytes below P+1 (RWP = region words rounded up to 4)
//   entries: n_slots rows of EW words: RW region words (masked, zero padded),
//       a word that is non-zero when a longer entry continues this one (ext), and
//       the queue id (0xFFFFFFFF for a virtual entry).  EW = RW + 2 is a multiple of 4, so a row is read with 128-bit
//       shared loads.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define KAS_HD __host__ __device__ __forceinline__
#else
#define KAS_HD inline
#endif

namespace kas {

constexpr uint32_t kTableMagic = 0x4B415335u;  // "KAS5"
constexpr uint32_t kTableHeaderWords = 8;
constexpr uint32_t kLiteralLen = 11;           // strlen("processing-")
// The compared "region" of a record starts at byte 8, i.e. at the last three bytes
// "ng-" of the literal: it then consists of whole 32-bit words of the record (no
// byte shuffling on the device), and "ng-" is simply part of every stored string.
constexpr uint32_t kRegionStart = 8;
constexpr uint32_t kRegionLiteral = kLiteralLen - kRegionStart;   // 3
constexpr uint32_t kVirtualQueue = 0x7FFFFFFEu;   // builder-side marker of an entry without a queue

KAS_HD constexpr uint32_t region_words(uint32_t rec_bytes) { return (rec_bytes - kRegionStart) / 4; }
KAS_HD constexpr uint32_t region_words_padded(uint32_t rec_bytes) { return (region_words(rec_bytes) + 3) / 4 * 4; }
KAS_HD constexpr uint32_t entry_words(uint32_t rec_bytes) { return region_words(rec_bytes) + 2; }
// one mask row per position of the closing ':' inside the region (3 + name length)
KAS_HD constexpr uint32_t length_rows(uint32_t rec_bytes) { return rec_bytes - kRegionStart; }

// Mask selecting the bytes of little-endian region word k that lie below nbytes.
KAS_HD uint32_t word_mask(int k, int nbytes) {
    int rem = nbytes - 4 * k;
    return rem >= 4 ? 0xFFFFFFFFu : (rem <= 0 ? 0u : (0xFFFFFFFFu >> (32 - 8 * rem)));
}

// Hash of the masked region words: a multiply-add chain (IMAD on the device)
// with one odd constant per word, then a finaliser.  `tag` is never 0.
KAS_HD uint32_t hash_word_constant(int k) { return 0x9E3779B1u + 0x3C6EF372u * (uint32_t)k; }
KAS_HD uint32_t hash_finish(uint32_t h, uint32_t nbytes, uint32_t seed) {
    h += nbytes * 0x85EBCA6Bu + seed;
    return h ^ (h >> 15);   // the slot functions multiply and take the top bits
}
// slot_shift = 32 - log2(n_slots); n_slots >= 8, so the shift is below 32
KAS_HD uint32_t hash_slot_a(uint32_t tag, uint32_t slot_shift) { return (tag * 0x27D4EB2Fu) >> slot_shift; }
KAS_HD uint32_t hash_slot_b(uint32_t tag, uint32_t slot_shift) { return (tag * 0x165667B1u + 0x9E3779B9u) >> slot_shift; }

}  // namespace kas
