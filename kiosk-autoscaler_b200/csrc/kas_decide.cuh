// kas_decide.cuh -- the replica decision (autoscaler.py:197-219) as device code,
// shared by the replay and the patch-event kernels.
#pragma once
#include "kas_common.cuh"

namespace kas {

struct PolDev {
    uint64_t magic;  // ceil(2^64 / |kpp|); 0 when |kpp| == 1
    int32_t mn, mx;
    int32_t neg;     // -1 when kpp < 0, else 0
    uint32_t k1;     // 1 when |kpp| == 1 (then magic == 0), else 0
    uint32_t m32;    // ceil(2^32 / kpp) - 1 when 1 <= kpp <= kSmallKppMax, else 0 (see quotient_small)
};
constexpr int32_t kSmallKppMax = 1 << 14;
constexpr uint32_t kSmallTallyTerm = 1u << 16;   // quotient_small: both terms of the tally below this

// Execution modes, chosen per launch by the host from the policies alone:
//   kGeneric  any kpp != 0, any min/max: exact 64-bit path.
//   kSimple   every policy has kpp >= 1 and min <= max and max >= 0 (every sane configuration):
//             a non-negative tally needs only a 32-bit multiply-high, the if/elif
//             clamp is max(min(d, max), min), "never scale down to a non-zero count"
//             is d > 0 ? max(d, cur) : d, and for a single-queue resource clipping
//             the one-term sum again (autoscaler.py:259) changes nothing.
//   kSimpleK1 kSimple with every kpp == 1 (scale.py:92 default): no division at all.
// Negative trace values take the exact path in every mode; all modes give the
// reference's integers.
enum ReplayMode { kGeneric = 0, kSimple = 1, kSimpleK1 = 2 };

// Host side: the device form of one policy (keys_per_pod != 0).
static inline PolDev make_poldev(const kas_policy &p) {
    PolDev d = {0, 0, 0, 0, 0, 0};
    const uint64_t ak = p.keys_per_pod < 0 ? (uint64_t)(-(int64_t)p.keys_per_pod) : (uint64_t)p.keys_per_pod;
    d.magic = ak == 1 ? 0ull : (~0ull / ak) + 1ull;   // ceil(2^64/ak) (exact when ak | 2^64 too)
    d.mn = p.min_pods;
    d.mx = p.max_pods;
    d.neg = p.keys_per_pod < 0 ? -1 : 0;
    d.k1 = ak == 1 ? 1u : 0u;
    if (p.keys_per_pod >= 1 && p.keys_per_pod <= kSmallKppMax)
        d.m32 = (uint32_t)((0xFFFFFFFFull + ak) / ak) - 1u;   // ceil(2^32 / ak) - 1
    return d;
}

// Execution mode for a set of policies (see kas_replay.cu); -1 when a keys_per_pod is 0.
static inline int replay_mode_for(const kas_policy *p, uint32_t n);

#ifdef __CUDACC__
// clip_pod_count (autoscaler.py:197-213), 32-bit form.  if/elif order kept.
__device__ __forceinline__ int32_t clip32(int32_t d, int32_t mn, int32_t mx, int32_t cur) {
    d = d > mx ? mx : (d < mn ? mn : d);
    return (d > 0 && d < cur) ? cur : d;
}
// Same with a 64-bit candidate (a sum of per-queue decisions, or a quotient).
__device__ __forceinline__ int32_t clip64(int64_t d, int32_t mn, int32_t mx, int32_t cur) {
    int32_t r = d > (int64_t)mx ? mx : (d < (int64_t)mn ? mn : (int32_t)d);
    return (r > 0 && r < cur) ? cur : r;
}

// tally // kpp with Python floor semantics for |tally| <= 2^32, any kpp != 0.
__device__ __forceinline__ int64_t floordiv_any(int64_t tally, const PolDev &p) {
    const int64_t t2 = (tally ^ (int64_t)p.neg) - (int64_t)p.neg;  // negate when kpp < 0
    const int64_t s = t2 >> 63;
    const uint64_t u = (uint64_t)(t2 ^ s);                          // floor(t/d) = ~((~t)/d), t < 0
    const uint64_t q = p.magic ? __umul64hi(p.magic, u) : u;
    return (int64_t)q ^ s;
}

// get_desired_pods (autoscaler.py:215-219) for one queue, exact for any input.
__device__ __forceinline__ int32_t decide_exact(int32_t llen, int32_t inprog, const PolDev &p, int32_t cur) {
    return clip64(floordiv_any((int64_t)llen + (int64_t)inprog, p), p.mn, p.mx, cur);
}

// The same for kpp >= 1, min <= max, max >= 0 and a tally in [0, 2^32), in two
// halves: the quotient depends on keys_per_pod alone (policies of a sweep that share
// it share the quotient), the rest on min / max / current pods.
template <bool K1>
__device__ __forceinline__ uint32_t quotient_simple(uint32_t tally, const PolDev &p) {
    if (K1) return tally;
    // floor(tally / kpp) = high 64 bits of magic * tally (the quotient fits 32 bits);
    // kpp == 1 has magic == 0 and k1 == 1, so the same multiply-adds give q = tally
    const uint32_t lo_hi = __umulhi((uint32_t)p.magic, tally);
    const uint32_t mh = (uint32_t)(((uint64_t)(uint32_t)(p.magic >> 32) * tally + lo_hi) >> 32);
    return tally * p.k1 + mh;
}
// floor(tally / kpp) for tally < 2^17 and 1 <= kpp <= 2^14 as ONE multiply-high:
// hi32((tally + 1) * m32) with m32 = ceil(2^32 / kpp) - 1 (the round-down magic number with
// the +1 on the dividend; exact over that whole range -- checked exhaustively in
// tests/test_host_logic.py).  The caller passes tally + 1.
__device__ __forceinline__ uint32_t quotient_small(uint32_t tally_plus_1, const PolDev &p) {
    return __umulhi(tally_plus_1, p.m32);
}
// The same quotient with the round-up magic number m_up = ceil(2^32 / kpp) = m32 + 1, which needs
// no + 1 on the dividend: exact for tally < 2^17 and 2 <= kpp <= 2^14 (the error term
// tally * (m_up * kpp - 2^32) stays below 2^32; checked exhaustively in tests/test_host_logic.py).
// kpp == 1 has no 32-bit m_up; its quotient is the tally.
__device__ __forceinline__ uint32_t quotient_up(uint32_t tally, uint32_t m_up) {
    return __umulhi(tally, m_up);
}
__device__ __forceinline__ int32_t finish_simple(uint32_t q, const PolDev &p, int32_t cur) {
    // if/elif clamp with 0 <= max and min <= max: the unsigned min already lands in
    // [0, max], so no saturation of q is needed before the signed max with min
    const int32_t d = max((int32_t)min(q, (uint32_t)p.mx), p.mn);
    return d > 0 ? max(d, cur) : d;
}
// finish_simple without the compare-and-select, for the policy sweep (integer ALU bound).
// d0 = max(min(q, max), min) is 0 exactly when min <= 0 and (q == 0 or max == 0); in that
// case the answer is 0, in every other case max(d0, cur).  With
//   keep = min(q + (min > 0), max > 0)            (0 or 1; shared by policies that agree on
//                                                  keys_per_pod, min > 0 and max > 0)
// both cases are max(min(q, max), min, cur * keep): one multiply on the FMA pipe and one
// three-input maximum (VIMNMX3) on the ALU pipe.
struct KeepTerm {
    uint32_t add;   // 1 when min_pods > 0
    uint32_t lim;   // 1 when max_pods > 0
};
__device__ __forceinline__ KeepTerm keep_term(const PolDev &p) {
    KeepTerm k;
    k.add = p.mn > 0 ? 1u : 0u;
    k.lim = p.mx > 0 ? 1u : 0u;
    return k;
}
__device__ __forceinline__ uint32_t keep_factor(uint32_t q, const KeepTerm &k, uint32_t one) {
    return min(q * one + k.add, k.lim);   // q <= 2^32 - 2, so q + 1 does not wrap
}
__device__ __forceinline__ int32_t finish_simple3(uint32_t q, uint32_t keep, const PolDev &p, int32_t cur) {
#ifdef KAS_NO_MAX3
    return max(max((int32_t)min(q, (uint32_t)p.mx), p.mn), cur * (int32_t)keep);
#else
    return __vimax3_s32((int32_t)min(q, (uint32_t)p.mx), p.mn, cur * (int32_t)keep);
#endif
}
template <bool K1>
__device__ __forceinline__ int32_t decide_simple(uint32_t tally, const PolDev &p, int32_t cur) {
    return finish_simple(quotient_simple<K1>(tally, p), p, cur);
}

// The `only_running` reading of a deployment (autoscaler.py:176-177) under the start-up
// model of kas.h: the specs of the last D ticks; what is available after a tick is the
// smallest of them and of the spec the tick leaves behind.
struct SpecWindow {
    // (a ring indexed by t % D; a shift register with constant indices, meant to keep it out of
    // local memory, measured 28 % slower: 14.1 against 11.0 ms per 2e9 queue-ticks)
    int32_t h[KAS_MAX_STARTUP_DELAY];
    __device__ __forceinline__ void init(int32_t spec0, uint32_t D) {
        for (uint32_t i = 0; i < D; ++i) h[i] = spec0;
    }
    // spec after tick t -> available replicas the next tick reads
    __device__ __forceinline__ int32_t push(int32_t spec, uint32_t t, uint32_t D) {
        int32_t avail = spec;
        for (uint32_t i = 0; i < D; ++i) avail = min(avail, h[i]);
        h[t % D] = spec;
        return avail;
    }
};

template <int MODE>
__device__ __forceinline__ int32_t decide(int32_t llen, int32_t inprog, const PolDev &p, int32_t cur) {
    if (MODE != kGeneric && ((llen | inprog) >= 0))
        return decide_simple<MODE == kSimpleK1>((uint32_t)llen + (uint32_t)inprog, p, cur);
    return decide_exact(llen, inprog, p, cur);
}

#endif  // __CUDACC__

static inline int replay_mode_for(const kas_policy *p, uint32_t n) {
    int mode = 2;  // kSimpleK1
    for (uint32_t s = 0; s < n; ++s) {
        if (p[s].keys_per_pod == 0) return -1;
        if (p[s].keys_per_pod < 0 || p[s].min_pods > p[s].max_pods || p[s].max_pods < 0) mode = 0;   // kGeneric
        else if (p[s].keys_per_pod != 1 && mode == 2) mode = 1;                        // kSimple
    }
    return mode;
}

}  // namespace kas
