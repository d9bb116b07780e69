// kas_replay_scan.cuh -- tick-parallel replay of single-queue resources (included by kas_replay.cu).
//
// The ticks of one resource are a sequential recurrence (cur[t+1] = desired[t]); a fleet hides that
// behind its width, a handful of resources cannot: one resource with a 10 000-tick trace is a
// 10 000-step dependency chain.  For a resource that owns ONE queue the chain can be cut.  With
// min_pods <= max_pods, keys_per_pod >= 1 and max_pods >= 0 (every sane policy) a tick maps the
// current pods c to
//     d = max(min(tally // kpp, max), min)                 (autoscaler.py:200-203, any sign of tally)
//     f_t(c) = d > 0 ? max(d, c) : d                       (:207-208; clipping the one-term sum again,
//                                                            :259, changes nothing)
// which is a CLAMP c -> max(lo, min(c, cap)):  (lo, cap) = (d, +inf) when d > 0, (d, d) otherwise.
// Clamps are closed under composition -- apply (lo1, cap1), then (lo2, cap2):
//     lo = clamp2(lo1), cap = clamp2(cap1)
// -- so the state after every tick is a prefix "sum" of an associative operator.  A block takes one
// (resource, policy): each thread composes the clamps of its run of consecutive ticks, a block-wide
// scan composes the runs, and each thread replays its run from the now known state before it,
// summing pod-ticks and counting patches (autoscaler.py:227).  Depth: 2 * T / threads + log2(threads)
// instead of T.  Resources with several queues keep the sequential kernels: their tick is a sum of
// such terms, whose compositions are not clamps.
#pragma once

namespace kas {

constexpr int kScanReplayThreads = 256;
constexpr uint32_t kScanReplayMinTicks = 1024;     // below this the warp-per-resource kernel is as fast

struct Clamp {
    int32_t lo, cap;       // c -> max(lo, min(c, cap)), lo <= cap
};
__device__ __forceinline__ int32_t clamp_apply(const Clamp &f, int32_t c) { return max(f.lo, min(c, f.cap)); }
// first f, then g
__device__ __forceinline__ Clamp clamp_then(const Clamp &f, const Clamp &g) {
    Clamp h;
    h.lo = clamp_apply(g, f.lo);
    h.cap = clamp_apply(g, f.cap);
    return h;
}
__device__ __forceinline__ Clamp clamp_identity() {
    Clamp f;
    f.lo = INT_MIN;
    f.cap = INT_MAX;
    return f;
}

// d of the comment above, exact for any int32 tally terms (the quotient is clamped in 64 bits).
__device__ __forceinline__ int32_t tick_target(int32_t llen, int32_t inprog, const PolDev &p) {
    const int64_t q = floordiv_any((int64_t)llen + (int64_t)inprog, p);
    return (int32_t)max(min(q, (int64_t)p.mx), (int64_t)p.mn);
}
__device__ __forceinline__ Clamp tick_clamp(int32_t d) {
    Clamp f;
    f.lo = d;
    f.cap = d > 0 ? INT_MAX : d;
    return f;
}

template <typename E>
__global__ void __launch_bounds__(kScanReplayThreads)
replay_scan_kernel(const __grid_constant__ ReplayParams P) {
    __shared__ Clamp s_warp[kScanReplayThreads / 32];
    __shared__ long long s_pt[kScanReplayThreads / 32];
    __shared__ uint32_t s_np[kScanReplayThreads / 32];
    __shared__ int32_t s_final;
    const uint64_t r = blockIdx.x / P.pol_count;
    const uint32_t slot = blockIdx.x % P.pol_count;
    const PolDev pol = P.pol[slot];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t T = P.T;
    const uint32_t per = (T + kScanReplayThreads - 1) / kScanReplayThreads;
    const uint32_t t_lo = min(T, tid * per), t_hi = min(T, t_lo + per);
    const E *lcol = static_cast<const E *>(P.llen) + r;
    const E *icol = static_cast<const E *>(P.inprog) + r;
    const int32_t ipq = P.inprog_mode == 1 ? static_cast<const int32_t *>(P.inprog)[r] : 0;
    auto target = [&](uint32_t t) {
        const int32_t l = ld_trace<E>(lcol + (uint64_t)t * P.Q);
        const int32_t ip = P.inprog_mode == 2 ? ld_trace<E>(icol + (uint64_t)t * P.Q) : ipq;
        return tick_target(l, ip, pol);
    };
    // 1. the clamp of this thread's run of ticks
    Clamp mine = clamp_identity();
    for (uint32_t t = t_lo; t < t_hi; ++t) mine = clamp_then(mine, tick_clamp(target(t)));
    // 2. inclusive scan over the threads (runs are in tick order: thread i before thread i + 1)
    Clamp incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        Clamp prev;
        prev.lo = __shfl_up_sync(0xFFFFFFFFu, incl.lo, o);
        prev.cap = __shfl_up_sync(0xFFFFFFFFu, incl.cap, o);
        if (lane >= (uint32_t)o) incl = clamp_then(prev, incl);
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    Clamp before = clamp_identity();                    // everything before this thread's run
    for (uint32_t w = 0; w < warp; ++w) before = clamp_then(before, s_warp[w]);
    {
        Clamp prev;
        prev.lo = __shfl_up_sync(0xFFFFFFFFu, incl.lo, 1);
        prev.cap = __shfl_up_sync(0xFFFFFFFFu, incl.cap, 1);
        if (lane > 0) before = clamp_then(before, prev);
    }
    // 3. replay the run from the state before it
    const int32_t c0 = initial_pods(P, r, slot);
    int32_t cur = clamp_apply(before, c0);
    long long pt = 0;
    uint32_t np = 0;
    const uint32_t index = P.pol_index[slot];
    for (uint32_t t = t_lo; t < t_hi; ++t) {
        const int32_t d = target(t);
        const int32_t want = d > 0 ? max(d, cur) : d;                    // autoscaler.py:255-259
        np += want != cur;                                               // :227
        cur = want;
        pt += want;
        if (P.dense) P.dense[((uint64_t)index * T + t) * P.R + r] = want;
    }
    // the last thread that owns a tick holds the final state
    if (t_lo < T && t_hi == T) s_final = cur;
    pt = warp_sum_ll(pt);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) np += __shfl_xor_sync(0xFFFFFFFFu, np, o);
    if (lane == 0) { s_pt[warp] = pt; s_np[warp] = np; }
    __syncthreads();
    if (tid == 0) {
        long long a = 0;
        uint32_t b = 0;
        for (int w = 0; w < kScanReplayThreads / 32; ++w) { a += s_pt[w]; b += s_np[w]; }
        const int32_t fin = s_final;
        if (P.totals) {
            kas_totals o;
            o.pod_ticks = a;
            o.n_patches = b;
            o.final_current = fin;
            P.totals[r * P.S + index] = o;
        }
        if (P.sums) {
            unsigned long long *o = P.sums + 3ull * index;
            atomicAdd(o + 0, (unsigned long long)a);
            atomicAdd(o + 1, (unsigned long long)b);
            atomicAdd(o + 2, (unsigned long long)(long long)fin);
        }
    }
}

// ---- any small resource: the tick chain cut by enumerating its few possible states ---------
// With 0 <= min_pods and max_pods <= 254 the pods of a resource only ever take the values
// 0..max_pods, or the starting count until the first change (clip_pod_count lands in [min, max] or
// keeps the current pods, autoscaler.py:197-213; a failed patch keeps them too, :267-273).  A run
// of ticks is then a map of that small set into itself, and such maps compose like any others:
// each thread of a block pushes EVERY state through its run of ticks (a table of max_pods + 2
// bytes), one thread chains the 256 tables from the starting pods, and each thread replays its run
// from the state it now knows, summing pod-ticks and counting patches.  This holds for resources
// with several queues as well (the reference's own QUEUES='predict,track' included): the per-queue
// targets of a tick do not depend on the state, only the hysteresis and the sum do.  Work grows
// with the number of states, depth shrinks by the number of threads: 10 000 ticks x 2 queues x 41
// policies took 5.1 ms on one warp per policy group and take 0.93 ms this way.
// A block whose policy or resource does not fit (min_pods < 0, max_pods > 255, more than
// kFsmMaxQueues queues) replays its trace on one thread, like a lane of the small-fleet kernel.
constexpr int kFsmThreads = 256;
constexpr uint32_t kFsmMaxQueues = 8;
constexpr int32_t kFsmMaxPods = 254;             // states 0..254 and the starting pods: an index fits a byte

template <typename E>
__global__ void __launch_bounds__(kFsmThreads)
replay_fsm_kernel(const __grid_constant__ ReplayParams P) {
    extern __shared__ uint8_t fsm_tab[];             // [kFsmThreads][stride]: state after the thread's run
    __shared__ uint16_t s_start[kFsmThreads];
    __shared__ long long s_pt[kFsmThreads / 32];
    __shared__ uint32_t s_np[kFsmThreads / 32];
    __shared__ int32_t s_final;
    const uint64_t r = blockIdx.x / P.pol_count;
    const uint32_t slot = blockIdx.x % P.pol_count;
    const PolDev pol = P.pol[slot];
    const uint32_t index = P.pol_index[slot];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t T = P.T;
    const uint64_t q0 = P.group_off ? P.group_off[r] : r;
    const uint64_t q1 = P.group_off ? P.group_off[r + 1] : r + 1;
    const uint32_t nq = q1 > q0 ? (uint32_t)min(q1 - q0, (uint64_t)0xFFFFFFFFu) : 0u;
    const int32_t c0 = initial_pods(P, r, slot);
    const bool fits = pol.mn >= 0 && pol.mx <= kFsmMaxPods && pol.mn <= pol.mx && nq <= kFsmMaxQueues &&
                      q1 - q0 <= kFsmMaxQueues;
    // the targets of one tick (autoscaler.py:200-203 per queue; exact for any count) ...
    auto targets = [&](uint32_t t, int32_t (&d)[kFsmMaxQueues]) {
        const E *lrow = static_cast<const E *>(P.llen) + (uint64_t)t * P.Q + q0;
        const E *irow = static_cast<const E *>(P.inprog) + (uint64_t)t * P.Q + q0;
        const int32_t *iq = static_cast<const int32_t *>(P.inprog) + q0;
#pragma unroll
        for (uint32_t j = 0; j < kFsmMaxQueues; ++j) {
            d[j] = 0;
            if (j < nq) {
                const int32_t l = ld_trace_cached<E>(lrow + j);
                const int32_t ip = P.inprog_mode == 2 ? ld_trace_cached<E>(irow + j) : (P.inprog_mode == 1 ? __ldg(iq + j) : 0);
                d[j] = tick_target(l, ip, pol);
            }
        }
    };
    // ... and the tick itself for current pods c: hysteresis per queue, sum, clip, a failed patch
    auto tick = [&](const int32_t (&d)[kFsmMaxQueues], bool failed, int32_t c) {
        long long sum = 0;                                                               // (c may be a huge cur0)
#pragma unroll
        for (uint32_t j = 0; j < kFsmMaxQueues; ++j)
            if (j < nq) sum += (d[j] > 0 && d[j] < c) ? c : d[j];                       // :207-208, :255-257
        const int32_t want = clip64(sum, pol.mn, pol.mx, c);                             // :259
        return (failed && want != c) ? c : want;                                         // :267-273
    };
    auto patch_failed = [&](uint32_t t) { return P.patch_fail && P.patch_fail[(uint64_t)t * P.R + r]; };

    long long pt = 0;
    uint32_t np = 0;
    if (fits) {
        const int32_t mx = pol.mx;
        const uint32_t n_states = (uint32_t)mx + 2u;       // 0..mx, and the starting pods
        const uint32_t stride = (n_states + 3u) & ~3u;
        const uint32_t per = (T + kFsmThreads - 1) / kFsmThreads;
        const uint32_t t_lo = min(T, tid * per), t_hi = min(T, t_lo + per);
        uint8_t *mine = fsm_tab + (size_t)tid * stride;
        auto value = [&](uint32_t k) { return k <= (uint32_t)mx ? (int32_t)k : c0; };
        auto state = [&](int32_t v) { return (v >= 0 && v <= mx) ? (uint32_t)v : (uint32_t)mx + 1u; };
        // 1. every state through this thread's run of ticks
        for (uint32_t k = 0; k < n_states; ++k) mine[k] = (uint8_t)k;
        for (uint32_t t = t_lo; t < t_hi; ++t) {
            int32_t d[kFsmMaxQueues];
            targets(t, d);
            const bool failed = patch_failed(t);
            // (in place: entry k is where state k stands after the ticks so far)
            for (uint32_t k = 0; k < n_states; ++k)
                mine[k] = (uint8_t)state(tick(d, failed, value(mine[k])));
        }
        __syncthreads();
        // 2. the state before every run: one thread walks the 256 tables
        if (tid == 0) {
            uint32_t k = state(c0);
            for (uint32_t i = 0; i < (uint32_t)kFsmThreads; ++i) {
                s_start[i] = (uint16_t)k;
                k = fsm_tab[(size_t)i * stride + k];
            }
        }
        __syncthreads();
        // 3. replay this run from its starting state
        int32_t cur = value(s_start[tid]);
        for (uint32_t t = t_lo; t < t_hi; ++t) {
            int32_t d[kFsmMaxQueues];
            targets(t, d);
            const bool failed = patch_failed(t);
            long long sum = 0;
#pragma unroll
            for (uint32_t j = 0; j < kFsmMaxQueues; ++j)
                if (j < nq) sum += (d[j] > 0 && d[j] < cur) ? cur : d[j];
            const int32_t want = clip64(sum, pol.mn, pol.mx, cur);
            np += want != cur;                                                           // the patch call is made (:227)
            if (!failed) cur = want;
            pt += cur;
            if (P.dense) P.dense[((uint64_t)index * T + t) * P.R + r] = cur;
        }
        if (t_lo < T && t_hi == T) s_final = cur;
    } else if (tid == 0) {
        // one thread, every tick in turn, exact 64-bit arithmetic (any policy, any group)
        int32_t cur = c0;
        for (uint32_t t = 0; t < T; ++t) {
            const E *lrow = static_cast<const E *>(P.llen) + (uint64_t)t * P.Q;
            const E *irow = static_cast<const E *>(P.inprog) + (uint64_t)t * P.Q;
            const int32_t *iq = static_cast<const int32_t *>(P.inprog);
            long long sum = 0;
            for (uint64_t q = q0; q < q1; ++q) {
                const int32_t l = ld_trace_cached<E>(lrow + q);
                const int32_t ip = P.inprog_mode == 2 ? ld_trace_cached<E>(irow + q) : (P.inprog_mode == 1 ? __ldg(iq + q) : 0);
                sum += decide_exact(l, ip, pol, cur);
            }
            const int32_t want = clip64(sum, pol.mn, pol.mx, cur);
            np += want != cur;
            if (!patch_failed(t)) cur = want;
            pt += cur;
            if (P.dense) P.dense[((uint64_t)index * T + t) * P.R + r] = cur;
        }
        s_final = cur;
    }
    pt = warp_sum_ll(pt);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) np += __shfl_xor_sync(0xFFFFFFFFu, np, o);
    if (lane == 0) { s_pt[warp] = pt; s_np[warp] = np; }
    __syncthreads();
    if (tid == 0) {
        long long a = 0;
        uint32_t b = 0;
        for (int w = 0; w < kFsmThreads / 32; ++w) { a += s_pt[w]; b += s_np[w]; }
        const int32_t fin = s_final;
        if (P.totals) {
            kas_totals o;
            o.pod_ticks = a;
            o.n_patches = b;
            o.final_current = fin;
            P.totals[r * P.S + index] = o;
        }
        if (P.sums) {
            unsigned long long *o = P.sums + 3ull * index;
            atomicAdd(o + 0, (unsigned long long)a);
            atomicAdd(o + 1, (unsigned long long)b);
            atomicAdd(o + 2, (unsigned long long)(long long)fin);
        }
    }
}

}  // namespace kas
