// kas_replay.cu -- K2: fused tally + desired-pods + clamp + hysteresis replay
// over a queue-depth trace, the element-wise decision kernel behind the scalar
// API, and the synthetic trace generator.
//
// Takes over T iterations of Autoscaler.scale (autoscaler.py:244-260) for R
// resources and S policies.  Pure integer work, HBM-bound at S = 1 (4 bytes of
// trace per decision) and ALU-bound under a policy sweep; nothing here is a
// contraction, so no tensor cores.  Ticks of one resource are sequentially
// dependent (cur[t+1] = desired[t]); resources and policies are independent,
// so a thread owns VEC adjacent resources x SP policies and marches down the
// tick-major trace, whose rows a warp reads as one contiguous 128/512-byte run.
//
// Kernels, chosen by replay_typed() at the bottom of this file:
//   replay_identity_kernel        single-queue resources, one policy (or a few, unvectorised)
//   replay_packed16_kernel        the same for a uint16 trace and a realistic policy: two
//                                 resources per register, 16-bit pair instructions
//   replay_sweep_kernel           single-queue resources under a policy sweep: trace tiles in
//                                 shared memory (TMA bulk copies), state 64 / 32 / 16 bits wide
//   replay_grouped_tiled_kernel   multi-queue resources, the block's run of queues staged in
//   replay_grouped_kernel         shared memory; the plain one also models the start-up delay
//   replay_small_kernel           few resources: a warp per resource, lanes are policies
//   replay_scan_kernel            few single-queue resources, long trace: the ticks in parallel
//                                 (kas_replay_scan.cuh)
//   decide_kernel, synth_trace_kernel
#include "kas_common.cuh"
#include "kas_decide.cuh"

#include <limits.h>
#include <algorithm>
#include <vector>
#include <type_traits>

namespace kas {

constexpr int kReplayThreads = 128;
constexpr int kMaxPoliciesPerLaunch = 32;
constexpr int kTickUnroll = 4;

struct ReplayParams {
    const void *llen;       // trace elements: int32 or uint16 (the kernels' template type E)
    const void *inprog;     // per-tick term: same element type; per-queue term: int32 [Q]
    const uint64_t *group_off;
    const int32_t *cur0;         // pods before tick 0: element (r, policy) at cur0[r * cur0_rstride + policy * cur0_sstride]
    uint32_t cur0_rstride;       // 1 / 0 for the caller's [R] array; 4 * S / 4 when it points into a kas_totals [R][S]
    uint32_t cur0_sstride;       //   array whose final_current fields a resumed replay starts from
    const uint8_t *patch_fail;   // [T][R] or null: the patch call of (tick, resource) raises and is swallowed
    kas_totals *totals;
    unsigned long long *sums;
    int32_t *dense;
    uint64_t Q, R;
    uint32_t T;
    uint32_t S;            // total policies (row stride of totals)
    uint32_t pol_first;    // global index of pol[0]
    uint32_t pol_count;    // policies in this launch
    int32_t inprog_mode;
    uint32_t startup_delay;   // ticks a replica needs to become available (grouped kernel only)
    int32_t one;           // the constant 1, opaque to the compiler (see pod-tick accumulation)
    uint16_t pol_index[kMaxPoliciesPerLaunch];   // slot -> index of the policy in the caller's list
    PolDev pol[kMaxPoliciesPerLaunch];
};

// Pods of resource r before the first tick under the policy in slot `slot` of this launch.
__device__ __forceinline__ int32_t initial_pods(const ReplayParams &P, uint64_t r, uint32_t slot) {
    if (!P.cur0) return 0;
    const uint32_t idx = P.pol_index[slot] == 0xFFFFu ? 0u : P.pol_index[slot];      // an empty slot is never written out
    return P.cur0[r * P.cur0_rstride + (uint64_t)idx * P.cur0_sstride];
}

// n += (a != b) as a compare and a predicated add (two instructions).
__device__ __forceinline__ void count_if_ne(uint32_t &n, int32_t a, int32_t b) {
    asm("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %1, %2;\n\t@p add.u32 %0, %0, 1;\n\t}" : "+r"(n) : "r"(a), "r"(b));
}

#ifndef KAS_PT_WIDE
#define KAS_PT_WIDE 1   // pod_ticks += want as one IMAD.WIDE (1) or as a 64-bit add (0)
#endif
__device__ __forceinline__ void add_pod_ticks(long long &pt, int32_t want, int32_t one) {
#if KAS_PT_WIDE
    pt = (long long)want * (long long)one + pt;
#else
    (void)one;
    pt += want;
#endif
}

__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}

// ---- every queue is its own resource (R == Q) ------------------------------
// E: storage type of the trace (int32_t or uint16_t).  IPMODE 0: no in-progress term, 1: one value per queue, 2: one per (tick, queue).
// Blocks are numbered policy-chunk fastest, so the chunks that read the same
// trace columns run together and share them through L2.
#ifndef KAS_REPLAY_SWEEP_SP
#define KAS_REPLAY_SWEEP_SP 2     // policies per thread under a sweep
#endif
#ifndef KAS_REPLAY_MINB_S1
#define KAS_REPLAY_MINB_S1 7      // resident blocks per SM (register budget), one policy per thread (6: 2.605 ms, 7: 2.450 ms, 8: 3.344 ms per 16 GB)
#endif
#ifndef KAS_REPLAY_MINB_SWEEP
#define KAS_REPLAY_MINB_SWEEP 4   // the same with several policies per thread
#endif
template <typename E, int VEC, int SP, int IPMODE, int MODE>
__global__ void __launch_bounds__(kReplayThreads,
                                  (SP != 1 ? KAS_REPLAY_MINB_SWEEP
                                           : (VEC * (IPMODE == 2 ? 2 : 1) >= 8 ? 4 : KAS_REPLAY_MINB_S1)))
replay_identity_kernel(const __grid_constant__ ReplayParams P) {
    constexpr int U = kTickUnroll;
    const uint32_t ny = (P.pol_count + SP - 1) / SP;
    const uint32_t pbase = (blockIdx.x % ny) * SP;   // index into P.pol
    const uint64_t r0 = ((uint64_t)(blockIdx.x / ny) * blockDim.x + threadIdx.x) * VEC;
    const bool valid = r0 < P.R;
    const uint32_t T = P.T;
    const uint64_t Q = P.Q;

    // this thread's policies, in registers; slots past the end repeat the last policy
    // and are never written out
    PolDev pol[SP];
    bool live[SP];
#pragma unroll
    for (int s = 0; s < SP; ++s) {
        live[s] = pbase + s < P.pol_count;
        pol[s] = P.pol[live[s] ? pbase + s : P.pol_count - 1];
    }
    const bool dense = P.dense != nullptr;
    const int32_t one = P.one;

    int32_t cur[SP][VEC];
    long long pt[SP][VEC];
    uint32_t np[SP][VEC];
    int32_t ipq[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
        ipq[v] = (valid && IPMODE == 1) ? static_cast<const int32_t *>(P.inprog)[r0 + v] : 0;
#pragma unroll
        for (int s = 0; s < SP; ++s) {
            cur[s][v] = (valid && live[s]) ? initial_pods(P, r0 + v, pbase + s) : 0;
            pt[s][v] = 0; np[s][v] = 0u;
        }
    }

    if (valid) {
        const E *lrow = static_cast<const E *>(P.llen) + r0;
        const E *irow = IPMODE == 2 ? static_cast<const E *>(P.inprog) + r0 : nullptr;
        // one tick of VEC resources under SP policies; FASTOK: no negative input in sight
        auto step = [&](uint32_t t, const int32_t (&l)[VEC], const int32_t (&ip)[VEC], auto fastok) {
            constexpr bool FASTOK = decltype(fastok)::value;
#pragma unroll
            for (int s = 0; s < SP; ++s) {
#pragma unroll
                for (int v = 0; v < VEC; ++v) {
                    const int32_t c = cur[s][v];
                    int32_t want;
                    if constexpr (FASTOK && MODE != kGeneric) {
                        const uint32_t tally = (uint32_t)l[v] + (IPMODE == 0 ? 0u : (uint32_t)ip[v]);
                        want = decide_simple<MODE == kSimpleK1>(tally, pol[s], c);        // :255-259
                    } else {
                        want = decide_exact(l[v], IPMODE == 0 ? 0 : ip[v], pol[s], c);    // :255-257
                        want = clip32(want, pol[s].mn, pol[s].mx, c);                     // :259
                    }
                    np[s][v] += (want != c);                                              // :227
                    cur[s][v] = want;
                    // pod_ticks += want as ONE 64-bit multiply-add on the FMA pipe (want * 1 + pt)
                    // instead of an add/add-with-carry pair on the integer ALU pipe, which is
                    // what limits this kernel under a policy sweep
                    pt[s][v] = (long long)want * (long long)one + pt[s][v];
                }
                if (dense && live[s]) {
                    int32_t *o = P.dense + ((uint64_t)P.pol_index[pbase + s] * T + t) * P.R + r0;
                    if constexpr (VEC > 1) {
#pragma unroll
                        for (int j = 0; j < VEC / 4; ++j)
                            reinterpret_cast<int4 *>(o)[j] =
                                make_int4(cur[s][4 * j], cur[s][4 * j + 1], cur[s][4 * j + 2], cur[s][4 * j + 3]);
                    } else {
                        o[0] = cur[s][0];
                    }
                }
            }
        };
        // rows t .. t+U-1 (clamped to the last tick) into registers
        auto load = [&](uint32_t t, int32_t (&l)[U][VEC], int32_t (&ip)[U][VEC]) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t tt = min(t + (uint32_t)u, T - 1);
                const uint64_t o = (uint64_t)tt * Q;
                if constexpr (VEC > 1) {
                    ld_trace_vec<E, VEC>(lrow + o, l[u]);
                    if constexpr (IPMODE == 2) ld_trace_vec<E, VEC>(irow + o, ip[u]);
                } else {
                    l[u][0] = ld_trace<E>(lrow + o);
                    if constexpr (IPMODE == 2) ip[u][0] = ld_trace<E>(irow + o);
                }
                if constexpr (IPMODE != 2) {
#pragma unroll
                    for (int v = 0; v < VEC; ++v) ip[u][v] = ipq[v];
                }
            }
        };
        // U ticks from registers; one sign test for the whole chunk picks the path
        auto run = [&](uint32_t t, const int32_t (&l)[U][VEC], const int32_t (&ip)[U][VEC]) {
            int32_t any = 0;
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int v = 0; v < VEC; ++v) any |= l[u][v] | (IPMODE == 0 ? 0 : ip[u][v]);
            if (MODE != kGeneric && any >= 0) {
#pragma unroll
                for (int u = 0; u < U; ++u) step(t + u, l[u], ip[u], std::true_type{});
            } else {
#pragma unroll
                for (int u = 0; u < U; ++u) step(t + u, l[u], ip[u], std::false_type{});
            }
        };
        // software pipeline: the next U rows are in flight while U rows are consumed
        int32_t la[U][VEC], ia[U][VEC], lb[U][VEC], ib[U][VEC];
        uint32_t t0 = 0;
        if (T >= 2 * U) {
            load(0, la, ia);
            for (; t0 + 2 * U <= T; t0 += 2 * U) {
                load(t0 + U, lb, ib);
                run(t0, la, ia);
                load(t0 + 2 * U, la, ia);
                run(t0 + U, lb, ib);
            }
        }
        for (; t0 < T; ++t0) {
            int32_t l[VEC], ip[VEC];
            const uint64_t o = (uint64_t)t0 * Q;
#pragma unroll
            for (int v = 0; v < VEC; ++v) {
                l[v] = ld_trace<E>(lrow + o + v);
                ip[v] = IPMODE == 2 ? ld_trace<E>(irow + o + v) : ipq[v];
            }
            step(t0, l, ip, std::false_type{});
        }
        if (P.totals) {
#pragma unroll
            for (int s = 0; s < SP; ++s)
                if (live[s])
#pragma unroll
                    for (int v = 0; v < VEC; ++v) {
                        kas_totals o;
                        o.pod_ticks = pt[s][v];
                        o.n_patches = np[s][v];
                        o.final_current = cur[s][v];
                        P.totals[(r0 + v) * P.S + P.pol_index[pbase + s]] = o;
                    }
        }
    }
    if (P.sums) {
#pragma unroll
        for (int s = 0; s < SP; ++s) {
            if (live[s]) {   // uniform across the block
                long long a = 0, b = 0, c = 0;
                if (valid) {
#pragma unroll
                    for (int v = 0; v < VEC; ++v) { a += pt[s][v]; b += np[s][v]; c += cur[s][v]; }
                }
                a = warp_sum_ll(a); b = warp_sum_ll(b); c = warp_sum_ll(c);
                if ((threadIdx.x & 31) == 0) {
                    unsigned long long *o = P.sums + 3ull * P.pol_index[pbase + s];
                    atomicAdd(o + 0, (unsigned long long)a);
                    atomicAdd(o + 1, (unsigned long long)b);
                    atomicAdd(o + 2, (unsigned long long)c);
                }
            }
        }
    }
}

// ---- one policy over a uint16 trace, packed arithmetic ---------------------
// The uint16 trace halves the bytes of the HBM-bound one-policy replay, but the kernel above
// then runs out of integer issue slots first (about 11 instructions per decision).  Here a
// 32-bit word of the trace -- two adjacent resources -- is never taken apart: sm_100's 16-bit
// pair instructions (VIMNMX.U16x2, VIMNMX3.S16x2, VIADD.16x2) make both decisions at once, and
// with keys_per_pod == 1 the loaded word already is the pair of quotients.  Same arithmetic as
// the sweep kernel's packed form (finish_simple3 in kas_decide.cuh); host-side conditions:
// uint16 trace, no in-progress term, no cur0, no failing patches, one policy with
// 0 <= min_pods, max_pods <= kPackedMaxPods and 1 <= keys_per_pod <= 2^14.
constexpr int32_t kPackedMaxPods = 1023;      // x kPackedFlushTicks stays below 2^16
constexpr uint32_t kPackedFlushTicks = 64;    // 16-bit patch / pod-tick sums are emptied this often

#ifndef KAS_PACKED16_MINB
#define KAS_PACKED16_MINB 8   /* 64 registers (a few spilled): latency hiding beats register comfort here (4: 1.322 ms, 8: 1.279 ms per 8 GB) */
#endif
template <bool K1>
__global__ void __launch_bounds__(kReplayThreads, KAS_PACKED16_MINB)
replay_packed16_kernel(const __grid_constant__ ReplayParams P) {
    constexpr int U = kTickUnroll, W = 4;       // ticks per batch; packed words (pairs of resources) per thread
    const uint64_t r0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * (2 * W);
    const bool valid = r0 < P.R;                // R is a multiple of 8 on this path
    const uint32_t T = P.T;
    const uint64_t Q = P.Q;
    const PolDev pol = P.pol[0];
    const uint32_t mx2 = (uint32_t)pol.mx * 0x00010001u, mn2 = (uint32_t)pol.mn * 0x00010001u;
    const bool min_positive = pol.mn > 0;       // then every decision is positive and keeps the current pods
    const bool dense = P.dense != nullptr;
    uint32_t cur2[W], np2[W], pa2[W];
    long long pt[2 * W];
    uint32_t np[2 * W];
#pragma unroll
    for (int w = 0; w < W; ++w) { cur2[w] = 0u; np2[w] = 0u; pa2[w] = 0u; }
#pragma unroll
    for (int v = 0; v < 2 * W; ++v) { pt[v] = 0; np[v] = 0u; }

    if (valid) {
        const uint16_t *lrow = static_cast<const uint16_t *>(P.llen) + r0;
        auto flush = [&]() {
#pragma unroll
            for (int w = 0; w < W; ++w) {
                pt[2 * w] += pa2[w] & 0xFFFFu;  pt[2 * w + 1] += pa2[w] >> 16;
                np[2 * w] += np2[w] & 0xFFFFu;  np[2 * w + 1] += np2[w] >> 16;
                pa2[w] = 0u; np2[w] = 0u;
            }
        };
        // one tick: the four words of `row` are this thread's eight resources
        auto step = [&](uint32_t t, const uint4 &row, auto densetag) {
            const uint32_t x[W] = {row.x, row.y, row.z, row.w};
#pragma unroll
            for (int w = 0; w < W; ++w) {
                uint32_t q2;
                if constexpr (K1) {
                    q2 = x[w];                                               // tally // 1, both halves
                } else {
                    const uint32_t qa = quotient_small((x[w] & 0xFFFFu) + 1u, pol);
                    const uint32_t qb = quotient_small((x[w] >> 16) + 1u, pol);
                    q2 = __byte_perm(qa, qb, 0x5410);
                }
                // keep factor as a mask: all ones when min_pods > 0, else 0xFFFF where q > 0
                const uint32_t km2 = min_positive ? 0xFFFFFFFFu : __vminu2(q2, 0x00010001u) * 0xFFFFu;
                const uint32_t c2 = cur2[w];
                const uint32_t w2 = __vimax3_s16x2(__vminu2(q2, mx2), mn2, c2 & km2);   // autoscaler.py:255-259
                np2[w] = __vadd2(np2[w], __vminu2(w2 ^ c2, 0x00010001u));            // :227
                pa2[w] = __vadd2(pa2[w], w2);
                cur2[w] = w2;
            }
            if constexpr (decltype(densetag)::value) {
                int32_t *o = P.dense + ((uint64_t)P.pol_index[0] * T + t) * P.R + r0;
                reinterpret_cast<int4 *>(o)[0] = make_int4((int32_t)(cur2[0] & 0xFFFFu), (int32_t)(cur2[0] >> 16),
                                                           (int32_t)(cur2[1] & 0xFFFFu), (int32_t)(cur2[1] >> 16));
                reinterpret_cast<int4 *>(o)[1] = make_int4((int32_t)(cur2[2] & 0xFFFFu), (int32_t)(cur2[2] >> 16),
                                                           (int32_t)(cur2[3] & 0xFFFFu), (int32_t)(cur2[3] >> 16));
            }
        };
        auto load = [&](uint32_t t, uint4 (&rows)[U]) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t tt = min(t + (uint32_t)u, T - 1);          // clamped; never used past T
                const int4 v = ldg_stream_i4(lrow + (uint64_t)tt * Q);
                rows[u] = make_uint4((uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w);
            }
        };
        auto run = [&](uint32_t t, const uint4 (&rows)[U]) {
            if (dense) {
#pragma unroll
                for (int u = 0; u < U; ++u) step(t + u, rows[u], std::true_type{});
            } else {
#pragma unroll
                for (int u = 0; u < U; ++u) step(t + u, rows[u], std::false_type{});
            }
        };
        // software pipeline: the next U rows are in flight while U rows are consumed
        uint4 ra[U], rb[U];
        uint32_t t0 = 0, since_flush = 0;
        if (T >= 2 * U) {
            load(0, ra);
            for (; t0 + 2 * U <= T; t0 += 2 * U) {
                load(t0 + U, rb);
                run(t0, ra);
                load(t0 + 2 * U, ra);
                run(t0 + U, rb);
                since_flush += 2 * U;
                if (since_flush + 2 * U > kPackedFlushTicks) { flush(); since_flush = 0; }
            }
        }
        flush();                                   // the tail below is shorter than 2 * U ticks
        for (; t0 < T; ++t0) {
            const int4 v = ldg_stream_i4(lrow + (uint64_t)t0 * Q);
            const uint4 row = make_uint4((uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w);
            if (dense) step(t0, row, std::true_type{}); else step(t0, row, std::false_type{});
        }
        flush();
        if (P.totals) {
#pragma unroll
            for (int v = 0; v < 2 * W; ++v) {
                kas_totals o;
                o.pod_ticks = pt[v];
                o.n_patches = np[v];
                o.final_current = (int32_t)((v & 1) ? cur2[v / 2] >> 16 : cur2[v / 2] & 0xFFFFu);
                P.totals[(r0 + v) * P.S + P.pol_index[0]] = o;
            }
        }
    }
    if (P.sums) {
        long long a = 0, b = 0, c = 0;
        if (valid) {
#pragma unroll
            for (int v = 0; v < 2 * W; ++v) {
                a += pt[v]; b += np[v];
                c += (long long)((v & 1) ? cur2[v / 2] >> 16 : cur2[v / 2] & 0xFFFFu);
            }
        }
        a = warp_sum_ll(a); b = warp_sum_ll(b); c = warp_sum_ll(c);
        if ((threadIdx.x & 31) == 0) {
            unsigned long long *o = P.sums + 3ull * P.pol_index[0];
            atomicAdd(o + 0, (unsigned long long)a);
            atomicAdd(o + 1, (unsigned long long)b);
            atomicAdd(o + 2, (unsigned long long)c);
        }
    }
}

// ---- resources that own several queues (CSR groups) ------------------------
template <typename E, int SP, int MODE, bool DELAY>
__global__ void __launch_bounds__(kReplayThreads)
replay_grouped_kernel(const __grid_constant__ ReplayParams P) {
    const uint32_t ny = (P.pol_count + SP - 1) / SP;
    const uint32_t pbase = (blockIdx.x % ny) * SP;
    const uint64_t r = (uint64_t)(blockIdx.x / ny) * blockDim.x + threadIdx.x;
    const bool valid = r < P.R;
    const uint32_t T = P.T;
    const uint64_t Q = P.Q;
    int32_t cur[SP];
    long long pt[SP];
    uint32_t np[SP];
    // DELAY (startup_delay > 0): cur[] is what the controller reads, the AVAILABLE replicas
    // (autoscaler.py:176-177); spec[] is what was last patched; win[] the specs of the last
    // D ticks.  Without it the two are the same number.
    const uint32_t D = DELAY ? P.startup_delay : 0u;
    int32_t spec[SP];
    SpecWindow win[DELAY ? SP : 1];
#pragma unroll
    for (int s = 0; s < SP; ++s) {
        const int32_t c0 = (valid && pbase + s < P.pol_count) ? initial_pods(P, r, pbase + s) : 0;
        cur[s] = c0; spec[s] = c0; pt[s] = 0; np[s] = 0u;
        if constexpr (DELAY) win[s].init(c0, D);
    }
    if (valid) {
        const uint64_t q0 = P.group_off ? P.group_off[r] : r;
        const uint64_t q1 = P.group_off ? P.group_off[r + 1] : r + 1;
        // a resource with ONE queue (and this kernel's extras: the start-up delay, failing patches):
        // the loads of the next ticks do not depend on the state, so eight ticks are fetched at
        // once instead of one load per tick at the head of the dependency chain
        constexpr uint32_t kAhead = 8;
        const bool single = q1 - q0 == 1;
        int32_t lbuf[kAhead], ibuf[kAhead];
        const int32_t *iq = static_cast<const int32_t *>(P.inprog);                  // per-queue term
        for (uint32_t t = 0; t < T; ++t) {
            const E *lrow = static_cast<const E *>(P.llen) + (uint64_t)t * Q;
            const E *irow = static_cast<const E *>(P.inprog) + (uint64_t)t * Q;      // per-tick term
            const bool failed = P.patch_fail && P.patch_fail[(uint64_t)t * P.R + r];
            long long sum[SP];
#pragma unroll
            for (int s = 0; s < SP; ++s) sum[s] = 0;
            if (single) {
                if (t % kAhead == 0) {
#pragma unroll
                    for (uint32_t u = 0; u < kAhead; ++u) {
                        const uint32_t tt = min(t + u, T - 1);
                        lbuf[u] = ld_trace<E>(static_cast<const E *>(P.llen) + (uint64_t)tt * Q + q0);
                        ibuf[u] = P.inprog_mode == 2 ? ld_trace<E>(static_cast<const E *>(P.inprog) + (uint64_t)tt * Q + q0)
                                                     : (P.inprog_mode == 1 ? __ldg(iq + q0) : 0);
                    }
                }
                int32_t l = lbuf[0], ip = ibuf[0];
#pragma unroll
                for (uint32_t u = 1; u < kAhead; ++u)
                    if (t % kAhead == u) { l = lbuf[u]; ip = ibuf[u]; }
#pragma unroll
                for (int s = 0; s < SP; ++s)
                    if (pbase + s < P.pol_count)
                        sum[s] += decide<MODE>(l, ip, P.pol[pbase + s], cur[s]);   // :255-257
            } else {
                for (uint64_t q = q0; q < q1; ++q) {
                    const int32_t l = ld_trace_cached<E>(lrow + q);
                    const int32_t ip = P.inprog_mode == 2 ? ld_trace_cached<E>(irow + q)
                                                          : (P.inprog_mode == 1 ? __ldg(iq + q) : 0);
#pragma unroll
                    for (int s = 0; s < SP; ++s)
                        if (pbase + s < P.pol_count)
                            sum[s] += decide<MODE>(l, ip, P.pol[pbase + s], cur[s]);   // :255-257
                }
            }
#pragma unroll
            for (int s = 0; s < SP; ++s) {
                if (pbase + s < P.pol_count) {
                    const PolDev &p = P.pol[pbase + s];
                    const int32_t want = clip64(sum[s], p.mn, p.mx, cur[s]);          // :259
                    if (want != cur[s]) {                                             // :227
                        np[s] += 1;                                                   // the patch call is made
                        if (!failed) spec[s] = want;                                  // raised and swallowed: :267-273
                    }
                    cur[s] = DELAY ? win[DELAY ? s : 0].push(spec[s], t, D) : spec[s];
                    pt[s] += spec[s];
                    if (P.dense) P.dense[((uint64_t)P.pol_index[pbase + s] * T + t) * P.R + r] = spec[s];
                }
            }
        }
        if (P.totals) {
#pragma unroll
            for (int s = 0; s < SP; ++s)
                if (pbase + s < P.pol_count) {
                    kas_totals o;
                    o.pod_ticks = pt[s];
                    o.n_patches = np[s];
                    o.final_current = spec[s];
                    P.totals[r * P.S + P.pol_index[pbase + s]] = o;
                }
        }
    }
    if (P.sums) {
#pragma unroll
        for (int s = 0; s < SP; ++s) {
            if (pbase + s < P.pol_count) {
                long long a = valid ? pt[s] : 0, b = valid ? (long long)np[s] : 0, c = valid ? (long long)spec[s] : 0;
                a = warp_sum_ll(a); b = warp_sum_ll(b); c = warp_sum_ll(c);
                if ((threadIdx.x & 31) == 0) {
                    unsigned long long *o = P.sums + 3ull * P.pol_index[pbase + s];
                    atomicAdd(o + 0, (unsigned long long)a);
                    atomicAdd(o + 1, (unsigned long long)b);
                    atomicAdd(o + 2, (unsigned long long)c);
                }
            }
        }
    }
}

// ---- policy sweeps over single-queue resources: the trace is read ONCE ---------
// Under a sweep every policy needs the same trace element.  A block owns a run of
// adjacent resources and ALL policies of the launch: tiles of the trace are staged in
// shared memory by TMA bulk copies landing on an mbarrier per stage (three tile
// buffers, one block barrier per tile) and every warp replays the tile for two policies and 128
// resources out of shared memory.  DRAM traffic is the trace itself, whatever the
// number of policies; the kernel is then bound by the integer pipes alone.
// A block always has about 16 warps (4 per scheduler): with C policy pairs it covers
// floor(16 / C) groups of 128 resources, and a tile holds 2048 / groups ticks.
constexpr int kSweepCols = 128;                 // resources per block (32 lanes x 4)
constexpr int kSweepTileRows = 64;              // ticks per tile when a block covers one group of columns
constexpr int kSweepStages = 3;                 // tiles in shared memory (one block barrier per tile)
constexpr int kSweepSP = 2;                     // policies per warp
constexpr int32_t kSweepNarrowBound = 1 << 24;  // x (at most 64 ticks per tile) stays below 2^32
constexpr int kSweepMaxWarps = kMaxPoliciesPerLaunch / kSweepSP;   // 16

__device__ __forceinline__ void cp_async_16(void *smem_dst, const void *gmem_src) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

#ifndef KAS_SWEEP_KEEPALL
#define KAS_SWEEP_KEEPALL 0    // packed form: a separate instance of the tick code for policy groups with min_pods > 0
                               // (one ALU instruction per pair and policy less, but 27.2 against 26.4 ms for int32 and
                               // 28.6 against 28.9 ms for uint16: the second copy of the tick code costs what it saves)
#endif
#ifndef KAS_SWEEP_FMA_ADDS
#define KAS_SWEEP_FMA_ADDS 1   // packed form, uint16 trace: patch / pod-tick sums of a tile as IMAD instead of VIADD.16x2
#endif
#ifndef KAS_SWEEP_TMA
#define KAS_SWEEP_TMA 1   // stage tiles with bulk asynchronous copies (TMA, cp.async.bulk + mbarrier); 0: cp.async
#endif
// mbarrier + 1-D bulk copy (global -> shared), the sm_90+/sm_100 TMA path for contiguous runs
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(bar);
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(a), "r"(parity) : "memory");
    }
}
__device__ __forceinline__ void bulk_copy_g2s(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src), "r"(bytes),
                   "r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}

// How the sweep kernel keeps its per-(resource, policy) state, chosen by the host:
//   kAccWide    pod-ticks in 64 bits every tick (any values)
//   kAccNarrow  32-bit pod-tick sums per tile (max_pods < kSweepNarrowBound, no cur0)
//   kAccPacked  two resources per register, 16 bits each: current pods, and per tile the
//               patches and pod-ticks (0 <= min_pods, max_pods <= kSweepPackedMaxPods, no
//               cur0, keys_per_pod <= 2^14); a batch of ticks with a count >= 2^14 is
//               replayed unpacked.  Every form gives the same integers.
enum SweepAcc { kAccWide = 0, kAccNarrow = 1, kAccPacked = 2 };
constexpr int32_t kSweepPackedMaxPods = 1023;           // x 64 ticks of a tile stays below 2^16
constexpr uint32_t kSweepPackedTallyTerm = 1u << 14;    // both terms below: the quotient fits 15 bits

template <typename E, bool TICKIP, int MODE, int ACC>
__global__ void __launch_bounds__(32 * kSweepMaxWarps, 1)
replay_sweep_kernel(const __grid_constant__ ReplayParams P) {
    constexpr int SP = kSweepSP, VEC = 4;
    constexpr bool NARROW = ACC == kAccNarrow;
    constexpr bool PACKED = ACC == kAccPacked;
    constexpr int kEPP = 16 / (int)sizeof(E);                 // trace elements per 16-byte piece
    constexpr int kGroupPieces = kSweepCols / kEPP;           // pieces per tick and column group
    extern __shared__ int4 sweep_smem[];            // [TICKIP ? 2 : 1][kSweepStages][ticks][groups][kGroupPieces] int4
    constexpr int kTileVec = kSweepTileRows * kGroupPieces;   // int4 per tile and array, whatever the split
    int4 *s_l = sweep_smem;                                  // llen tiles
    int4 *s_i = sweep_smem + kSweepStages * kTileVec;        // in-progress tiles (TICKIP only)

    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t chunks = (P.pol_count + SP - 1) / SP;          // policy pairs in this launch
    const uint32_t groups = blockDim.x / (32u * chunks);          // column groups of the block
    const uint32_t tile_ticks = kSweepTileRows / groups;
    const uint32_t pbase = (warp % chunks) * SP;
    const uint32_t cg = warp / chunks;
    const uint64_t col0 = (uint64_t)blockIdx.x * groups * kSweepCols;     // first resource of the block
    const uint64_t r0 = col0 + (uint64_t)cg * kSweepCols + 4ull * lane;
    const bool valid = r0 < P.R;
    const uint32_t T = P.T;
    const uint64_t Q = P.Q;
    const uint32_t row_vec = groups * kGroupPieces;                          // int4 per staged tick
    const uint32_t vec_here = (uint32_t)(min((uint64_t)groups * kSweepCols, P.R - col0) / kEPP);   // R % kEPP == 0

    PolDev pol[SP];
    bool live[SP];
#pragma unroll
    for (int s = 0; s < SP; ++s) {
        live[s] = pbase + s < P.pol_count && P.pol_index[pbase + s] != 0xFFFFu;   // not an empty slot
        pol[s] = P.pol[pbase + s < P.pol_count ? pbase + s : P.pol_count - 1];
    }
    const bool dense = P.dense != nullptr;
    const int32_t one = P.one;
    // The host pairs policies that agree on keys_per_pod, min > 0 and max > 0 (an odd group
    // is closed by an empty slot), so a warp's two policies share the quotient and the keep
    // factor: both come from pol[0].
    const KeepTerm keep = keep_term(pol[0]);
    const bool small_kpp = pol[0].m32 != 0u;
    int32_t cur[SP][VEC];
    long long pt[SP][VEC];
    uint32_t np[SP][VEC];
    int32_t ipq[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
        ipq[v] = (valid && P.inprog_mode == 1) ? static_cast<const int32_t *>(P.inprog)[r0 + v] : 0;
#pragma unroll
        for (int s = 0; s < SP; ++s) {
            cur[s][v] = (valid && live[s]) ? initial_pods(P, r0 + v, pbase + s) : 0;
            pt[s][v] = 0; np[s][v] = 0u;
        }
    }
    int32_t ipq_sign = 0;     // negative when a per-queue in-progress count is
#pragma unroll
    for (int v = 0; v < VEC; ++v) ipq_sign |= ipq[v];
    // NARROW: pod-ticks of the ticks of one tile are summed in 32 bits and added to the
    // 64-bit totals once per tile.  The host picks it when that cannot wrap: under the
    // simple arithmetic a decision is at most max(max_pods, the pods at the start), so
    // max_pods < kSweepNarrowBound for every policy and no cur0 suffices.  (One kernel per
    // choice: with both forms in one kernel the compiler merges them into the slower.)
    uint32_t pa[SP][VEC];
#pragma unroll
    for (int s = 0; s < SP; ++s)
#pragma unroll
        for (int v = 0; v < VEC; ++v) pa[s][v] = 0u;
    // PACKED: resources (0,1) and (2,3) of this lane share a register, low and high half.
    // cur2 is THE current-pods state of the kernel (cur[][] only lives inside a batch that is
    // replayed unpacked and after the last tile); np2 / pa2 are emptied once per tile.
    uint32_t cur2[SP][VEC / 2], np2[SP][VEC / 2], pa2[SP][VEC / 2], mx2[SP], mn2[SP];
#pragma unroll
    for (int s = 0; s < SP; ++s) {
        mx2[s] = (uint32_t)pol[s].mx * 0x00010001u;
        mn2[s] = (uint32_t)pol[s].mn * 0x00010001u;
#pragma unroll
        for (int p = 0; p < VEC / 2; ++p) { cur2[s][p] = 0u; np2[s][p] = 0u; pa2[s][p] = 0u; }
    }
    const uint32_t add2 = keep.add * 0x00010001u;
    // PACKED: 65536 as a value the compiler cannot see through, so that taking a 32-bit word of
    // the uint16 trace apart (w >> 16, w - (w >> 16 << 16)) and putting two quotients together
    // (hi * 65536 + lo) are multiplies on the FMA pipe: the integer ALU pipe is this kernel's limit
    const uint32_t k64k = 65536u * (uint32_t)one;
    const uint32_t m_up = pol[0].m32 + 1u;      // ceil(2^32 / keys_per_pod), keys_per_pod >= 2 (see quotient_up)
    const bool kpp_one = pol[0].k1 != 0u;
    const bool keep_all = keep.add != 0u;       // min_pods > 0 (and then max_pods > 0): the keep factor is always 1
    // per-queue in-progress counts as packed pairs, and whether they fit the packed form at all
    uint32_t ipq2[VEC / 2];
#pragma unroll
    for (int p = 0; p < VEC / 2; ++p) ipq2[p] = (uint32_t)ipq[2 * p] | ((uint32_t)ipq[2 * p + 1] << 16);

    const uint32_t n_tiles = (T + tile_ticks - 1) / tile_ticks;
#if KAS_SWEEP_TMA
    // stage tile `k` into buffer k % kSweepStages: one thread issues a bulk copy per
    // tick (the block's columns of a tick are one contiguous run) and the bytes land
    // on the stage's mbarrier; everybody waits on that barrier's phase.  (A variant
    // without the block barrier -- warps count themselves out of a tile and the last
    // one refills it -- measured 2 % slower.)
    __shared__ uint64_t bars[kSweepStages];
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < kSweepStages; ++i) mbar_init(&bars[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const uint32_t row_bytes_here = vec_here * 16u;
    auto stage = [&](uint32_t k) {
        if (threadIdx.x == 0 && k < n_tiles) {
            const uint32_t t0 = k * tile_ticks;
            const uint32_t nt = min(tile_ticks, T - t0);
            uint64_t *bar = &bars[k % kSweepStages];
            int4 *dl = s_l + (k % kSweepStages) * kTileVec;
            int4 *di = s_i + (k % kSweepStages) * kTileVec;
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // earlier reads of this buffer are done
            mbar_expect_tx(bar, nt * row_bytes_here * (TICKIP ? 2u : 1u));
            for (uint32_t t = 0; t < nt; ++t) {
                const uint64_t g = (uint64_t)(t0 + t) * Q + col0;
                bulk_copy_g2s(dl + t * row_vec, static_cast<const E *>(P.llen) + g, row_bytes_here, bar);
                if (TICKIP) bulk_copy_g2s(di + t * row_vec, static_cast<const E *>(P.inprog) + g, row_bytes_here, bar);
            }
        }
    };
    stage(0);
    stage(1);
    for (uint32_t k = 0; k < n_tiles; ++k) {
        mbar_wait(&bars[k % kSweepStages], (k / kSweepStages) & 1u);   // tile k has landed
        __syncthreads();               // all are done with tile k-1
        stage(k + 2);                  // refill the buffer tile k-1 occupied
#else
    // stage tile `k` into buffer k % kSweepStages: every thread copies 16-byte pieces
    auto stage = [&](uint32_t k) {
        if (k < n_tiles) {
            const uint32_t t0 = k * tile_ticks;
            const uint32_t nt = min(tile_ticks, T - t0);
            int4 *dl = s_l + (k % kSweepStages) * kTileVec;
            int4 *di = s_i + (k % kSweepStages) * kTileVec;
            for (uint32_t e = threadIdx.x; e < nt * row_vec; e += blockDim.x) {
                const uint32_t t = e / row_vec, c = e - t * row_vec;
                if (c < vec_here) {
                    const uint64_t g = (uint64_t)(t0 + t) * Q + col0 + (uint64_t)kEPP * c;
                    cp_async_16(dl + e, static_cast<const E *>(P.llen) + g);
                    if (TICKIP) cp_async_16(di + e, static_cast<const E *>(P.inprog) + g);
                }
            }
        }
        cp_async_commit();
    };
    stage(0);
    stage(1);
    for (uint32_t k = 0; k < n_tiles; ++k) {
        cp_async_wait<1>();            // this thread's pieces of tile k have landed
        __syncthreads();               // ... and everybody else's; all are done with tile k-1
        stage(k + 2);                  // refill the buffer tile k-1 occupied
#endif
        const uint32_t t0 = k * tile_ticks;
        const uint32_t nt = min(tile_ticks, T - t0);
        // this lane's four elements of a staged tick: element (cg * 128 + 4 * lane) of the row
        const E *tl = reinterpret_cast<const E *>(s_l + (k % kSweepStages) * kTileVec) + cg * kSweepCols + 4u * lane;
        const E *ti = reinterpret_cast<const E *>(s_i + (k % kSweepStages) * kTileVec) + cg * kSweepCols + 4u * lane;
        const uint32_t row_elems = row_vec * kEPP;
        const uint32_t working = __ballot_sync(0xFFFFFFFFu, valid && (live[0] || live[SP - 1]));
        if (valid && (live[0] || live[SP - 1])) {
            // four consecutive elements out of shared memory (one 128- or 64-bit load)
            auto read4 = [](const E *p, int32_t (&o)[VEC]) {
                if constexpr (sizeof(E) == 4) {
                    const int4 a = *reinterpret_cast<const int4 *>(p);
                    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w;
                } else {
                    const uint2 a = *reinterpret_cast<const uint2 *>(p);
                    o[0] = (int32_t)(a.x & 0xFFFFu); o[1] = (int32_t)(a.x >> 16);
                    o[2] = (int32_t)(a.y & 0xFFFFu); o[3] = (int32_t)(a.y >> 16);
                }
            };
            // G ticks at a time: their rows are read from shared memory first, one sign
            // test picks the arithmetic for all of them, then the recurrences run
            auto ticks = [&](uint32_t t, auto gtag, auto densetag) {
                constexpr int G = decltype(gtag)::value;
                constexpr bool DENSE = decltype(densetag)::value;   // chosen per tile, so the tick code has no branch
                // PACKED: two resources per instruction (min / max3 / add on 16-bit halves, sm_100's
                // VIMNMX.U16x2, VIMNMX3.S16x2, VIADDMNMX.U16x2, VIADD.16x2).  Per pair of resources the
                // policies of the warp share the quotients, packed q2, and the keep factors as a mask km2
                // (0xFFFF where keep == 1); per policy: min with max_pods, the three-input maximum with
                // min_pods and the kept current pods, and the two sums of the tile.
                // KEEPALL: the warp's policies have min_pods > 0, so every decision is positive and the
                // current pods are always kept (keep factor 1): no mask to build, no AND.  (Two plain
                // lambdas from one macro: a generic lambda with a tag parameter here crashes cicc 12.9.)
#define KAS_RUN_PACKED(NAME, KEEPALL)                                                                                  \
                auto NAME = [&](const uint32_t (&q2)[G][VEC / 2]) {                                                    \
                    _Pragma("unroll")                                                                                  \
                    for (int g = 0; g < G; ++g) {                                                                      \
                        uint32_t km2[VEC / 2];                                                                         \
                        _Pragma("unroll")                                                                              \
                        for (int p = 0; p < VEC / 2; ++p)                                                              \
                            km2[p] = KEEPALL ? 0xFFFFFFFFu : __viaddmin_u16x2(q2[g][p], add2, 0x00010001u) * 0xFFFFu;  \
                        _Pragma("unroll")                                                                              \
                        for (int s = 0; s < SP; ++s) {                                                                 \
                            _Pragma("unroll")                                                                          \
                            for (int p = 0; p < VEC / 2; ++p) {                                                        \
                                const uint32_t c2 = cur2[s][p];                                                        \
                                const uint32_t w2 = __vimax3_s16x2(__vminu2(q2[g][p], mx2[s]), mn2[s],                 \
                                                                   KEEPALL ? c2 : (c2 & km2[p])); /* :255-259 */      \
                                if constexpr (KAS_SWEEP_FMA_ADDS && sizeof(E) == 2) {                                  \
                                    /* the two sums of the tile as multiply-adds (no carry crosses the halves): */    \
                                    /* with a uint16 trace the FMA pipe has room (28.9 against 30.2 ms), with an */   \
                                    /* int32 trace it makes no difference (26.6 / 26.4 ms) */                         \
                                    np2[s][p] = __vminu2(w2 ^ c2, 0x00010001u) * (uint32_t)one + np2[s][p]; /* :227 */ \
                                    pa2[s][p] = w2 * (uint32_t)one + pa2[s][p];                                        \
                                } else {                                                                               \
                                    np2[s][p] = __vadd2(np2[s][p], __vminu2(w2 ^ c2, 0x00010001u));         /* :227 */ \
                                    pa2[s][p] = __vadd2(pa2[s][p], w2);                                                \
                                }                                                                                      \
                                cur2[s][p] = w2;                                                                       \
                            }                                                                                          \
                            if (DENSE && live[s]) {                                                                    \
                                int32_t *o = P.dense + ((uint64_t)P.pol_index[pbase + s] * T + t0 + t + g) * P.R + r0; \
                                *reinterpret_cast<int4 *>(o) = make_int4((int32_t)(cur2[s][0] & 0xFFFFu), (int32_t)(cur2[s][0] >> 16), \
                                                                         (int32_t)(cur2[s][1] & 0xFFFFu), (int32_t)(cur2[s][1] >> 16)); \
                            }                                                                                          \
                        }                                                                                              \
                    }                                                                                                  \
                }
                KAS_RUN_PACKED(run_packed_mask, false);
                KAS_RUN_PACKED(run_packed_keep, true);
#undef KAS_RUN_PACKED
                auto run_packed = [&](const uint32_t (&q2)[G][VEC / 2]) {
                    if (KAS_SWEEP_KEEPALL && keep_all) run_packed_keep(q2);
                    else run_packed_mask(q2);
                };
                if constexpr (PACKED) {
                    // The common batch -- every count below 2^14 -- never leaves the packed form: the
                    // tallies are formed as pairs straight from the staged rows (a 32-bit word of a
                    // uint16 trace IS a pair), the quotients pair by pair.
                    uint32_t q2[G][VEC / 2];
                    uint32_t big = (uint32_t)ipq_sign;
                    if constexpr (sizeof(E) == 2) {
                        uint32_t w[G][VEC / 2];
#pragma unroll
                        for (int g = 0; g < G; ++g) {
                            const uint2 a = *reinterpret_cast<const uint2 *>(tl + (t + g) * row_elems);
                            w[g][0] = a.x; w[g][1] = a.y;
                            big |= (a.x | a.y) & 0xC000C000u;
                            if (TICKIP) {
                                const uint2 b = *reinterpret_cast<const uint2 *>(ti + (t + g) * row_elems);
                                big |= (b.x | b.y) & 0xC000C000u;
                                w[g][0] += b.x; w[g][1] += b.y;            // halves below 2^15: no carry between them
                            } else {
                                w[g][0] += ipq2[0]; w[g][1] += ipq2[1];    // (zero without a per-queue term)
                            }
                        }
                        if (__all_sync(working, big < kSweepPackedTallyTerm)) {
#pragma unroll
                            for (int g = 0; g < G; ++g)
#pragma unroll
                                for (int p = 0; p < VEC / 2; ++p) {
                                    if (kpp_one) {
                                        q2[g][p] = w[g][p];                                    // tally // 1, both halves
                                    } else {
                                        const uint32_t hi = __umulhi(w[g][p], k64k);           // w >> 16
                                        const uint32_t lo = w[g][p] - hi * k64k;
                                        q2[g][p] = quotient_up(hi, m_up) * k64k + quotient_up(lo, m_up);
                                    }
                                }
                            run_packed(q2);
                            return;
                        }
                    } else {
                        uint32_t x[G][VEC];
#pragma unroll
                        for (int g = 0; g < G; ++g) {
                            const int4 a = *reinterpret_cast<const int4 *>(tl + (t + g) * row_elems);
                            x[g][0] = (uint32_t)a.x; x[g][1] = (uint32_t)a.y; x[g][2] = (uint32_t)a.z; x[g][3] = (uint32_t)a.w;
                            big |= (uint32_t)(a.x | a.y | a.z | a.w);
                            if (TICKIP) {
                                const int4 b = *reinterpret_cast<const int4 *>(ti + (t + g) * row_elems);
                                big |= (uint32_t)(b.x | b.y | b.z | b.w);
                                x[g][0] += (uint32_t)b.x; x[g][1] += (uint32_t)b.y; x[g][2] += (uint32_t)b.z; x[g][3] += (uint32_t)b.w;
                            } else {
                                // (zero without a per-queue term; skipping the adds under a uniform
                                // branch measured 9 % slower)
#pragma unroll
                                for (int v = 0; v < VEC; ++v) x[g][v] += (uint32_t)ipq[v];
                            }
                        }
                        if (__all_sync(working, big < kSweepPackedTallyTerm)) {
#pragma unroll
                            for (int g = 0; g < G; ++g)
#pragma unroll
                                for (int p = 0; p < VEC / 2; ++p) {
                                    if (kpp_one) q2[g][p] = x[g][2 * p + 1] * k64k + x[g][2 * p];
                                    else q2[g][p] = quotient_up(x[g][2 * p + 1], m_up) * k64k + quotient_up(x[g][2 * p], m_up);
                                }
                            run_packed(q2);
                            return;
                        }
                    }
                    // a count too large (or negative): this batch unpacked
#pragma unroll
                    for (int s = 0; s < SP; ++s)
#pragma unroll
                        for (int p = 0; p < VEC / 2; ++p) {
                            cur[s][2 * p] = (int32_t)(cur2[s][p] & 0xFFFFu);
                            cur[s][2 * p + 1] = (int32_t)(cur2[s][p] >> 16);
                        }
                }
                int32_t l[G][VEC], ip[G][VEC];
                int32_t any = ipq_sign;
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    read4(tl + (t + g) * row_elems, l[g]);
                    if (TICKIP) {
                        read4(ti + (t + g) * row_elems, ip[g]);
                        if constexpr (sizeof(E) == 4 || PACKED) {
#pragma unroll
                            for (int v = 0; v < VEC; ++v) any |= ip[g][v];
                        }
                    } else {
#pragma unroll
                        for (int v = 0; v < VEC; ++v) ip[g][v] = ipq[v];
                    }
                    if constexpr (sizeof(E) == 4 || PACKED) {   // a uint16 element is neither negative nor >= 2^16
#pragma unroll
                        for (int v = 0; v < VEC; ++v) any |= l[g][v];
                    }
                }
                auto run = [&](auto fasttag, auto smalltag) {
                    constexpr bool FAST = decltype(fasttag)::value;
                    constexpr bool SMALLQ = decltype(smalltag)::value;   // one multiply-high per quotient
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        uint32_t q0[VEC], k0[VEC];
                        if constexpr (FAST) {
#pragma unroll
                            for (int v = 0; v < VEC; ++v) {
                                if constexpr (SMALLQ)
                                    q0[v] = quotient_small((uint32_t)l[g][v] + (uint32_t)ip[g][v] + 1u, pol[0]);
                                else
                                    q0[v] = quotient_simple<false>((uint32_t)l[g][v] + (uint32_t)ip[g][v], pol[0]);
                                k0[v] = keep_factor(q0[v], keep, (uint32_t)one);
                            }
                        }
#pragma unroll
                        for (int s = 0; s < SP; ++s) {
#pragma unroll
                            for (int v = 0; v < VEC; ++v) {
                                const int32_t c = cur[s][v];
                                int32_t want;
                                if constexpr (FAST) {
                                    want = finish_simple3(q0[v], k0[v], pol[s], c);                                   // :255-259
                                } else {
                                    want = decide_exact(l[g][v], ip[g][v], pol[s], c);                                 // :255-257
                                    want = clip32(want, pol[s].mn, pol[s].mx, c);                                     // :259
                                }
                                count_if_ne(np[s][v], want, c);                                                       // :227
                                cur[s][v] = want;
                                // (issuing these two adds on the FMA pipe instead measured 1-8 % slower)
                                if constexpr (NARROW && FAST) pa[s][v] += (uint32_t)want;
                                else add_pod_ticks(pt[s][v], want, one);   // also a PACKED kernel's unpacked batch
                            }
                            if (DENSE && live[s]) {
                                int32_t *o = P.dense + ((uint64_t)P.pol_index[pbase + s] * T + t0 + t + g) * P.R + r0;
                                *reinterpret_cast<int4 *>(o) = make_int4(cur[s][0], cur[s][1], cur[s][2], cur[s][3]);
                            }
                        }
                    }
                };
                // every count of the batch below 2^16 (and keys_per_pod <= 2^14): short quotient
                // (a whole-warp choice, so that one large count does not make the warp run both forms)
                if (MODE != kGeneric && small_kpp && __all_sync(working, (uint32_t)any < kSmallTallyTerm))
                    run(std::true_type{}, std::true_type{});
                else if (MODE != kGeneric && any >= 0) run(std::true_type{}, std::false_type{});
                else run(std::false_type{}, std::false_type{});
                if constexpr (PACKED) {
#pragma unroll
                    for (int s = 0; s < SP; ++s)
#pragma unroll
                        for (int p = 0; p < VEC / 2; ++p)
                            cur2[s][p] = (uint32_t)cur[s][2 * p] | ((uint32_t)cur[s][2 * p + 1] << 16);
                }
            };
            uint32_t t = 0;
            if (dense) {
                for (; t + 4 <= nt; t += 4) ticks(t, std::integral_constant<int, 4>{}, std::true_type{});
                for (; t < nt; ++t) ticks(t, std::integral_constant<int, 1>{}, std::true_type{});
            } else {
                for (; t + 4 <= nt; t += 4) ticks(t, std::integral_constant<int, 4>{}, std::false_type{});
                for (; t < nt; ++t) ticks(t, std::integral_constant<int, 1>{}, std::false_type{});
            }
            if constexpr (NARROW) {
#pragma unroll
                for (int s = 0; s < SP; ++s)
#pragma unroll
                    for (int v = 0; v < VEC; ++v) { pt[s][v] += pa[s][v]; pa[s][v] = 0u; }
            }
            if constexpr (PACKED) {
#pragma unroll
                for (int s = 0; s < SP; ++s)
#pragma unroll
                    for (int p = 0; p < VEC / 2; ++p) {
                        pt[s][2 * p] += pa2[s][p] & 0xFFFFu;  pt[s][2 * p + 1] += pa2[s][p] >> 16;
                        np[s][2 * p] += np2[s][p] & 0xFFFFu;  np[s][2 * p + 1] += np2[s][p] >> 16;
                        pa2[s][p] = 0u; np2[s][p] = 0u;
                    }
            }
        }
    }
#if !KAS_SWEEP_TMA
    cp_async_wait<0>();
#endif
    if constexpr (PACKED) {
#pragma unroll
        for (int s = 0; s < SP; ++s)
#pragma unroll
            for (int p = 0; p < VEC / 2; ++p) {
                cur[s][2 * p] = (int32_t)(cur2[s][p] & 0xFFFFu);
                cur[s][2 * p + 1] = (int32_t)(cur2[s][p] >> 16);
            }
    }

    if (valid && P.totals) {
#pragma unroll
        for (int s = 0; s < SP; ++s)
            if (live[s])
#pragma unroll
                for (int v = 0; v < VEC; ++v) {
                    kas_totals o;
                    o.pod_ticks = pt[s][v];
                    o.n_patches = np[s][v];
                    o.final_current = cur[s][v];
                    P.totals[(r0 + v) * P.S + P.pol_index[pbase + s]] = o;
                }
    }
    if (P.sums) {
#pragma unroll
        for (int s = 0; s < SP; ++s) {
            if (live[s]) {   // uniform across the warp
                long long a = 0, b = 0, c = 0;
                if (valid) {
#pragma unroll
                    for (int v = 0; v < VEC; ++v) { a += pt[s][v]; b += np[s][v]; c += cur[s][v]; }
                }
                a = warp_sum_ll(a); b = warp_sum_ll(b); c = warp_sum_ll(c);
                if (lane == 0) {
                    unsigned long long *o = P.sums + 3ull * P.pol_index[pbase + s];
                    atomicAdd(o + 0, (unsigned long long)a);
                    atomicAdd(o + 1, (unsigned long long)b);
                    atomicAdd(o + 2, (unsigned long long)c);
                }
            }
        }
    }
}

// ---- small fleets: one warp per resource, lanes are policies ------------------
// A handful of resources (one reference pod manages ONE) cannot fill the machine
// with a thread per resource, and a lone thread pays a full memory round trip per
// tick.  Here a warp owns one resource: the lanes load a tile of ticks of its
// queues into shared memory together, then every lane replays the tile for its own
// policy out of shared memory (all lanes read the same word: a broadcast).
constexpr int kSmallTileElems = 2048;      // trace elements per tile and warp (8 KB, twice with a per-tick in-progress trace)
constexpr uint64_t kSmallMaxResources = 4096;

template <typename E, int MODE>
__global__ void __launch_bounds__(32)
replay_small_kernel(const __grid_constant__ ReplayParams P) {
    __shared__ int32_t s_l[kSmallTileElems];
    __shared__ int32_t s_i[kSmallTileElems];
    const uint64_t r = blockIdx.x;
    const uint32_t lane = threadIdx.x;
    const bool live = lane < P.pol_count;
    const PolDev pol = P.pol[live ? lane : 0];
    const uint64_t q0 = P.group_off ? P.group_off[r] : r;
    const uint32_t nq = (uint32_t)((P.group_off ? P.group_off[r + 1] : r + 1) - q0);
    const uint32_t T = P.T;
    const uint32_t tile_ticks = nq ? kSmallTileElems / nq : T;
    int32_t cur = live ? initial_pods(P, r, lane) : 0;
    long long pt = 0;
    uint32_t np = 0;
    for (uint32_t t0 = 0; t0 < T; t0 += tile_ticks) {
        const uint32_t nt = min(tile_ticks, T - t0);
        const uint32_t elems = nt * nq;
        __syncwarp();
        for (uint32_t e = lane; e < elems; e += 32) {
            const uint32_t t = e / nq, j = e - t * nq;
            const uint64_t g = (uint64_t)(t0 + t) * P.Q + q0 + j;
            s_l[e] = ld_trace<E>(static_cast<const E *>(P.llen) + g);
            s_i[e] = P.inprog_mode == 2 ? ld_trace<E>(static_cast<const E *>(P.inprog) + g)
                                        : (P.inprog_mode == 1 ? __ldg(static_cast<const int32_t *>(P.inprog) + q0 + j) : 0);
        }
        __syncwarp();
        if (live) {
            for (uint32_t t = 0; t < nt; ++t) {
                long long sum = 0;
                for (uint32_t j = 0; j < nq; ++j)
                    sum += decide<MODE>(s_l[t * nq + j], s_i[t * nq + j], pol, cur);      // :255-257
                const int32_t want = clip64(sum, pol.mn, pol.mx, cur);                    // :259
                if (want != cur) {                                                        // :227
                    np += 1;                                                              // the patch call is made
                    if (!(P.patch_fail && P.patch_fail[(uint64_t)(t0 + t) * P.R + r])) cur = want;   // :267-273
                }
                pt += cur;
                if (P.dense) P.dense[((uint64_t)P.pol_index[lane] * T + t0 + t) * P.R + r] = cur;
            }
        }
    }
    if (live) {
        if (P.totals) {
            kas_totals o;
            o.pod_ticks = pt;
            o.n_patches = np;
            o.final_current = cur;
            P.totals[r * P.S + P.pol_index[lane]] = o;
        }
        if (P.sums) {
            unsigned long long *o = P.sums + 3ull * P.pol_index[lane];
            atomicAdd(o + 0, (unsigned long long)pt);
            atomicAdd(o + 1, (unsigned long long)np);
            atomicAdd(o + 2, (unsigned long long)(long long)cur);
        }
    }
}

// ---- element-wise decision (scalar API) ------------------------------------
__device__ __forceinline__ long long clip_ll(long long d, long long mn, long long mx, long long cur) {
    d = d > mx ? mx : (d < mn ? mn : d);
    return (d > 0 && d < cur) ? cur : d;
}

__global__ void __launch_bounds__(256)
decide_kernel(const long long *__restrict__ a, const long long *__restrict__ kpp,
              const long long *__restrict__ mn, const long long *__restrict__ mx,
              const long long *__restrict__ cur, uint64_t n, int mode,
              long long *__restrict__ out, int *__restrict__ status) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (uint64_t)gridDim.x * blockDim.x) {
        long long d = a[i];
        if (mode == 0) {
            const long long k = kpp[i];
            if (k == 0) {            // reference: ZeroDivisionError (autoscaler.py:217)
                *status = 1;
                out[i] = 0;
                continue;
            }
            if (d == LLONG_MIN && k == -1) {
                d = LLONG_MAX;       // true quotient 2^63 exceeds every max_pods: the clamp picks max
            } else {
                long long q = d / k;
                if ((d % k != 0) && ((d < 0) != (k < 0))) q -= 1;   // Python //
                d = q;
            }
        }
        out[i] = clip_ll(d, mn[i], mx[i], cur[i]);
    }
}

// ---- totals of consecutive tick chunks (resume) ------------------------------
__global__ void __launch_bounds__(256)
totals_chain_kernel(kas_totals *__restrict__ acc, const kas_totals *__restrict__ chunk, uint64_t n,
                    long long *__restrict__ sums_acc, const long long *__restrict__ sums_chunk, uint32_t S) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        kas_totals a = acc[i];
        const kas_totals c = chunk[i];
        a.pod_ticks += c.pod_ticks;
        a.n_patches += c.n_patches;
        a.final_current = c.final_current;
        acc[i] = a;
    }
    if (sums_acc && blockIdx.x == 0 && threadIdx.x < S) {
        long long *a = sums_acc + 3 * threadIdx.x;
        const long long *c = sums_chunk + 3 * threadIdx.x;
        a[0] += c[0];
        a[1] += c[1];
        a[2] = c[2];
    }
}

// ---- synthetic trace (spec in DESIGN.md; twin in oracle/kas_oracle.c) ------
__device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

__global__ void __launch_bounds__(256)
synth_trace_kernel(uint64_t seed, uint32_t T, uint64_t q_first, uint64_t n_q, int32_t *__restrict__ out) {
    const uint64_t total = (uint64_t)T * n_q;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t t = (uint32_t)(i / n_q);
        const uint64_t q = q_first + (i % n_q);
        const uint64_t epoch = t >> 6;
        const uint64_t he = mix64(mix64(seed ^ (q * 0xD1342543DE82EF95ULL)) + epoch);
        const uint64_t ht = mix64(he ^ ((uint64_t)t * 0x2545F4914F6CDD1DULL));
        const uint32_t active = (uint32_t)(he % 100u) < 45u;
        const uint32_t level = (uint32_t)((he >> 8) % 400u);
        const uint32_t jitter = (uint32_t)(ht & 0xFFu);
        int32_t v = 0;
        if (active) v = (int32_t)((level * jitter) >> 7);
        else if (jitter < 10u) v = (int32_t)((ht >> 8) & 3u);
        out[i] = v;
    }
}

// ---- resources that own several queues, trace staged in shared memory ------
// replay_grouped_kernel above reads the trace element by element: a thread's loads of one
// tick depend on nothing, but nothing makes them early either, and neighbouring threads
// touch the same sectors again and again (measured 0.67-0.86 TB/s).  Here the block -- 128
// consecutive resources, whose queues are one contiguous run of the queue axis -- copies that
// run for a tile of ticks into shared memory with 16-byte cp.async pieces, two tiles in
// flight, and the threads then replay their resource out of shared memory.  The run is
// widened to 16-byte boundaries, which needs 16-byte aligned rows (host-checked).  A block
// whose run does not fit a tile even for one tick takes the direct loads.
#ifndef KAS_GROUPED_TILE_KB
#define KAS_GROUPED_TILE_KB 12   /* 8 blocks per SM: this kernel lives on latency hiding (24 KB: 8.1 ms, 12 KB: 6.7 ms) */
#endif
constexpr int kGroupedTileBytes = KAS_GROUPED_TILE_KB * 1024;     // per stage and array
constexpr int kGroupedStages = 2;
constexpr uint32_t kGroupedMaxTileTicks = 32;

template <typename E, int SP, int MODE, bool TICKIP>
__global__ void __launch_bounds__(kReplayThreads)
replay_grouped_tiled_kernel(const __grid_constant__ ReplayParams P) {
    constexpr uint32_t kEPP = 16 / sizeof(E);
    constexpr uint32_t kTileVec = kGroupedTileBytes / 16;     // int4 per stage and array
    extern __shared__ int4 gsm[];                              // [stages][TICKIP ? 2 : 1][kTileVec]
    const uint32_t ny = (P.pol_count + SP - 1) / SP;
    const uint32_t pbase = (blockIdx.x % ny) * SP;
    const uint64_t rb = (uint64_t)(blockIdx.x / ny) * blockDim.x;      // first resource of the block
    const uint64_t r = rb + threadIdx.x;
    const bool valid = r < P.R;
    const uint32_t T = P.T;
    const uint64_t Q = P.Q;
    // the block's run of the queue axis, widened to whole 16-byte pieces
    const uint64_t r_end = min(rb + (uint64_t)blockDim.x, P.R);
    const uint64_t qa = P.group_off[rb], qb = P.group_off[r_end];
    const uint64_t a0 = qa & ~(uint64_t)(kEPP - 1);
    const uint64_t a1 = (qb + kEPP - 1) & ~(uint64_t)(kEPP - 1);       // <= Q: Q is a multiple of kEPP
    const uint64_t row_vec64 = (a1 - a0) / kEPP;                        // int4 per tick
    const uint32_t tile_ticks = row_vec64 == 0 ? kGroupedMaxTileTicks
                                : (uint32_t)min((uint64_t)kGroupedMaxTileTicks, (uint64_t)kTileVec / row_vec64);
    const uint32_t row_vec = (uint32_t)row_vec64;
    int32_t cur[SP];
    long long pt[SP];
    uint32_t np[SP];
#pragma unroll
    for (int s = 0; s < SP; ++s) {
        cur[s] = (valid && pbase + s < P.pol_count) ? initial_pods(P, r, pbase + s) : 0;
        pt[s] = 0; np[s] = 0u;
    }
    const uint64_t q0 = valid ? P.group_off[r] : 0, q1 = valid ? P.group_off[r + 1] : 0;
    // The staged run covers this block's queues only if the offsets do not go back or leave
    // [0, Q]; a block that sees such offsets reads element by element, like the plain kernel.
    const bool stray = !(qa <= qb && qb <= Q) || (valid && !(qa <= q0 && q0 <= q1 && q1 <= qb));
    const bool tiled = !__syncthreads_or(stray ? 1 : 0) && tile_ticks > 0;   // uniform over the block
    const int32_t *iq = static_cast<const int32_t *>(P.inprog);          // per-queue term
    // this thread's policies in registers; slots past the end repeat the last one, unwritten
    PolDev pol[SP];
    bool live[SP];
#pragma unroll
    for (int s = 0; s < SP; ++s) {
        live[s] = pbase + s < P.pol_count;
        pol[s] = P.pol[live[s] ? pbase + s : P.pol_count - 1];
    }
    const uint32_t n_own = (uint32_t)min(q1 - q0, (uint64_t)0xFFFFFFFFu);   // queues of this resource
    const bool huge = q1 - q0 > 0xFFFFFFFFull;                              // (never tiled then)

    // one tick of this thread's resource; LOAD(q) -> (llen, inprog) of queue q at this tick
    auto tick = [&](uint32_t t, auto load) {
        const bool failed = P.patch_fail && P.patch_fail[(uint64_t)t * P.R + r];
        long long sum[SP];
#pragma unroll
        for (int s = 0; s < SP; ++s) sum[s] = 0;
        // Simple arithmetic first, with no test per queue: the signs of everything read are
        // collected, and a tick that saw a negative count is summed again the exact way.
        int32_t any = 0;
        auto one_queue = [&](uint64_t q, auto fasttag) {
            int32_t l, ip;
            load(q, l, ip);
            if constexpr (decltype(fasttag)::value) {
                any |= l | ip;
#pragma unroll
                for (int s = 0; s < SP; ++s)
                    sum[s] += decide_simple<false>((uint32_t)l + (uint32_t)ip, pol[s], cur[s]);   // autoscaler.py:255-257
            } else {
#pragma unroll
                for (int s = 0; s < SP; ++s) sum[s] += decide_exact(l, ip, pol[s], cur[s]);
            }
        };
        auto all_queues = [&](auto fasttag) {
            if (!huge) {
                for (uint32_t j = 0; j < n_own; ++j) one_queue(q0 + j, fasttag);
            } else {
                for (uint64_t q = q0; q < q1; ++q) one_queue(q, fasttag);
            }
        };
        if constexpr (MODE != kGeneric) {
            all_queues(std::true_type{});
            if (any < 0) {
#pragma unroll
                for (int s = 0; s < SP; ++s) sum[s] = 0;
                all_queues(std::false_type{});
            }
        } else {
            all_queues(std::false_type{});
        }
#pragma unroll
        for (int s = 0; s < SP; ++s) {
            if (live[s]) {
                const PolDev &p = pol[s];
                const int32_t want = clip64(sum[s], p.mn, p.mx, cur[s]);          // :259
                if (want != cur[s]) {                                             // :227
                    np[s] += 1;                                                   // the patch call is made
                    if (!failed) cur[s] = want;                                   // raised and swallowed: :267-273
                }
                pt[s] += cur[s];
                if (P.dense) P.dense[((uint64_t)P.pol_index[pbase + s] * T + t) * P.R + r] = cur[s];
            }
        }
    };

    if (tiled) {
        const uint32_t n_tiles = (T + tile_ticks - 1) / tile_ticks;
        auto stage = [&](uint32_t k) {
            if (k < n_tiles) {
                const uint32_t t0 = k * tile_ticks;
                const uint32_t nt = min(tile_ticks, T - t0);
                int4 *dl = gsm + (k % kGroupedStages) * (TICKIP ? 2 : 1) * kTileVec;
                int4 *di = dl + kTileVec;
                for (uint32_t e = threadIdx.x; e < nt * row_vec; e += blockDim.x) {
                    const uint32_t t = e / row_vec, c = e - t * row_vec;
                    const uint64_t g = (uint64_t)(t0 + t) * Q + a0 + (uint64_t)kEPP * c;
                    cp_async_16(dl + e, static_cast<const E *>(P.llen) + g);
                    if (TICKIP) cp_async_16(di + e, static_cast<const E *>(P.inprog) + g);
                }
            }
            cp_async_commit();
        };
        stage(0);
        for (uint32_t k = 0; k < n_tiles; ++k) {
            stage(k + 1);                  // its buffer was released by the barrier that ended tile k-1
            cp_async_wait<1>();            // this thread's pieces of tile k have landed
            __syncthreads();               // ... and everybody else's
            if (valid) {
                const uint32_t t0 = k * tile_ticks;
                const uint32_t nt = min(tile_ticks, T - t0);
                const E *tl = reinterpret_cast<const E *>(gsm + (k % kGroupedStages) * (TICKIP ? 2 : 1) * kTileVec);
                const E *ti = tl + (size_t)kTileVec * kEPP;
                const uint32_t row_elems = row_vec * kEPP;
                for (uint32_t t = 0; t < nt; ++t) {
                    const E *lrow = tl + (size_t)t * row_elems;
                    const E *irow = ti + (size_t)t * row_elems;
                    tick(t0 + t, [&](uint64_t q, int32_t &l, int32_t &ip) {
                        const uint32_t j = (uint32_t)(q - a0);          // place inside the staged run
                        l = (int32_t)lrow[j];
                        ip = TICKIP ? (int32_t)irow[j] : (P.inprog_mode == 1 ? __ldg(iq + q) : 0);
                    });
                }
            }
            __syncthreads();               // everybody is done with tile k
        }
        cp_async_wait<0>();
    } else if (valid) {
        for (uint32_t t = 0; t < T; ++t) {
            const E *lrow = static_cast<const E *>(P.llen) + (uint64_t)t * Q;
            const E *irow = static_cast<const E *>(P.inprog) + (uint64_t)t * Q;
            tick(t, [&](uint64_t q, int32_t &l, int32_t &ip) {
                l = ld_trace_cached<E>(lrow + q);
                ip = TICKIP ? ld_trace_cached<E>(irow + q) : (P.inprog_mode == 1 ? __ldg(iq + q) : 0);
            });
        }
    }
    if (valid && P.totals) {
#pragma unroll
        for (int s = 0; s < SP; ++s)
            if (live[s]) {
                kas_totals o;
                o.pod_ticks = pt[s];
                o.n_patches = np[s];
                o.final_current = cur[s];
                P.totals[r * P.S + P.pol_index[pbase + s]] = o;
            }
    }
    if (P.sums) {
#pragma unroll
        for (int s = 0; s < SP; ++s) {
            if (live[s]) {
                long long a = valid ? pt[s] : 0, b = valid ? (long long)np[s] : 0, c = valid ? (long long)cur[s] : 0;
                a = warp_sum_ll(a); b = warp_sum_ll(b); c = warp_sum_ll(c);
                if ((threadIdx.x & 31) == 0) {
                    unsigned long long *o = P.sums + 3ull * P.pol_index[pbase + s];
                    atomicAdd(o + 0, (unsigned long long)a);
                    atomicAdd(o + 1, (unsigned long long)b);
                    atomicAdd(o + 2, (unsigned long long)c);
                }
            }
        }
    }
}

// ---- multi-queue resources, the lean form ------------------------------------
// The tiled kernel above spends about 40 instructions per (queue, tick) with 20 of 32 lanes
// busy: neighbouring threads own groups of different sizes, so a warp runs as long as its
// largest group, and everything is carried in 64 bits.  This kernel serves the common case --
// realistic policies (1 <= keys_per_pod <= 2^14, 0 <= max_pods < 2^20, min <= max), no cur0, no
// failing patches, no dense output -- with
//   * threads SORTED by group size inside the block (a counting sort in shared memory), so the
//     lanes of a warp own groups of one size and the per-queue loop is unrolled for it; the
//     warps of a block are rotated by a hash of the block index, because the heavy warps of
//     all blocks would otherwise meet on one scheduler;
//   * the block's run of the queue axis staged tile by tile with one TMA bulk copy per tick row
//     (three stages on mbarriers): no thread spends an instruction on moving the trace;
//   * 32-bit arithmetic: the quotient is one multiply-high (quotient_small), clamp and
//     hysteresis are min / keep factor / three-input maximum as in the sweep kernel
//     (kas_decide.cuh), for the per-queue decisions (autoscaler.py:255-257) and for the clip of
//     their sum (:259) alike.  A tick that holds a count outside [0, 2^16) is replayed by the
//     exact 64-bit code.
// Offsets that are not non-decreasing inside [0, Q], a group of 2048 queues or more, a run too
// wide for a tile: that block reads the trace element by element with the exact arithmetic.
// Block sizes: 512 resources per block for fleets that fill the machine (a larger block sorts
// a larger population, so more warps hold a single size class: 30.8 of 32 lanes busy against
// 22.3 with 128; measured 1.37 ms against 1.72 ms per 8 GB), 128 for smaller fleets.  Stages:
// 96 bytes per thread and stage, two stages for the large block, three for the small one.
#ifndef KAS_GS_BIG_THREADS
#define KAS_GS_BIG_THREADS 512
#endif
constexpr int kGsSmallThreads = 128;
constexpr int kGsBigThreads = KAS_GS_BIG_THREADS;
constexpr uint32_t kGsStageBytesPerThread = 96;
constexpr uint32_t kGsMaxTileTicks = 32;
constexpr uint32_t kGsMaxOwn = 2048;          // queues of one resource: x kGsMaxPods stays below 2^31
constexpr int32_t kGsMaxPods = 1 << 20;

constexpr uint32_t gs_stage_bytes(int threads) { return (uint32_t)threads * kGsStageBytesPerThread; }
constexpr int gs_stages(int threads) { return threads >= 512 ? 2 : 3; }

template <typename E, int SP, int IPMODE, bool K1, int kGsThreads>
__global__ void __launch_bounds__(kGsThreads)
replay_grouped_sorted_kernel(const __grid_constant__ ReplayParams P) {
    constexpr int kGsStages = gs_stages(kGsThreads);          // tile buffers; kGsStages - 1 tiles are in flight
    constexpr uint32_t kGsStageBytes = gs_stage_bytes(kGsThreads);   // per stage and array
    constexpr uint32_t kEPP = 16 / sizeof(E);
    constexpr uint32_t kStageVec = kGsStageBytes / 16;
    constexpr uint32_t kArrays = IPMODE == 2 ? 2u : 1u;
    constexpr bool kCheck = sizeof(E) == 4;     // a uint16 element is neither negative nor >= 2^16
    extern __shared__ int4 gsm[];               // [stages][kArrays][kStageVec]; IPMODE 1: + int32 [kStageVec * kEPP]
    __shared__ uint64_t bars[kGsStages];
    __shared__ uint32_t s_cnt[33];
    __shared__ uint16_t s_perm[kGsThreads];

    const uint32_t tid = threadIdx.x;
    const uint32_t ny = (P.pol_count + SP - 1) / SP;
    const uint32_t pbase = (blockIdx.x % ny) * SP;
    const uint32_t bx = blockIdx.x / ny;
    const uint64_t rb = (uint64_t)bx * kGsThreads;                       // first resource of the block
    const uint32_t nres = (uint32_t)min((uint64_t)kGsThreads, P.R - rb);
    const uint32_t T = P.T;
    const uint64_t Q = P.Q;

    // the block's run of the queue axis, widened to whole 16-byte pieces
    const uint64_t qa = P.group_off[rb], qb = P.group_off[rb + nres];
    bool bad = !(qa <= qb && qb <= Q);
    const uint64_t a0 = qa & ~(uint64_t)(kEPP - 1);
    const uint64_t a1 = (qb + kEPP - 1) & ~(uint64_t)(kEPP - 1);          // <= Q: Q is a multiple of kEPP
    const uint64_t row_vec64 = bad ? 0 : (a1 - a0) / kEPP;                // int4 per tick
    if (row_vec64 > kStageVec) bad = true;

    // counting sort of the block's resources by group size
    uint32_t key = 0, pos = 0;
    if (tid < nres) {
        const uint64_t g0 = P.group_off[rb + tid], g1 = P.group_off[rb + tid + 1];
        if (!(g0 <= g1 && qa <= g0 && g1 <= qb) || g1 - g0 >= kGsMaxOwn) bad = true;
        else key = (uint32_t)min(g1 - g0, (uint64_t)32);
    }
    if (tid < 33) s_cnt[tid] = 0u;
    s_perm[tid] = 0xFFFFu;
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < kGsStages; ++i) mbar_init(&bars[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid < nres) pos = atomicAdd(&s_cnt[key], 1u);
    __syncthreads();
    if (tid < 32) {
        const uint32_t c = s_cnt[tid];
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (tid >= (uint32_t)o) incl += v;
        }
        const uint32_t tot = __shfl_sync(0xFFFFFFFFu, incl, 31);
        s_cnt[tid] = incl - c;            // first place of size class `tid`
        if (tid == 0) s_cnt[32] = tot;    // ... and of the class "32 queues or more"
    }
    __syncthreads();
    if (tid < nres) {
        const uint32_t p = s_cnt[key] + pos;
        const uint32_t rot = (bx * 2654435761u) >> 27;                   // a pseudo-random turn of the warps
        s_perm[((((p >> 5) + rot) & (kGsThreads / 32 - 1)) << 5) | (p & 31u)] = (uint16_t)tid;
    }
    // the per-queue in-progress counts of the run, once (IPMODE 1)
    int32_t *s_ipq = reinterpret_cast<int32_t *>(gsm + kGsStages * kArrays * kStageVec);
    if constexpr (IPMODE == 1) {
        if (row_vec64 <= kStageVec) {
            const uint32_t row_elems0 = (uint32_t)row_vec64 * kEPP;
            const int32_t *iq = static_cast<const int32_t *>(P.inprog) + a0;
            for (uint32_t i = tid; i < row_elems0; i += kGsThreads) {
                const int32_t v = __ldg(iq + i);
                s_ipq[i] = v;
                if ((uint32_t)v >= kSmallTallyTerm && a0 + i >= qa && a0 + i < qb) bad = true;
            }
        }
    }
    const bool tiled = !__syncthreads_or(bad ? 1 : 0);                   // uniform over the block

    const uint32_t mine = s_perm[tid];
    const bool valid = tiled ? mine != 0xFFFFu : tid < nres;
    const uint64_t r = rb + (tiled ? (valid ? mine : 0u) : tid);
    const uint64_t q0 = valid ? P.group_off[r] : 0, q1 = valid ? P.group_off[r + 1] : 0;

    PolDev pol[SP];
    bool live[SP];
    KeepTerm keep[SP];
#pragma unroll
    for (int s = 0; s < SP; ++s) {
        live[s] = pbase + s < P.pol_count;
        pol[s] = P.pol[live[s] ? pbase + s : P.pol_count - 1];
        keep[s] = keep_term(pol[s]);
    }
    int32_t cur[SP];
    long long pt[SP];
    uint32_t np[SP];
#pragma unroll
    for (int s = 0; s < SP; ++s) { cur[s] = 0; pt[s] = 0; np[s] = 0u; }
    if (tiled) {
        const uint32_t row_vec = (uint32_t)row_vec64;
        const uint32_t row_elems = row_vec * kEPP;
        const uint32_t row_bytes = row_vec * 16u;
        const uint32_t tile_ticks = row_vec == 0 ? kGsMaxTileTicks : min(kGsMaxTileTicks, kStageVec / row_vec);
        const uint32_t n_tiles = (T + tile_ticks - 1) / tile_ticks;
        const uint32_t n_own = (uint32_t)(q1 - q0);
        const uint32_t j0 = valid ? (uint32_t)(q0 - a0) : 0u;             // place inside the staged run
        // the lanes of a warp own groups of (nearly) one size: the loop over queues is unrolled for
        // the largest of them, lanes with fewer queues sit out the last rounds under a predicate
        const uint32_t n_max = __reduce_max_sync(0xFFFFFFFFu, valid ? n_own : 0u);
        // one bulk copy per tick row, issued by the lanes of warp 0 side by side (a single thread
        // issuing them one after another holds its warp, and with it the block's barrier, back)
        auto stage = [&](uint32_t k) {
            if (tid < 32 && k < n_tiles) {
                const uint32_t t0 = k * tile_ticks;
                const uint32_t nt = min(tile_ticks, T - t0);
                uint64_t *bar = &bars[k % kGsStages];
                int4 *dl = gsm + (k % kGsStages) * kArrays * kStageVec;
                int4 *di = dl + kStageVec;
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");       // earlier reads of this buffer are done
                if (tid == 0) mbar_expect_tx(bar, nt * row_bytes * kArrays);
                __syncwarp();
                if (row_bytes) {
                    for (uint32_t t = tid; t < nt; t += 32) {
                        const uint64_t g = (uint64_t)(t0 + t) * Q + a0;
                        bulk_copy_g2s(dl + t * row_vec, static_cast<const E *>(P.llen) + g, row_bytes, bar);
                        if (IPMODE == 2) bulk_copy_g2s(di + t * row_vec, static_cast<const E *>(P.inprog) + g, row_bytes, bar);
                    }
                }
            }
        };
#pragma unroll
        for (int i = 0; i < kGsStages - 1; ++i) stage((uint32_t)i);
        for (uint32_t k = 0; k < n_tiles; ++k) {
            mbar_wait(&bars[k % kGsStages], (k / kGsStages) & 1u);       // tile k has landed
            __syncthreads();                                              // all are done with tile k-1
            stage(k + kGsStages - 1);                                     // refill the buffer tile k-1 occupied
            if (!valid) continue;
            const uint32_t t0 = k * tile_ticks;
            const uint32_t nt = min(tile_ticks, T - t0);
            const E *tl = reinterpret_cast<const E *>(gsm + (k % kGsStages) * kArrays * kStageVec) + j0;
            const E *ti = tl + (size_t)kStageVec * kEPP;
            const int32_t *iq = s_ipq + j0;
            // the exact form of one tick (a count outside [0, 2^16) was seen)
            auto slow_tick = [&](const E *lrow, const E *irow, int32_t (&want)[SP]) {
                long long sum[SP];
#pragma unroll
                for (int s = 0; s < SP; ++s) sum[s] = 0;
                for (uint32_t j = 0; j < n_own; ++j) {
                    const int32_t l = (int32_t)lrow[j];
                    const int32_t ip = IPMODE == 2 ? (int32_t)irow[j] : (IPMODE == 1 ? iq[j] : 0);
#pragma unroll
                    for (int s = 0; s < SP; ++s) sum[s] += decide_exact(l, ip, pol[s], cur[s]);   // autoscaler.py:255-257
                }
#pragma unroll
                for (int s = 0; s < SP; ++s) want[s] = clip64(sum[s], pol[s].mn, pol[s].mx, cur[s]);   // :259
            };
            uint32_t own[8];                 // own[j] = 1 when queue j of the unrolled loop is this resource's
#pragma unroll
            for (int j = 0; j < 8; ++j) own[j] = (uint32_t)j < n_own ? (uint32_t)P.one : 0u;
            auto run_tile = [&](auto ntag) {
                constexpr int N = decltype(ntag)::value;                  // queues per resource, 0: not known here
                for (uint32_t t = 0; t < nt; ++t) {
                    const E *lrow = tl + (size_t)t * row_elems;
                    const E *irow = ti + (size_t)t * row_elems;
                    uint32_t sum[SP];
#pragma unroll
                    for (int s = 0; s < SP; ++s) sum[s] = 0u;
                    uint32_t any = 0u;
                    // `live_q`: the queue is this resource's own (a lane with fewer queues than the
                    // warp's largest group reads on into its neighbour's and adds nothing)
                    // (the add is a multiply-add by 0 or 1: the integer ALU pipe, which runs at half the
                    // FMA pipe's rate, is what this kernel fills first)
                    auto one_queue = [&](uint32_t j, uint32_t live_q) {
                        uint32_t tally = (uint32_t)lrow[j];
                        if constexpr (kCheck) any |= tally;
                        if constexpr (IPMODE == 1) tally += (uint32_t)iq[j];          // range-checked when staged
                        if constexpr (IPMODE == 2) {
                            const uint32_t ip = (uint32_t)irow[j];
                            if constexpr (kCheck) any |= ip;
                            tally += ip;
                        }
#pragma unroll
                        for (int s = 0; s < SP; ++s) {
                            const uint32_t q = K1 ? tally : quotient_small(tally + 1u, pol[s]);
                            const uint32_t v = (uint32_t)finish_simple3(q, __viaddmin_u32(q, keep[s].add, keep[s].lim),
                                                                        pol[s], cur[s]);                     // :255-257
                            sum[s] = v * live_q + sum[s];
                        }
                    };
                    if constexpr (N > 0) {
#pragma unroll
                        for (int j = 0; j < N; ++j) one_queue((uint32_t)j, own[j < 8 ? j : 0]);
                    } else {
                        for (uint32_t j = 0; j < n_own; ++j) one_queue(j, 1u);
                    }
                    // (checking once per tile instead -- fast code for the whole tile, then a replay of
                    // the tile from saved state if a count was out of range -- measured 32 % slower for
                    // one policy over int32, 1.72 against 1.30 ms: the saved state costs registers)
                    int32_t want[SP];
                    if (kCheck && any >= kSmallTallyTerm) {
                        slow_tick(lrow, irow, want);
                    } else {
#pragma unroll
                        for (int s = 0; s < SP; ++s)                                                        // :259
                            want[s] = finish_simple3(sum[s], __viaddmin_u32(sum[s], keep[s].add, keep[s].lim), pol[s], cur[s]);
                    }
#pragma unroll
                    for (int s = 0; s < SP; ++s) {
                        count_if_ne(np[s], want[s], cur[s]);                                                // :227
                        cur[s] = want[s];
                        pt[s] += want[s];
                    }
                }
            };
            if (SP == 1 || (SP == 2 && sizeof(E) == 2)) {
                switch (n_max) {
                    case 1: run_tile(std::integral_constant<int, 1>{}); break;
                    case 2: run_tile(std::integral_constant<int, 2>{}); break;
                    case 3: run_tile(std::integral_constant<int, 3>{}); break;
                    case 4: run_tile(std::integral_constant<int, 4>{}); break;
                    case 5: run_tile(std::integral_constant<int, 5>{}); break;
                    case 6: run_tile(std::integral_constant<int, 6>{}); break;
                    case 7: run_tile(std::integral_constant<int, 7>{}); break;
                    case 8: run_tile(std::integral_constant<int, 8>{}); break;
                    default: run_tile(std::integral_constant<int, 0>{}); break;
                }
            } else {
                run_tile(std::integral_constant<int, 0>{});
            }
        }
    } else if (valid) {
        // element-wise loads, exact arithmetic: any offsets, any group size
        const int32_t *iqg = static_cast<const int32_t *>(P.inprog);
        for (uint32_t t = 0; t < T; ++t) {
            const E *lrow = static_cast<const E *>(P.llen) + (uint64_t)t * Q;
            const E *irow = static_cast<const E *>(P.inprog) + (uint64_t)t * Q;
            long long sum[SP];
#pragma unroll
            for (int s = 0; s < SP; ++s) sum[s] = 0;
            for (uint64_t q = q0; q < q1; ++q) {
                const int32_t l = ld_trace_cached<E>(lrow + q);
                const int32_t ip = IPMODE == 2 ? ld_trace_cached<E>(irow + q) : (IPMODE == 1 ? __ldg(iqg + q) : 0);
#pragma unroll
                for (int s = 0; s < SP; ++s) sum[s] += decide_exact(l, ip, pol[s], cur[s]);
            }
#pragma unroll
            for (int s = 0; s < SP; ++s) {
                const int32_t want = clip64(sum[s], pol[s].mn, pol[s].mx, cur[s]);
                count_if_ne(np[s], want, cur[s]);
                cur[s] = want;
                pt[s] += want;
            }
        }
    }
    if (valid && P.totals) {
#pragma unroll
        for (int s = 0; s < SP; ++s)
            if (live[s]) {
                kas_totals o;
                o.pod_ticks = pt[s];
                o.n_patches = np[s];
                o.final_current = cur[s];
                P.totals[r * P.S + P.pol_index[pbase + s]] = o;
            }
    }
    if (P.sums) {
#pragma unroll
        for (int s = 0; s < SP; ++s) {
            if (live[s]) {
                long long a = valid ? pt[s] : 0, b = valid ? (long long)np[s] : 0, c = valid ? (long long)cur[s] : 0;
                a = warp_sum_ll(a); b = warp_sum_ll(b); c = warp_sum_ll(c);
                if ((tid & 31u) == 0) {
                    unsigned long long *o = P.sums + 3ull * P.pol_index[pbase + s];
                    atomicAdd(o + 0, (unsigned long long)a);
                    atomicAdd(o + 1, (unsigned long long)b);
                    atomicAdd(o + 2, (unsigned long long)c);
                }
            }
        }
    }
}

}  // namespace kas
#include "kas_replay_scan.cuh"
namespace kas {

template <typename E, int VEC, int SP>
static void launch_identity(const ReplayParams &P, int mode, cudaStream_t stream) {
    const uint64_t threads = (P.R + VEC - 1) / VEC;
    const uint64_t nx = (threads + kReplayThreads - 1) / kReplayThreads;
    const uint32_t ny = (P.pol_count + SP - 1) / SP;
    const unsigned grid = (unsigned)(nx * ny);
#define KAS_LAUNCH_ID(IPM)                                                                        \
    do {                                                                                          \
        if (mode == kSimpleK1) replay_identity_kernel<E, VEC, SP, IPM, kSimpleK1><<<grid, kReplayThreads, 0, stream>>>(P); \
        else if (mode == kSimple) replay_identity_kernel<E, VEC, SP, IPM, kSimple><<<grid, kReplayThreads, 0, stream>>>(P); \
        else replay_identity_kernel<E, VEC, SP, IPM, kGeneric><<<grid, kReplayThreads, 0, stream>>>(P);     \
    } while (0)
    if (P.inprog_mode == 0) KAS_LAUNCH_ID(0);
    else if (P.inprog_mode == 1) KAS_LAUNCH_ID(1);
    else KAS_LAUNCH_ID(2);
#undef KAS_LAUNCH_ID
}

template <typename E>
static int launch_sweep(const ReplayParams &P, int mode, cudaStream_t stream) {
    const bool tickip = P.inprog_mode == 2;
    const unsigned chunks = (P.pol_count + kSweepSP - 1) / kSweepSP;          // <= 16
    const unsigned groups = kSweepMaxWarps / chunks;                           // >= 1
    const unsigned warps = chunks * groups;
    const uint64_t cols_per_block = (uint64_t)groups * kSweepCols;
    const unsigned grid = (unsigned)((P.R + cols_per_block - 1) / cols_per_block);
    const size_t smem = (size_t)(tickip ? 2 : 1) * kSweepStages * kSweepTileRows * kSweepCols * sizeof(E);
#define KAS_LAUNCH_SWEEP(TI, MD, AC)                                                               \
    do {                                                                                           \
        auto kern = replay_sweep_kernel<E, TI, MD, AC>;                                            \
        /* dynamic tiles + the static mbarriers can pass the 48 KB default */                      \
        KAS_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,      \
                                          (int)smem));                                             \
        kern<<<grid, 32 * warps, smem, stream>>>(P);                                               \
    } while (0)
    // state width (SweepAcc): what the policies of this launch and the absence of cur0 allow
    int acc = kAccWide;
    if (mode != kGeneric && P.cur0 == nullptr) {
        bool narrow = true, packed = true;
        for (uint32_t s = 0; s < P.pol_count; ++s) {
            narrow = narrow && P.pol[s].mx < kSweepNarrowBound;
            packed = packed && P.pol[s].mn >= 0 && P.pol[s].mx <= kSweepPackedMaxPods && P.pol[s].m32 != 0u;
        }
        acc = packed ? kAccPacked : (narrow ? kAccNarrow : kAccWide);
    }
    if (mode == kGeneric) {
        if (tickip) KAS_LAUNCH_SWEEP(true, kGeneric, kAccWide); else KAS_LAUNCH_SWEEP(false, kGeneric, kAccWide);
    } else if (acc == kAccPacked) {
        if (tickip) KAS_LAUNCH_SWEEP(true, kSimple, kAccPacked); else KAS_LAUNCH_SWEEP(false, kSimple, kAccPacked);
    } else if (acc == kAccNarrow) {
        if (tickip) KAS_LAUNCH_SWEEP(true, kSimple, kAccNarrow); else KAS_LAUNCH_SWEEP(false, kSimple, kAccNarrow);
    } else {
        if (tickip) KAS_LAUNCH_SWEEP(true, kSimple, kAccWide); else KAS_LAUNCH_SWEEP(false, kSimple, kAccWide);
    }
#undef KAS_LAUNCH_SWEEP
    return KAS_OK;
}

template <typename E>
static void launch_small(const ReplayParams &P, int mode, cudaStream_t stream) {
    if (mode != kGeneric) replay_small_kernel<E, kSimple><<<(unsigned)P.R, 32, 0, stream>>>(P);
    else replay_small_kernel<E, kGeneric><<<(unsigned)P.R, 32, 0, stream>>>(P);
}

template <typename E, int SP>
static int launch_grouped(const ReplayParams &P, int mode, cudaStream_t stream) {
    const uint64_t nx = (P.R + kReplayThreads - 1) / kReplayThreads;
    const uint32_t ny = (P.pol_count + SP - 1) / SP;
    const unsigned grid = (unsigned)(nx * ny);
    if (P.startup_delay) {
        if (mode != kGeneric) replay_grouped_kernel<E, SP, kSimple, true><<<grid, kReplayThreads, 0, stream>>>(P);
        else replay_grouped_kernel<E, SP, kGeneric, true><<<grid, kReplayThreads, 0, stream>>>(P);
        return KAS_OK;
    }
    // rows that start on 16-byte boundaries and real groups: the shared-memory tiled kernel
    constexpr uint64_t kEPP = 16 / sizeof(E);
    const bool tickip = P.inprog_mode == 2;
    const bool tileable = P.group_off != nullptr && P.Q % kEPP == 0 && aligned(P.llen, 16) &&
                          (!tickip || aligned(P.inprog, 16));
    // realistic policies and nothing exotic asked for: the lean kernel (sorted threads, TMA tiles)
    bool lean = tileable && mode != kGeneric && !P.cur0 && !P.patch_fail && !P.dense;
    for (uint32_t s = 0; s < P.pol_count && lean; ++s)
        lean = P.pol[s].m32 != 0u && P.pol[s].mx < kGsMaxPods;
#ifdef KAS_NO_LEAN_GROUPED
    lean = false;
#endif
    if (lean) {
        const bool k1 = SP == 1 && P.pol[0].k1 != 0u;
        const bool big = P.R >= (uint64_t)kGsBigThreads * 148u;
#define KAS_LAUNCH_GS(IPM, KONE, TH)                                                               \
    do {                                                                                           \
        /* (+ 64 bytes: a lane may read up to seven elements past the last staged row) */          \
        const size_t smem = (size_t)gs_stages(TH) * (tickip ? 2 : 1) * gs_stage_bytes(TH) +       \
                            (P.inprog_mode == 1 ? (size_t)gs_stage_bytes(TH) / sizeof(E) * 4 : 0) + 64; \
        const unsigned gs_grid = (unsigned)(((P.R + TH - 1) / TH) * ny);                          \
        auto kern = replay_grouped_sorted_kernel<E, SP, IPM, KONE, TH>;                            \
        KAS_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,      \
                                          (int)smem));                                             \
        kern<<<gs_grid, TH, smem, stream>>>(P);                                                    \
    } while (0)
#define KAS_LAUNCH_GS_T(IPM, KONE)                                                                 \
    do {                                                                                           \
        if (big) KAS_LAUNCH_GS(IPM, KONE, kGsBigThreads); else KAS_LAUNCH_GS(IPM, KONE, kGsSmallThreads); \
    } while (0)
        if constexpr (SP == 1) {
            if (k1) {
                if (P.inprog_mode == 0) KAS_LAUNCH_GS_T(0, true);
                else if (P.inprog_mode == 1) KAS_LAUNCH_GS_T(1, true);
                else KAS_LAUNCH_GS_T(2, true);
                return KAS_OK;
            }
        }
        if (P.inprog_mode == 0) KAS_LAUNCH_GS_T(0, false);
        else if (P.inprog_mode == 1) KAS_LAUNCH_GS_T(1, false);
        else KAS_LAUNCH_GS_T(2, false);
#undef KAS_LAUNCH_GS_T
#undef KAS_LAUNCH_GS
        return KAS_OK;
    }
    if (tileable) {
        const size_t smem = (size_t)kGroupedStages * (tickip ? 2 : 1) * kGroupedTileBytes;
#define KAS_LAUNCH_GT(MD, TI)                                                                      \
    do {                                                                                           \
        auto kern = replay_grouped_tiled_kernel<E, SP, MD, TI>;                                    \
        KAS_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,      \
                                          (int)smem));                                             \
        kern<<<grid, kReplayThreads, smem, stream>>>(P);                                           \
    } while (0)
        if (mode != kGeneric) {
            if (tickip) KAS_LAUNCH_GT(kSimple, true); else KAS_LAUNCH_GT(kSimple, false);
        } else {
            if (tickip) KAS_LAUNCH_GT(kGeneric, true); else KAS_LAUNCH_GT(kGeneric, false);
        }
#undef KAS_LAUNCH_GT
        return KAS_OK;
    }
    if (mode != kGeneric) replay_grouped_kernel<E, SP, kSimple, false><<<grid, kReplayThreads, 0, stream>>>(P);
    else replay_grouped_kernel<E, SP, kGeneric, false><<<grid, kReplayThreads, 0, stream>>>(P);
    return KAS_OK;
}

}  // namespace kas

using namespace kas;

// Everything after argument checking, for one element type of the trace.
template <typename E>
static int replay_typed(const kas_replay_args *a, int mode, cudaStream_t stream) {
    constexpr int kVec = 16 / (int)sizeof(E);                 // elements per 128-bit load: 4 or 8
    const bool identity = a->d_group_off == nullptr;
    const bool tickip = a->inprog_mode == 2;
    const bool vec = identity && (a->n_queues % kVec == 0) && aligned(a->d_llen, 16) &&
                     (!tickip || aligned(a->d_inprog, 16)) &&
                     (!a->d_dense || aligned(a->d_dense, 16)) && a->n_resources >= 4096;

    // few resources: a warp per resource with its queues staged in shared memory
    // (needs every resource's queues of one tick to fit a tile)
    const bool delayed = a->startup_delay != 0;     // the availability model lives in the grouped kernel
    const bool small = !delayed && a->n_resources <= kSmallMaxResources &&
                       (identity || a->n_queues <= (uint64_t)kSmallTileElems);
    // ... and when each of them owns ONE queue and the trace is long, the ticks themselves are
    // replayed in parallel (kas_replay_scan.cuh)
    const bool scan = small && identity && mode != kGeneric && !a->d_patch_fail &&
                      a->n_ticks >= kScanReplayMinTicks;
    // ... any other small fleet with a long trace: the few possible states are enumerated
    // (it multiplies the work by the number of states to divide the depth by 256: only worth it
    // while resources x policies cannot fill the machine with sequential chains)
    const bool fsm = small && !scan && a->n_ticks >= kScanReplayMinTicks &&
                     a->n_resources * (uint64_t)a->n_policies <= 2048;
    // Slots of a launch in the caller's order, except under a sweep of single-queue
    // resources: there policies are grouped by what a warp's two policies can share
    // (keys_per_pod -> the quotient; min > 0 and max > 0 -> the keep factor), and a group
    // of odd size is followed by an empty slot so that no pair straddles two groups.
    // Results are written at the caller's indices either way.
    constexpr uint32_t kEmptySlot = 0xFFFFu;           // n_policies <= 65535, so never an index
    std::vector<uint32_t> slots(a->n_policies);
    for (uint32_t i = 0; i < a->n_policies; ++i) slots[i] = i;
    const bool sweep_path = !small && !delayed && identity && vec && a->n_policies > 1 && !a->d_patch_fail;
    if (sweep_path) {
        auto group_less = [&](uint32_t x, uint32_t y) {
            const kas_policy &px = a->policies[x], &py = a->policies[y];
            if (px.keys_per_pod != py.keys_per_pod) return px.keys_per_pod < py.keys_per_pod;
            if ((px.min_pods > 0) != (py.min_pods > 0)) return py.min_pods > 0;
            return (px.max_pods > 0) && !(py.max_pods > 0);
        };
        std::vector<uint32_t> order(slots);
        std::stable_sort(order.begin(), order.end(), group_less);
        slots.clear();
        for (size_t i = 0; i < order.size(); ++i) {
            const bool new_group = i > 0 && group_less(order[i - 1], order[i]);
            if (new_group && (slots.size() & 1u)) slots.push_back(kEmptySlot);
            slots.push_back(order[i]);
        }
    }
    for (size_t first = 0; first < slots.size(); first += kMaxPoliciesPerLaunch) {
        ReplayParams P;
        P.llen = a->d_llen;
        P.inprog = a->d_inprog;
        P.group_off = a->d_group_off;
        P.cur0 = a->d_cur0;
        P.cur0_rstride = 1;
        P.cur0_sstride = 0;
        if (a->d_resume) {       // start from the final pods of an earlier replay's totals [R][S]
            P.cur0 = reinterpret_cast<const int32_t *>(a->d_resume) + 3;      // kas_totals.final_current
            P.cur0_rstride = 4 * a->n_policies;
            P.cur0_sstride = 4;
        }
        P.patch_fail = a->d_patch_fail;
        P.totals = a->d_totals;
        P.sums = reinterpret_cast<unsigned long long *>(a->d_sums);
        P.dense = a->d_dense;
        P.Q = a->n_queues;
        P.R = a->n_resources;
        P.T = a->n_ticks;
        P.S = a->n_policies;
        P.pol_first = (uint32_t)first;
        P.pol_count = (uint32_t)std::min<size_t>(slots.size() - first, kMaxPoliciesPerLaunch);
        P.inprog_mode = a->inprog_mode;
        P.startup_delay = a->startup_delay;
        P.one = 1;
        for (uint32_t s = 0; s < (uint32_t)kMaxPoliciesPerLaunch; ++s) {
            PolDev d = {0, 0, 0, 0, 0u, 0u};
            uint32_t idx = kEmptySlot;
            if (s < P.pol_count) {
                idx = slots[first + s];
                // an empty slot repeats its partner (always the slot before it)
                d = make_poldev(a->policies[idx != kEmptySlot ? idx : slots[first + s - 1]]);
            }
            P.pol[s] = d;
            P.pol_index[s] = (uint16_t)idx;
        }
        if (scan) {
            replay_scan_kernel<E><<<(unsigned)(P.R * P.pol_count), kScanReplayThreads, 0, stream>>>(P);
        } else if (fsm) {
            const size_t smem = (size_t)kFsmThreads * 256;
            KAS_CUDA_TRY(cudaFuncSetAttribute(replay_fsm_kernel<E>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            replay_fsm_kernel<E><<<(unsigned)(P.R * P.pol_count), kFsmThreads, smem, stream>>>(P);
        } else if (small) {
            launch_small<E>(P, mode, stream);
        } else if (identity && !a->d_patch_fail && !delayed) {
            if (P.pol_count == 1) {
                const PolDev &pd = P.pol[0];
                const bool packable = sizeof(E) == 2 && vec && mode != kGeneric && a->inprog_mode == 0 &&
                                      !P.cur0 && pd.mn >= 0 && pd.mx <= kPackedMaxPods && pd.m32 != 0u;
                if (packable) {
                    const unsigned grid = (unsigned)((P.R / 8 + kReplayThreads - 1) / kReplayThreads);
                    if (pd.k1) replay_packed16_kernel<true><<<grid, kReplayThreads, 0, stream>>>(P);
                    else replay_packed16_kernel<false><<<grid, kReplayThreads, 0, stream>>>(P);
                } else if (vec) launch_identity<E, kVec, 1>(P, mode, stream);
                else launch_identity<E, 1, 1>(P, mode, stream);
            } else {
                if (vec) {
                    const int st = launch_sweep<E>(P, mode, stream);
                    if (st != KAS_OK) return st;
                } else {
                    launch_identity<E, 1, 4>(P, mode, stream);
                }
            }
        } else {
            // several policies: four per thread over an int32 trace (fewer passes over HBM), two per
            // thread with the queue loop unrolled over a uint16 trace (half the bytes per pass; 36.0
            // against 51.7 ms for 41 policies x 2e9 queue-ticks; for int32 the same costs 59 against 54.5)
            constexpr int kMultiSP = sizeof(E) == 2 ? 2 : 4;
            const int st = P.pol_count == 1 ? launch_grouped<E, 1>(P, mode, stream)
                                            : launch_grouped<E, kMultiSP>(P, mode, stream);
            if (st != KAS_OK) return st;
        }
        KAS_CUDA_TRY(cudaGetLastError());
    }
    return KAS_OK;
}

extern "C" {


int kas_replay(const kas_replay_args *a, void *stream_v) {
    if (!a || !a->policies || a->n_policies == 0 || a->n_policies > 65535) return KAS_ERR_INVALID;
    if (a->n_queues == 0 || a->n_resources == 0 || a->n_ticks == 0 || !a->d_llen) return KAS_ERR_INVALID;
    if (a->inprog_mode < 0 || a->inprog_mode > 2) return KAS_ERR_INVALID;
    if (a->inprog_mode != 0 && !a->d_inprog) return KAS_ERR_INVALID;
    if (a->trace_dtype != KAS_TRACE_INT32 && a->trace_dtype != KAS_TRACE_UINT16) return KAS_ERR_INVALID;
    if (a->startup_delay > KAS_MAX_STARTUP_DELAY) return KAS_ERR_INVALID;
    if (!a->d_group_off && a->n_resources != a->n_queues) return KAS_ERR_INVALID;
    if (!a->d_totals && !a->d_sums && !a->d_dense) return KAS_ERR_INVALID;
    if (a->d_resume && (a->d_cur0 || a->startup_delay)) return KAS_ERR_INVALID;   // one source of starting pods; the
                                                                                   // availability window is not carried
    if (a->d_resume && a->d_resume == a->d_totals) return KAS_ERR_INVALID;        // read while being written
    if (!aligned(a->d_resume, 8)) return KAS_ERR_ALIGNMENT;
    if (a->n_resources > (0x7FFFFFFFull / kMaxPoliciesPerLaunch) * kReplayThreads) return KAS_ERR_UNSUPPORTED;
    const size_t elem = a->trace_dtype == KAS_TRACE_UINT16 ? 2 : 4;
    if (!aligned(a->d_llen, elem) || !aligned(a->d_inprog, a->inprog_mode == 2 ? elem : 4) ||
        !aligned(a->d_dense, 4) || !aligned(a->d_totals, 8) || !aligned(a->d_sums, 8) ||
        !aligned(a->d_group_off, 8) || !aligned(a->d_cur0, 4))
        return KAS_ERR_ALIGNMENT;
    const int mode = replay_mode_for(a->policies, a->n_policies);
    if (mode < 0) return KAS_ERR_ZERO_DIVISION;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (a->d_sums)
        KAS_CUDA_TRY(cudaMemsetAsync(a->d_sums, 0, sizeof(int64_t) * 3 * a->n_policies, stream));
    return a->trace_dtype == KAS_TRACE_UINT16 ? replay_typed<uint16_t>(a, mode, stream)
                                              : replay_typed<int32_t>(a, mode, stream);
}

int kas_totals_chain(kas_totals *d_acc, const kas_totals *d_chunk, uint64_t n_resources, uint32_t n_policies,
                     int64_t *d_sums_acc, const int64_t *d_sums_chunk, void *stream_v) {
    if (!d_acc || !d_chunk || n_policies == 0 || (d_sums_acc != nullptr) != (d_sums_chunk != nullptr))
        return KAS_ERR_INVALID;
    if (n_policies > 256 && d_sums_acc) return KAS_ERR_UNSUPPORTED;
    if (!aligned(d_acc, 8) || !aligned(d_chunk, 8) || !aligned(d_sums_acc, 8) || !aligned(d_sums_chunk, 8))
        return KAS_ERR_ALIGNMENT;
    const uint64_t n = n_resources * n_policies;
    if (n == 0) return KAS_OK;
    const int sms = sm_count();
    if (sms <= 0) return KAS_ERR_CUDA;
    uint64_t grid = (n + 255) / 256;
    if (grid > (uint64_t)sms * 16) grid = (uint64_t)sms * 16;
    totals_chain_kernel<<<(unsigned)grid, 256, 0, static_cast<cudaStream_t>(stream_v)>>>(
        d_acc, d_chunk, n, reinterpret_cast<long long *>(d_sums_acc),
        reinterpret_cast<const long long *>(d_sums_chunk), n_policies);
    KAS_CUDA_TRY(cudaGetLastError());
    return KAS_OK;
}

int kas_decide(const int64_t *d_a, const int64_t *d_kpp, const int64_t *d_mn, const int64_t *d_mx,
               const int64_t *d_cur, uint64_t n, int mode, int64_t *d_out, int32_t *d_status,
               void *stream_v) {
    if (mode != 0 && mode != 1) return KAS_ERR_INVALID;
    if (n && (!d_a || !d_mn || !d_mx || !d_cur || !d_out || !d_status)) return KAS_ERR_INVALID;
    if (n && mode == 0 && !d_kpp) return KAS_ERR_INVALID;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (d_status) KAS_CUDA_TRY(cudaMemsetAsync(d_status, 0, sizeof(int32_t), stream));
    if (n == 0) return KAS_OK;
    const int sms = sm_count();
    if (sms <= 0) return KAS_ERR_CUDA;
    uint64_t grid = (n + 255) / 256;
    if (grid > (uint64_t)sms * 8) grid = (uint64_t)sms * 8;
    decide_kernel<<<(unsigned)grid, 256, 0, stream>>>(
        reinterpret_cast<const long long *>(d_a), reinterpret_cast<const long long *>(d_kpp),
        reinterpret_cast<const long long *>(d_mn), reinterpret_cast<const long long *>(d_mx),
        reinterpret_cast<const long long *>(d_cur), n, mode, reinterpret_cast<long long *>(d_out), d_status);
    KAS_CUDA_TRY(cudaGetLastError());
    return KAS_OK;
}

int kas_synth_trace(uint64_t seed, uint32_t n_ticks, uint64_t q_first, uint64_t n_q, int32_t *d_out,
                    void *stream_v) {
    if (!d_out && n_ticks && n_q) return KAS_ERR_INVALID;
    if (n_ticks == 0 || n_q == 0) return KAS_OK;
    const int sms = sm_count();
    if (sms <= 0) return KAS_ERR_CUDA;
    const uint64_t total = (uint64_t)n_ticks * n_q;
    uint64_t grid = (total + 255) / 256;
    if (grid > (uint64_t)sms * 16) grid = (uint64_t)sms * 16;
    synth_trace_kernel<<<(unsigned)grid, 256, 0, static_cast<cudaStream_t>(stream_v)>>>(seed, n_ticks, q_first,
                                                                                       n_q, d_out);
    KAS_CUDA_TRY(cudaGetLastError());
    return KAS_OK;
}

}  // extern "C"
