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
#include "kas_common.cuh"
#include "kas_decide.cuh"

#include <limits.h>
#include <type_traits>

namespace kas {

constexpr int kReplayThreads = 128;
constexpr int kMaxPoliciesPerLaunch = 32;
constexpr int kTickUnroll = 4;

struct ReplayParams {
    const int32_t *llen;
    const int32_t *inprog;
    const uint64_t *group_off;
    const int32_t *cur0;
    kas_totals *totals;
    unsigned long long *sums;
    int32_t *dense;
    uint64_t Q, R;
    uint32_t T;
    uint32_t S;            // total policies (row stride of totals)
    uint32_t pol_first;    // global index of pol[0]
    uint32_t pol_count;    // policies in this launch
    int32_t inprog_mode;
    int32_t one;           // the constant 1, opaque to the compiler (see pod-tick accumulation)
    PolDev pol[kMaxPoliciesPerLaunch];
};

__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}

// ---- every queue is its own resource (R == Q) ------------------------------
// IPMODE 0: no in-progress term, 1: one value per queue, 2: one per (tick, queue).
// Blocks are numbered policy-chunk fastest, so the chunks that read the same
// trace columns run together and share them through L2.
#ifndef KAS_REPLAY_SWEEP_SP
#define KAS_REPLAY_SWEEP_SP 2     // policies per thread under a sweep
#endif
#ifndef KAS_REPLAY_MINB_S1
#define KAS_REPLAY_MINB_S1 6      // resident blocks per SM (register budget), one policy per thread
#endif
#ifndef KAS_REPLAY_MINB_SWEEP
#define KAS_REPLAY_MINB_SWEEP 4   // the same with several policies per thread
#endif
template <int VEC, int SP, int IPMODE, int MODE>
__global__ void __launch_bounds__(kReplayThreads, (SP == 1 ? KAS_REPLAY_MINB_S1 : KAS_REPLAY_MINB_SWEEP))
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
        const int32_t c0 = (valid && P.cur0) ? P.cur0[r0 + v] : 0;
        ipq[v] = (valid && IPMODE == 1) ? P.inprog[r0 + v] : 0;
#pragma unroll
        for (int s = 0; s < SP; ++s) { cur[s][v] = c0; pt[s][v] = 0; np[s][v] = 0u; }
    }

    if (valid) {
        const int32_t *lrow = P.llen + r0;
        const int32_t *irow = IPMODE == 2 ? P.inprog + r0 : nullptr;
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
                    int32_t *o = P.dense + ((uint64_t)(P.pol_first + pbase + s) * T + t) * P.R + r0;
                    if constexpr (VEC == 4)
                        *reinterpret_cast<int4 *>(o) = make_int4(cur[s][0], cur[s][1], cur[s][2], cur[s][3]);
                    else
                        o[0] = cur[s][0];
                }
            }
        };
        // rows t .. t+U-1 (clamped to the last tick) into registers
        auto load = [&](uint32_t t, int32_t (&l)[U][VEC], int32_t (&ip)[U][VEC]) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t tt = min(t + (uint32_t)u, T - 1);
                const uint64_t o = (uint64_t)tt * Q;
                if constexpr (VEC == 4) {
                    const int4 a = ldg_stream_i4(lrow + o);
                    l[u][0] = a.x; l[u][1] = a.y; l[u][2] = a.z; l[u][3] = a.w;
                    if constexpr (IPMODE == 2) {
                        const int4 b = ldg_stream_i4(irow + o);
                        ip[u][0] = b.x; ip[u][1] = b.y; ip[u][2] = b.z; ip[u][3] = b.w;
                    }
                } else {
                    l[u][0] = ldg_stream_i32(lrow + o);
                    if constexpr (IPMODE == 2) ip[u][0] = ldg_stream_i32(irow + o);
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
                l[v] = ldg_stream_i32(lrow + o + v);
                ip[v] = IPMODE == 2 ? ldg_stream_i32(irow + o + v) : ipq[v];
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
                        P.totals[(r0 + v) * P.S + P.pol_first + pbase + s] = o;
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
                    unsigned long long *o = P.sums + 3ull * (P.pol_first + pbase + s);
                    atomicAdd(o + 0, (unsigned long long)a);
                    atomicAdd(o + 1, (unsigned long long)b);
                    atomicAdd(o + 2, (unsigned long long)c);
                }
            }
        }
    }
}

// ---- resources that own several queues (CSR groups) ------------------------
template <int SP, int MODE>
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
    const int32_t c0 = (valid && P.cur0) ? P.cur0[r] : 0;
#pragma unroll
    for (int s = 0; s < SP; ++s) { cur[s] = c0; pt[s] = 0; np[s] = 0u; }
    if (valid) {
        const uint64_t q0 = P.group_off ? P.group_off[r] : r;
        const uint64_t q1 = P.group_off ? P.group_off[r + 1] : r + 1;
        for (uint32_t t = 0; t < T; ++t) {
            const int32_t *lrow = P.llen + (uint64_t)t * Q;
            const int32_t *irow = P.inprog_mode == 2 ? P.inprog + (uint64_t)t * Q : P.inprog;
            long long sum[SP];
#pragma unroll
            for (int s = 0; s < SP; ++s) sum[s] = 0;
            for (uint64_t q = q0; q < q1; ++q) {
                const int32_t l = ldg_stream_i32(lrow + q);
                const int32_t ip = P.inprog_mode ? __ldg(irow + q) : 0;
#pragma unroll
                for (int s = 0; s < SP; ++s)
                    if (pbase + s < P.pol_count)
                        sum[s] += decide<MODE>(l, ip, P.pol[pbase + s], cur[s]);   // :255-257
            }
#pragma unroll
            for (int s = 0; s < SP; ++s) {
                if (pbase + s < P.pol_count) {
                    const PolDev &p = P.pol[pbase + s];
                    const int32_t want = clip64(sum[s], p.mn, p.mx, cur[s]);          // :259
                    np[s] += (want != cur[s]);
                    cur[s] = want;
                    pt[s] += want;
                    if (P.dense) P.dense[((uint64_t)(P.pol_first + pbase + s) * T + t) * P.R + r] = want;
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
                    o.final_current = cur[s];
                    P.totals[r * P.S + P.pol_first + pbase + s] = o;
                }
        }
    }
    if (P.sums) {
#pragma unroll
        for (int s = 0; s < SP; ++s) {
            if (pbase + s < P.pol_count) {
                long long a = valid ? pt[s] : 0, b = valid ? (long long)np[s] : 0, c = valid ? (long long)cur[s] : 0;
                a = warp_sum_ll(a); b = warp_sum_ll(b); c = warp_sum_ll(c);
                if ((threadIdx.x & 31) == 0) {
                    unsigned long long *o = P.sums + 3ull * (P.pol_first + pbase + s);
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

template <int MODE>
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
    int32_t cur = P.cur0 ? P.cur0[r] : 0;
    long long pt = 0;
    uint32_t np = 0;
    for (uint32_t t0 = 0; t0 < T; t0 += tile_ticks) {
        const uint32_t nt = min(tile_ticks, T - t0);
        const uint32_t elems = nt * nq;
        __syncwarp();
        for (uint32_t e = lane; e < elems; e += 32) {
            const uint32_t t = e / nq, j = e - t * nq;
            const uint64_t g = (uint64_t)(t0 + t) * P.Q + q0 + j;
            s_l[e] = ldg_stream_i32(P.llen + g);
            s_i[e] = P.inprog_mode == 2 ? ldg_stream_i32(P.inprog + g)
                                        : (P.inprog_mode == 1 ? __ldg(P.inprog + q0 + j) : 0);
        }
        __syncwarp();
        if (live) {
            for (uint32_t t = 0; t < nt; ++t) {
                long long sum = 0;
                for (uint32_t j = 0; j < nq; ++j)
                    sum += decide<MODE>(s_l[t * nq + j], s_i[t * nq + j], pol, cur);      // :255-257
                const int32_t want = clip64(sum, pol.mn, pol.mx, cur);                    // :259
                np += (want != cur);                                                      // :227
                cur = want;
                pt += want;
                if (P.dense) P.dense[((uint64_t)(P.pol_first + lane) * T + t0 + t) * P.R + r] = want;
            }
        }
    }
    if (live) {
        if (P.totals) {
            kas_totals o;
            o.pod_ticks = pt;
            o.n_patches = np;
            o.final_current = cur;
            P.totals[r * P.S + P.pol_first + lane] = o;
        }
        if (P.sums) {
            unsigned long long *o = P.sums + 3ull * (P.pol_first + lane);
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

template <int VEC, int SP>
static void launch_identity(const ReplayParams &P, int mode, cudaStream_t stream) {
    const uint64_t threads = (P.R + VEC - 1) / VEC;
    const uint64_t nx = (threads + kReplayThreads - 1) / kReplayThreads;
    const uint32_t ny = (P.pol_count + SP - 1) / SP;
    const unsigned grid = (unsigned)(nx * ny);
#define KAS_LAUNCH_ID(IPM)                                                                        \
    do {                                                                                          \
        if (mode == kSimpleK1) replay_identity_kernel<VEC, SP, IPM, kSimpleK1><<<grid, kReplayThreads, 0, stream>>>(P); \
        else if (mode == kSimple) replay_identity_kernel<VEC, SP, IPM, kSimple><<<grid, kReplayThreads, 0, stream>>>(P); \
        else replay_identity_kernel<VEC, SP, IPM, kGeneric><<<grid, kReplayThreads, 0, stream>>>(P);     \
    } while (0)
    if (P.inprog_mode == 0) KAS_LAUNCH_ID(0);
    else if (P.inprog_mode == 1) KAS_LAUNCH_ID(1);
    else KAS_LAUNCH_ID(2);
#undef KAS_LAUNCH_ID
}

static void launch_small(const ReplayParams &P, int mode, cudaStream_t stream) {
    if (mode != kGeneric) replay_small_kernel<kSimple><<<(unsigned)P.R, 32, 0, stream>>>(P);
    else replay_small_kernel<kGeneric><<<(unsigned)P.R, 32, 0, stream>>>(P);
}

template <int SP>
static void launch_grouped(const ReplayParams &P, int mode, cudaStream_t stream) {
    const uint64_t nx = (P.R + kReplayThreads - 1) / kReplayThreads;
    const uint32_t ny = (P.pol_count + SP - 1) / SP;
    const unsigned grid = (unsigned)(nx * ny);
    if (mode != kGeneric) replay_grouped_kernel<SP, kSimple><<<grid, kReplayThreads, 0, stream>>>(P);
    else replay_grouped_kernel<SP, kGeneric><<<grid, kReplayThreads, 0, stream>>>(P);
}

}  // namespace kas

using namespace kas;

extern "C" {

int kas_replay(const kas_replay_args *a, void *stream_v) {
    if (!a || !a->policies || a->n_policies == 0) return KAS_ERR_INVALID;
    if (a->n_queues == 0 || a->n_resources == 0) return KAS_ERR_INVALID;
    if (a->n_ticks && !a->d_llen) return KAS_ERR_INVALID;
    if (a->inprog_mode < 0 || a->inprog_mode > 2) return KAS_ERR_INVALID;
    if (a->inprog_mode != 0 && !a->d_inprog) return KAS_ERR_INVALID;
    if (!a->d_group_off && a->n_resources != a->n_queues) return KAS_ERR_INVALID;
    if (!a->d_totals && !a->d_sums && !a->d_dense) return KAS_ERR_INVALID;
    if (a->n_resources > (0x7FFFFFFFull / kMaxPoliciesPerLaunch) * kReplayThreads) return KAS_ERR_UNSUPPORTED;
    if (!aligned(a->d_llen, 4) || !aligned(a->d_inprog, 4) || !aligned(a->d_dense, 4) ||
        !aligned(a->d_totals, 8) || !aligned(a->d_sums, 8) || !aligned(a->d_group_off, 8) ||
        !aligned(a->d_cur0, 4))
        return KAS_ERR_ALIGNMENT;
    const int mode = replay_mode_for(a->policies, a->n_policies);
    if (mode < 0) return KAS_ERR_ZERO_DIVISION;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    if (a->d_sums)
        KAS_CUDA_TRY(cudaMemsetAsync(a->d_sums, 0, sizeof(int64_t) * 3 * a->n_policies, stream));

    const bool identity = a->d_group_off == nullptr;
    const bool tickip = a->inprog_mode == 2;
    if (a->n_ticks == 0) return KAS_ERR_INVALID;
    const bool vec4 = identity && (a->n_queues % 4 == 0) && aligned(a->d_llen, 16) &&
                      (!tickip || aligned(a->d_inprog, 16)) &&
                      (a->inprog_mode != 1 || aligned(a->d_inprog, 4)) &&
                      (!a->d_dense || aligned(a->d_dense, 16)) && a->n_resources >= 4096;

    // few resources: a warp per resource with its queues staged in shared memory
    // (needs every resource's queues of one tick to fit a tile)
    const bool small = a->n_resources <= kSmallMaxResources &&
                       (identity || a->n_queues <= (uint64_t)kSmallTileElems);
    for (uint32_t first = 0; first < a->n_policies; first += kMaxPoliciesPerLaunch) {
        ReplayParams P;
        P.llen = a->d_llen;
        P.inprog = a->d_inprog;
        P.group_off = a->d_group_off;
        P.cur0 = a->d_cur0;
        P.totals = a->d_totals;
        P.sums = reinterpret_cast<unsigned long long *>(a->d_sums);
        P.dense = a->d_dense;
        P.Q = a->n_queues;
        P.R = a->n_resources;
        P.T = a->n_ticks;
        P.S = a->n_policies;
        P.pol_first = first;
        P.pol_count = a->n_policies - first < (uint32_t)kMaxPoliciesPerLaunch
                          ? a->n_policies - first : (uint32_t)kMaxPoliciesPerLaunch;
        P.inprog_mode = a->inprog_mode;
        P.one = 1;
        for (uint32_t s = 0; s < (uint32_t)kMaxPoliciesPerLaunch; ++s) {
            PolDev d = {0, 0, 0, 0, 0u};
            if (s < P.pol_count) d = make_poldev(a->policies[first + s]);
            P.pol[s] = d;
        }
        if (small) {
            launch_small(P, mode, stream);
        } else if (identity) {
            if (P.pol_count == 1) {
                if (vec4) launch_identity<4, 1>(P, mode, stream);
                else launch_identity<1, 1>(P, mode, stream);
            } else {
                if (vec4) launch_identity<4, KAS_REPLAY_SWEEP_SP>(P, mode, stream);
                else launch_identity<1, 4>(P, mode, stream);
            }
        } else {
            if (P.pol_count == 1) launch_grouped<1>(P, mode, stream);
            else launch_grouped<4>(P, mode, stream);
        }
        KAS_CUDA_TRY(cudaGetLastError());
    }
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
