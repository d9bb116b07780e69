// kas_events.cu -- the patch-event stream of a replay.
//
// Stands where Autoscaler.scale_resource (autoscaler.py:221-242) would have
// called patch_namespaced_deployment / patch_namespaced_job: an event is a tick
// on which desired != current (:227), with the value before and the value
// patched ({'spec': {'replicas' | 'parallelism': to_pods}}, :232,:236).  Events
// come out grouped by resource and in tick order inside a resource, as a CSR
// (event_off[R+1], events[total]) -- the per-resource patch log.
//
// Three steps: count per resource (the replay recurrence, nothing written but
// the count), exclusive scan of the counts (tile sums in a scratch buffer),
// and the same recurrence again writing each event at its final place.  The
// order is fixed by the scan, not by atomics, so the stream is reproducible.
#include "kas_common.cuh"
#include "kas_decide.cuh"

#include <mutex>

namespace kas {

constexpr int kEventThreads = 128;
constexpr int kScanThreads = 256;
constexpr int kScanPerThread = 16;
constexpr int kScanTile = kScanThreads * kScanPerThread;   // 4096 counts per block

struct EventParams {
    const void *llen;       // trace elements: int32 or uint16 (template type E)
    const void *inprog;     // per-tick term: same type; per-queue term: int32 [Q]
    const uint64_t *group_off;
    const int32_t *cur0;             // pods before tick 0 of resource r: cur0[r * cur0_stride]
    uint64_t cur0_stride;            // 1, or 4 * S when it points into the kas_totals [R][S] of a resumed replay
    const uint8_t *patch_fail;       // [T][R] or null
    unsigned long long *event_off;   // [R+1]; counts live in [1..R] between the steps
    kas_patch_event *events;
    uint64_t capacity;
    uint64_t Q, R;
    uint32_t T;
    int32_t inprog_mode;
    uint32_t startup_delay;          // kas_replay_args.startup_delay
    bool u16;
    PolDev pol;
};

// One thread per resource; EMIT = false counts, EMIT = true writes.
template <typename E, int MODE, bool EMIT>
__global__ void __launch_bounds__(kEventThreads)
events_kernel(const __grid_constant__ EventParams P) {
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P.R) return;
    const PolDev pol = P.pol;
    const uint64_t q0 = P.group_off ? P.group_off[r] : r;
    const uint64_t q1 = P.group_off ? P.group_off[r + 1] : r + 1;
    int32_t cur = P.cur0 ? P.cur0[r * P.cur0_stride] : 0;     // what the controller reads
    int32_t spec = cur;                       // what was last patched (differs under startup_delay)
    const uint32_t D = P.startup_delay;
    SpecWindow win;
    win.init(spec, D);
    unsigned long long n = 0;
    const unsigned long long base = EMIT ? P.event_off[r] : 0ull;
    for (uint32_t t = 0; t < P.T; ++t) {
        const E *lrow = static_cast<const E *>(P.llen) + (uint64_t)t * P.Q;
        const E *irow = static_cast<const E *>(P.inprog) + (uint64_t)t * P.Q;          // per-tick term
        const int32_t *iq = static_cast<const int32_t *>(P.inprog);                  // per-queue term
        long long sum = 0;
        for (uint64_t q = q0; q < q1; ++q) {
            const int32_t l = ld_trace<E>(lrow + q);
            const int32_t ip = P.inprog_mode == 2 ? ld_trace<E>(irow + q)
                                                  : (P.inprog_mode == 1 ? __ldg(iq + q) : 0);
            sum += decide<MODE>(l, ip, pol, cur);                       // autoscaler.py:255-257
        }
        const int32_t want = clip64(sum, pol.mn, pol.mx, cur);           // :259
        if (want != cur) {                                               // :227
            if (EMIT && base + n < P.capacity) {
                kas_patch_event e;
                e.tick = t;
                e.from_pods = cur;
                e.to_pods = want;
                P.events[base + n] = e;
            }
            ++n;
            if (!(P.patch_fail && P.patch_fail[(uint64_t)t * P.R + r])) spec = want;   // :267-273
        }
        cur = D ? win.push(spec, t, D) : spec;
    }
    if (!EMIT) {
        P.event_off[r + 1] = n;
        if (r == 0) P.event_off[0] = 0ull;
    }
}

// ---- time to the first pod after a burst ------------------------------------
// The same recurrence once more, keeping per resource what a cost report needs beside
// pod-hours and churn: how long work that arrives at an idle resource waits for its first
// pod (definition in include/kas.h).  One thread per
// resource; sums[0..4] = served, unserved, wait_sum, max wait, pending over all resources.
template <typename E, int MODE>
__global__ void __launch_bounds__(kEventThreads)
first_pod_kernel(const __grid_constant__ EventParams P, kas_first_pod *__restrict__ out,
                 unsigned long long *__restrict__ sums) {
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = r < P.R;
    uint32_t served = 0, unserved = 0, wmax = 0, pending = 0;
    unsigned long long wsum = 0;
    if (valid) {
        const PolDev pol = P.pol;
        const uint64_t q0 = P.group_off ? P.group_off[r] : r;
        const uint64_t q1 = P.group_off ? P.group_off[r + 1] : r + 1;
        int32_t cur = P.cur0 ? P.cur0[r * P.cur0_stride] : 0;     // what the controller reads
        int32_t spec = cur;
        const uint32_t D = P.startup_delay;
        SpecWindow win;
        win.init(spec, D);
        long long prev_tally = 0;
        int64_t waiting = -1;
        for (uint32_t t = 0; t < P.T; ++t) {
            const E *lrow = static_cast<const E *>(P.llen) + (uint64_t)t * P.Q;
            const E *irow = static_cast<const E *>(P.inprog) + (uint64_t)t * P.Q;
            const int32_t *iq = static_cast<const int32_t *>(P.inprog);
            long long sum = 0, total = 0;
            for (uint64_t q = q0; q < q1; ++q) {
                const int32_t l = ld_trace<E>(lrow + q);
                const int32_t ip = P.inprog_mode == 2 ? ld_trace<E>(irow + q)
                                                      : (P.inprog_mode == 1 ? __ldg(iq + q) : 0);
                total += (long long)l + (long long)ip;                    // autoscaler.py:73
                sum += decide<MODE>(l, ip, pol, cur);                     // :255-257
            }
            const int32_t read = cur;                                     // :252
            const int32_t want = clip64(sum, pol.mn, pol.mx, cur);        // :259
            if (want != cur && !(P.patch_fail && P.patch_fail[(uint64_t)t * P.R + r])) spec = want;
            cur = D ? win.push(spec, t, D) : spec;                        // running after the tick
            if (waiting < 0 && total > 0 && prev_tally <= 0 && read <= 0) waiting = t;
            if (waiting >= 0) {
                if (cur > 0) {
                    const uint32_t w = (uint32_t)(t - waiting);
                    ++served; wsum += w; wmax = max(wmax, w);
                    waiting = -1;
                } else if (total <= 0) {
                    ++unserved;
                    waiting = -1;
                }
            }
            prev_tally = total;
        }
        if (waiting >= 0) pending = (uint32_t)(P.T - waiting);
        if (out) {
            kas_first_pod o;
            o.n_served = served; o.n_unserved = unserved; o.wait_max = wmax; o.pending = pending;
            o.wait_sum = wsum;
            out[r] = o;
        }
    }
    if (sums) {
        unsigned long long a = served, b = unserved, c = wsum, d = wmax, e = pending;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a += __shfl_xor_sync(0xFFFFFFFFu, a, o);
            b += __shfl_xor_sync(0xFFFFFFFFu, b, o);
            c += __shfl_xor_sync(0xFFFFFFFFu, c, o);
            d = max(d, __shfl_xor_sync(0xFFFFFFFFu, d, o));
            e += __shfl_xor_sync(0xFFFFFFFFu, e, o);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(sums + 0, a);
            atomicAdd(sums + 1, b);
            atomicAdd(sums + 2, c);
            atomicMax(sums + 3, d);
            atomicAdd(sums + 4, e);
        }
    }
}

// ---- exclusive scan of the counts, in place over event_off[1..R] -------------
// (event_off[i+1] becomes the sum of counts 0..i, i.e. the end of resource i.)
__global__ void __launch_bounds__(kScanThreads)
scan_tile_sums_kernel(const unsigned long long *__restrict__ v, uint64_t n,
                      unsigned long long *__restrict__ tile_sums) {
    __shared__ unsigned long long warp_part[kScanThreads / 32];
    const uint64_t base = (uint64_t)blockIdx.x * kScanTile;
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < kScanPerThread; ++k) {
        const uint64_t i = base + (uint64_t)k * kScanThreads + threadIdx.x;
        if (i < n) s += v[i];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
    if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < kScanThreads / 32; ++w) t += warp_part[w];
        tile_sums[blockIdx.x] = t;
    }
}

// One block turns the tile sums into exclusive prefixes (serial over chunks of
// blockDim, a warp-shuffle scan inside a chunk).
__global__ void __launch_bounds__(1024)
scan_tile_prefix_kernel(unsigned long long *__restrict__ tile_sums, uint64_t n_tiles) {
    __shared__ unsigned long long warp_tot[32];
    __shared__ unsigned long long carry_s;
    if (threadIdx.x == 0) carry_s = 0ull;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint64_t c = 0; c < n_tiles; c += blockDim.x) {
        const uint64_t i = c + threadIdx.x;
        const unsigned long long mine = i < n_tiles ? tile_sums[i] : 0ull;
        unsigned long long x = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if (lane >= (uint32_t)o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            unsigned long long w = warp_tot[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned long long y = __shfl_up_sync(0xFFFFFFFFu, w, o);
                if (lane >= (uint32_t)o) w += y;
            }
            warp_tot[lane] = w;   // inclusive over warps
        }
        __syncthreads();
        const unsigned long long before = carry_s + (warp ? warp_tot[warp - 1] : 0ull) + (x - mine);
        if (i < n_tiles) tile_sums[i] = before;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry_s = before + mine;
        __syncthreads();
    }
}

// Inclusive scan inside each tile plus the tile's prefix.  A thread owns
// kScanPerThread CONSECUTIVE counts so the running sum stays in registers.
__global__ void __launch_bounds__(kScanThreads)
scan_apply_kernel(unsigned long long *__restrict__ v, uint64_t n,
                  const unsigned long long *__restrict__ tile_prefix) {
    __shared__ unsigned long long warp_tot[kScanThreads / 32];
    const uint64_t base = (uint64_t)blockIdx.x * kScanTile + (uint64_t)threadIdx.x * kScanPerThread;
    unsigned long long x[kScanPerThread];
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < kScanPerThread; ++k) {
        x[k] = base + k < n ? v[base + k] : 0ull;
        s += x[k];
    }
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned long long incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= (uint32_t)o) incl += y;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    unsigned long long before = tile_prefix ? tile_prefix[blockIdx.x] : 0ull;
    for (uint32_t w = 0; w < warp; ++w) before += warp_tot[w];
    before += incl - s;
#pragma unroll
    for (int k = 0; k < kScanPerThread; ++k) {
        before += x[k];
        if (base + k < n) v[base + k] = before;
    }
}

static uint64_t scan_tiles(uint64_t n) { return (n + kScanTile - 1) / kScanTile; }

// Shared by the entry point that takes the caller's scratch buffer (and by any
// other that brings its own): counts -> offsets.
int scan_counts_in_place(unsigned long long *d_counts, uint64_t n, unsigned long long *d_tile_sums,
                         cudaStream_t stream) {
    if (n == 0) return KAS_OK;
    const uint64_t tiles = scan_tiles(n);
    if (tiles > 1) {
        scan_tile_sums_kernel<<<(unsigned)tiles, kScanThreads, 0, stream>>>(d_counts, n, d_tile_sums);
        scan_tile_prefix_kernel<<<1, 1024, 0, stream>>>(d_tile_sums, tiles);
    }
    scan_apply_kernel<<<(unsigned)tiles, kScanThreads, 0, stream>>>(d_counts, n, tiles > 1 ? d_tile_sums : nullptr);
    KAS_CUDA_TRY(cudaGetLastError());
    return KAS_OK;
}

static int fill_event_params(const kas_replay_args *a, uint32_t policy_index, EventParams &P, int &mode) {
    if (!a || !a->policies || policy_index >= a->n_policies) return KAS_ERR_INVALID;
    if (a->n_queues == 0 || a->n_resources == 0 || a->n_ticks == 0 || !a->d_llen) return KAS_ERR_INVALID;
    if (a->inprog_mode < 0 || a->inprog_mode > 2 || (a->inprog_mode != 0 && !a->d_inprog)) return KAS_ERR_INVALID;
    if (!a->d_group_off && a->n_resources != a->n_queues) return KAS_ERR_INVALID;
    if (a->n_resources > 0x7FFFFFFFull * kEventThreads) return KAS_ERR_UNSUPPORTED;
    if (a->trace_dtype != KAS_TRACE_INT32 && a->trace_dtype != KAS_TRACE_UINT16) return KAS_ERR_INVALID;
    if (a->startup_delay > KAS_MAX_STARTUP_DELAY) return KAS_ERR_INVALID;
    const size_t elem = a->trace_dtype == KAS_TRACE_UINT16 ? 2 : 4;
    if (!aligned(a->d_llen, elem) || !aligned(a->d_inprog, a->inprog_mode == 2 ? elem : 4) ||
        !aligned(a->d_group_off, 8) || !aligned(a->d_cur0, 4))
        return KAS_ERR_ALIGNMENT;
    const kas_policy &p = a->policies[policy_index];
    mode = replay_mode_for(&p, 1);
    if (mode < 0) return KAS_ERR_ZERO_DIVISION;
    P.llen = a->d_llen;
    P.inprog = a->d_inprog;
    P.group_off = a->d_group_off;
    P.cur0 = a->d_cur0;
    P.cur0_stride = 1;
    if (a->d_resume) {
        if (a->d_cur0 || a->startup_delay) return KAS_ERR_INVALID;
        if (!aligned(a->d_resume, 8)) return KAS_ERR_ALIGNMENT;
        P.cur0 = reinterpret_cast<const int32_t *>(a->d_resume) + 3 + 4ull * policy_index;   // .final_current
        P.cur0_stride = 4ull * a->n_policies;
    }
    P.patch_fail = a->d_patch_fail;
    P.Q = a->n_queues;
    P.R = a->n_resources;
    P.T = a->n_ticks;
    P.inprog_mode = a->inprog_mode;
    P.startup_delay = a->startup_delay;
    P.pol = make_poldev(p);
    P.u16 = a->trace_dtype == KAS_TRACE_UINT16;
    P.event_off = nullptr;
    P.events = nullptr;
    P.capacity = 0;
    return KAS_OK;
}

template <bool EMIT>
static int launch_events(const EventParams &P, int mode, cudaStream_t stream) {
    const unsigned grid = (unsigned)((P.R + kEventThreads - 1) / kEventThreads);
    if (P.u16) {
        if (mode == kGeneric) events_kernel<uint16_t, kGeneric, EMIT><<<grid, kEventThreads, 0, stream>>>(P);
        else events_kernel<uint16_t, kSimple, EMIT><<<grid, kEventThreads, 0, stream>>>(P);
    } else {
        if (mode == kGeneric) events_kernel<int32_t, kGeneric, EMIT><<<grid, kEventThreads, 0, stream>>>(P);
        else events_kernel<int32_t, kSimple, EMIT><<<grid, kEventThreads, 0, stream>>>(P);
    }
    KAS_CUDA_TRY(cudaGetLastError());
    return KAS_OK;
}

int patch_event_offsets_impl(const kas_replay_args *args, uint32_t policy_index, uint64_t *d_event_off,
                             unsigned long long *d_tile_sums, cudaStream_t stream) {
    EventParams P;
    int mode = 0;
    const int st = fill_event_params(args, policy_index, P, mode);
    if (st != KAS_OK) return st;
    P.event_off = reinterpret_cast<unsigned long long *>(d_event_off);
    const int st2 = launch_events<false>(P, mode, stream);
    if (st2 != KAS_OK) return st2;
    return scan_counts_in_place(P.event_off + 1, P.R, d_tile_sums, stream);
}

// ---- built-in scratch for callers that must not allocate --------------------
// The scan's tile sums, as a fixed array inside the library image: every GPU that
// loads libkas holds these 64 KB until the process exits.  In host code the name
// below is only a handle for cudaGetSymbolAddress, never a pointer to pass to a
// kernel; the device address is looked up once per device and remembered.
constexpr uint64_t kStaticTiles = KAS_STATIC_SCRATCH_MAX_RESOURCES / kScanTile;
static_assert(KAS_STATIC_SCRATCH_MAX_RESOURCES % kScanTile == 0, "limit must be a whole number of scan tiles");
static __device__ __align__(256) unsigned long long g_static_tile_sums[kStaticTiles];

constexpr int kMaxDevices = 64;
static std::mutex g_static_mu;
static unsigned long long *g_static_addr[kMaxDevices] = {nullptr};

// Device address of the built-in scratch on the CURRENT device.  The first call on
// a device asks the runtime for it (no allocation: the array came with the module).
static int static_tile_sums(unsigned long long **out) {
    int dev = -1;
    KAS_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return KAS_ERR_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (g_static_addr[dev] == nullptr) {
        void *p = nullptr;
        KAS_CUDA_TRY(cudaGetSymbolAddress(&p, g_static_tile_sums));
        if (p == nullptr) return KAS_ERR_CUDA;
        g_static_addr[dev] = static_cast<unsigned long long *>(p);
    }
    *out = g_static_addr[dev];
    return KAS_OK;
}

}  // namespace kas

using namespace kas;

extern "C" {

size_t kas_patch_events_scratch_bytes(uint64_t n_resources) {
    const uint64_t tiles = scan_tiles(n_resources);
    return tiles > 1 ? tiles * sizeof(unsigned long long) : 0;
}

int kas_patch_event_offsets(const kas_replay_args *args, uint32_t policy_index, uint64_t *d_event_off,
                            void *d_scratch, size_t scratch_bytes, void *stream) {
    if (!args || !d_event_off) return KAS_ERR_INVALID;
    if (!aligned(d_event_off, 8) || !aligned(d_scratch, 8)) return KAS_ERR_ALIGNMENT;
    const size_t need = kas_patch_events_scratch_bytes(args->n_resources);
    if (need && (!d_scratch || scratch_bytes < need)) return KAS_ERR_TOO_SMALL;
    return patch_event_offsets_impl(args, policy_index, d_event_off,
                                    static_cast<unsigned long long *>(d_scratch),
                                    static_cast<cudaStream_t>(stream));
}

int kas_patch_event_emit(const kas_replay_args *args, uint32_t policy_index, const uint64_t *d_event_off,
                         kas_patch_event *d_events, uint64_t capacity, void *stream) {
    if (!args || !d_event_off || (capacity && !d_events)) return KAS_ERR_INVALID;
    if (!aligned(d_event_off, 8) || !aligned(d_events, 4)) return KAS_ERR_ALIGNMENT;
    EventParams P;
    int mode = 0;
    const int st = fill_event_params(args, policy_index, P, mode);
    if (st != KAS_OK) return st;
    P.event_off = const_cast<unsigned long long *>(reinterpret_cast<const unsigned long long *>(d_event_off));
    P.events = d_events;
    P.capacity = capacity;
    if (capacity == 0) return KAS_OK;
    return launch_events<true>(P, mode, static_cast<cudaStream_t>(stream));
}

int kas_first_pod_latency(const kas_replay_args *args, uint32_t policy_index, kas_first_pod *d_out,
                          int64_t *d_sums, void *stream_v) {
    if (!args || (!d_out && !d_sums)) return KAS_ERR_INVALID;
    if (!aligned(d_out, 8) || !aligned(d_sums, 8)) return KAS_ERR_ALIGNMENT;
    EventParams P;
    int mode = 0;
    const int st = fill_event_params(args, policy_index, P, mode);
    if (st != KAS_OK) return st;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    unsigned long long *sums = reinterpret_cast<unsigned long long *>(d_sums);
    if (sums) KAS_CUDA_TRY(cudaMemsetAsync(sums, 0, 5 * sizeof(unsigned long long), stream));
    const unsigned grid = (unsigned)((P.R + kEventThreads - 1) / kEventThreads);
    if (P.u16) {
        if (mode == kGeneric) first_pod_kernel<uint16_t, kGeneric><<<grid, kEventThreads, 0, stream>>>(P, d_out, sums);
        else first_pod_kernel<uint16_t, kSimple><<<grid, kEventThreads, 0, stream>>>(P, d_out, sums);
    } else {
        if (mode == kGeneric) first_pod_kernel<int32_t, kGeneric><<<grid, kEventThreads, 0, stream>>>(P, d_out, sums);
        else first_pod_kernel<int32_t, kSimple><<<grid, kEventThreads, 0, stream>>>(P, d_out, sums);
    }
    KAS_CUDA_TRY(cudaGetLastError());
    return KAS_OK;
}

int kas_patch_event_offsets_static(const kas_replay_args *args, uint32_t policy_index,
                                   uint64_t *d_event_off, void *stream) {
    if (!args || !d_event_off) return KAS_ERR_INVALID;
    if (!aligned(d_event_off, 8)) return KAS_ERR_ALIGNMENT;
    if (args->n_resources > KAS_STATIC_SCRATCH_MAX_RESOURCES) return KAS_ERR_STATIC_LIMIT;
    unsigned long long *tile_sums = nullptr;
    const int st = static_tile_sums(&tile_sums);
    if (st != KAS_OK) return st;
    // Nothing in the array has to start at zero: every tile sum the scan reads in this
    // call is written by this call first (scan_tile_sums_kernel), whatever an earlier
    // call left behind.
    return patch_event_offsets_impl(args, policy_index, d_event_off, tile_sums,
                                    static_cast<cudaStream_t>(stream));
}

}  // extern "C"
