// kas_common.cuh -- shared helpers for libkas (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/kas.h"

namespace kas {

void set_cuda_error(cudaError_t e, const char *where);
void set_text_error(const char *where, const char *what);   // same slot, for non-CUDA libraries (NCCL)

#define KAS_CUDA_TRY(expr)                                    \
    do {                                                      \
        cudaError_t _e = (expr);                              \
        if (_e != cudaSuccess) {                              \
            ::kas::set_cuda_error(_e, #expr);                 \
            return KAS_ERR_CUDA;                              \
        }                                                     \
    } while (0)

// SM count of the current device (cached per device by the caller's thread).
int sm_count();
// SMs the persistent kernels leave free (kas_set_reserved_sms; 0 by default).
int reserved_sms();

static inline bool aligned(const void *p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

#ifdef __CUDACC__
// Streaming 128-bit load: read-once data, keep it out of L1.
__device__ __forceinline__ uint4 ldg_stream_128(const void *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// 128-bit load that may allocate in L1 (used when the neighbouring 16 bytes of
// the same sector are fetched by the next instruction of the same thread).
__device__ __forceinline__ uint4 ldg_nc_128(const void *p) {
    uint4 r;
    asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// 128-bit shared-memory load at a 32-bit shared address (the address arithmetic stays 32-bit
// and the window base is not re-derived at every use).
__device__ __forceinline__ uint4 lds_128(uint32_t saddr) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(saddr));
    return r;
}
// sm_100 256-bit load (LDG.E.256): one whole 32-byte sector per thread.
__device__ __forceinline__ void ldg_stream_256(const void *p, uint32_t (&r)[8]) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]),
                   "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(p));
}
__device__ __forceinline__ int4 ldg_stream_i4(const void *p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ int ldg_stream_i32(const void *p) {
    int r;
    asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
// One element of a queue-depth trace (int32 or uint16 storage) as an int32, streaming.
template <typename E>
__device__ __forceinline__ int ld_trace(const E *p);
template <>
__device__ __forceinline__ int ld_trace<int32_t>(const int32_t *p) { return ldg_stream_i32(p); }
template <>
__device__ __forceinline__ int ld_trace<uint16_t>(const uint16_t *p) {
    unsigned short r;
    asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(r) : "l"(p));
    return (int)r;
}
// The same through the cached path (small, re-read inputs).
template <typename E>
__device__ __forceinline__ int ld_trace_cached(const E *p) { return (int)__ldg(p); }
// N = 16 / sizeof(E) consecutive trace elements with one 128-bit streaming load.
template <typename E, int N>
__device__ __forceinline__ void ld_trace_vec(const E *p, int (&out)[N]) {
    static_assert(N * sizeof(E) == 16, "one 128-bit load");
    const int4 w = ldg_stream_i4(p);
    if constexpr (sizeof(E) == 4) {
        out[0] = w.x; out[1] = w.y; out[2] = w.z; out[3] = w.w;
    } else {
        const unsigned x[4] = {(unsigned)w.x, (unsigned)w.y, (unsigned)w.z, (unsigned)w.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            out[2 * k] = (int)(x[k] & 0xFFFFu);
            out[2 * k + 1] = (int)(x[k] >> 16);
        }
    }
}
#endif

}  // namespace kas
