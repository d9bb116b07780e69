// kas_glob.cu -- general Redis-glob match and count over a variable-length key
// store, for queue names that contain glob metacharacters.
//
// autoscaler.py:69 builds the SCAN pattern as 'processing-{}:*'.format(q) without
// escaping q, so a queue name with * ? [ or \ is a live glob in the reference.
// The prefix kernel (kas_match.cu) refuses such names; this kernel evaluates them
// with the matcher semantics of `SCAN ... MATCH` (redis-server util.c
// stringmatchlen, restated from its published behaviour -- see oracle/oracle.py,
// parity unpinned upstream):
//   *  any run of bytes (also empty)      ?  any one byte
//   [abc] [a-z] [^x]  classes; a backslash inside a class escapes the next byte;
//   a reversed range is swapped; an unterminated class ends with the pattern
//   \x  the byte x; a trailing lone backslash is a literal backslash
// One thread per key, all lanes of a warp walk the same pattern at the same time,
// per-pattern counts by warp ballot + popc.  The matcher is iterative: a mismatch
// after a '*' resumes one byte further behind the LAST star, which is equivalent
// to the server's recursive search for these pattern elements.
// The keys of a block's batch (256 consecutive keys = one contiguous run of the blob) are
// copied to shared memory with coalesced 16-byte loads first, so the byte-by-byte walk of
// every pattern reads shared memory instead of issuing dependent one-byte global loads; a
// batch whose bytes do not fit (very long keys) is matched straight from global memory.
#include "kas_common.cuh"

namespace kas {

constexpr int kGlobThreads = 256;
constexpr uint32_t kGlobMaxPatternBytes = 16 * 1024;
constexpr uint32_t kGlobStageBytes = 32 * 1024;     // key bytes of one batch staged in shared memory

// One key byte: from shared memory (a 32-bit shared address) or from global memory.
template <bool SHARED>
__device__ __forceinline__ uint8_t key_byte(const uint8_t *s, uint32_t saddr, uint64_t i) {
    if constexpr (SHARED) {
        uint32_t v;
        asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(saddr + (uint32_t)i));
        return (uint8_t)v;
    } else {
        return __ldg(s + i);
    }
}

// Does pattern element at p[pi..] match byte c?  Advances pi past the element.
// Mirrors the class / escape / literal cases of the server's matcher.
__device__ __forceinline__ bool glob_element(const uint8_t *p, uint32_t plen, uint32_t &pi, uint8_t c) {
    const uint8_t e = p[pi];
    if (e == '?') { pi += 1; return true; }
    if (e == '[') {
        uint32_t i = pi + 1;
        const bool neg = i < plen && p[i] == '^';
        if (neg) ++i;
        bool hit = false;
        for (;;) {
            if (i + 1 < plen && p[i] == '\\') {          // escaped member
                ++i;
                if (p[i] == c) hit = true;
            } else if (i < plen && p[i] == ']') {
                break;
            } else if (i >= plen) {                      // unterminated: class ends with the pattern
                --i;
                break;
            } else if (i + 2 < plen && p[i + 1] == '-') {
                uint8_t lo = p[i], hi = p[i + 2];
                if (lo > hi) { const uint8_t t = lo; lo = hi; hi = t; }
                i += 2;
                if (c >= lo && c <= hi) hit = true;
            } else if (p[i] == c) {
                hit = true;
            }
            ++i;
        }
        pi = i + 1;
        return neg ? !hit : hit;
    }
    if (e == '\\' && pi + 1 < plen) {
        const bool ok = p[pi + 1] == c;
        pi += 2;
        return ok;
    }
    pi += 1;
    return e == c;
}

// `lit`: length of the pattern's leading run of plain bytes (nothing of * ? [ \).  That run can
// only match the same bytes at the start of the key, so it is compared in a tight loop first --
// most (key, pattern) pairs end there -- and the general walk starts behind it.
template <bool SHARED>
__device__ bool glob_match(const uint8_t *p, uint32_t plen, uint32_t lit, const uint8_t *s, uint32_t saddr,
                           uint64_t slen) {
    if (slen < lit) return false;
    for (uint32_t i = 0; i < lit; ++i)
        if (key_byte<SHARED>(s, saddr, i) != p[i]) return false;
    uint32_t pi = lit, star_pi = 0xFFFFFFFFu;
    uint64_t si = lit, star_si = 0;
    while (si < slen) {
        if (pi < plen && p[pi] == '*') {
            while (pi + 1 < plen && p[pi + 1] == '*') ++pi;
            if (pi + 1 == plen) return true;             // trailing star swallows the rest
            star_pi = pi;
            star_si = si;
            ++pi;
            continue;
        }
        if (pi < plen) {
            uint32_t next = pi;
            if (glob_element(p, plen, next, key_byte<SHARED>(s, saddr, si))) {
                pi = next;
                ++si;
                continue;
            }
        }
        if (star_pi == 0xFFFFFFFFu) return false;
        pi = star_pi + 1;                                // let the last star take one more byte
        si = ++star_si;
    }
    while (pi < plen && p[pi] == '*') ++pi;
    return pi == plen;
}

__global__ void __launch_bounds__(kGlobThreads)
glob_count_kernel(const uint8_t *__restrict__ blob, const uint64_t *__restrict__ off, uint64_t n_keys,
                  uint32_t sep_bytes, const uint8_t *__restrict__ patterns,
                  const uint32_t *__restrict__ pat_off, uint32_t n_patterns,
                  unsigned long long *__restrict__ counts) {
    extern __shared__ uint8_t gsm[];
    // [n_patterns+1] offsets, [n_patterns] histogram, [n_patterns] literal-prefix lengths, then the pattern bytes
    uint32_t *s_off = reinterpret_cast<uint32_t *>(gsm);
    uint32_t *s_hist = s_off + n_patterns + 1;
    uint32_t *s_lit = s_hist + n_patterns;
    uint8_t *s_pat = reinterpret_cast<uint8_t *>(s_lit + n_patterns);
    for (uint32_t i = threadIdx.x; i <= n_patterns; i += blockDim.x) s_off[i] = pat_off[i];
    for (uint32_t i = threadIdx.x; i < n_patterns; i += blockDim.x) s_hist[i] = 0u;
    __syncthreads();
    const uint32_t total = s_off[n_patterns];
    for (uint32_t i = threadIdx.x; i < total; i += blockDim.x) s_pat[i] = patterns[i];
    __syncthreads();
    for (uint32_t q = threadIdx.x; q < n_patterns; q += blockDim.x) {
        const uint32_t pa = s_off[q], pl = s_off[q + 1] - pa;
        uint32_t n = 0;
        while (n < pl) {
            const uint8_t c = s_pat[pa + n];
            if (c == '*' || c == '?' || c == '[' || c == '\\') break;
            ++n;
        }
        s_lit[q] = n;
    }
    __syncthreads();

    const uint32_t lane = threadIdx.x & 31u;
    // the staging area: 16-byte aligned, behind the pattern bytes
    uint8_t *s_keys = gsm + (((3 * n_patterns + 1) * 4 + total + 15u) & ~15u);
    const uint32_t s_keys_addr = (uint32_t)__cvta_generic_to_shared(s_keys);
    for (uint64_t base = (uint64_t)blockIdx.x * blockDim.x; base < n_keys; base += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t i = base + threadIdx.x;
        const bool live = i < n_keys;
        uint64_t a = 0, len = 0;
        if (live) {
            a = off[i];
            const uint64_t span = off[i + 1] - a;
            len = span > sep_bytes ? span - sep_bytes : 0;
        }
        // the batch's run of the blob, widened to 16-byte boundaries
        const uint64_t last = min(base + (uint64_t)blockDim.x, n_keys);
        const uint64_t b0 = off[base], b1 = off[last];
        const uintptr_t g0 = (reinterpret_cast<uintptr_t>(blob) + b0) & ~(uintptr_t)15;
        const uintptr_t g1 = (reinterpret_cast<uintptr_t>(blob) + b1 + 15) & ~(uintptr_t)15;
        const bool staged = b1 >= b0 && g1 - g0 <= kGlobStageBytes;          // uniform over the block
        if (staged) {
            const uint32_t pieces = (uint32_t)((g1 - g0) / 16);
            for (uint32_t k = threadIdx.x; k < pieces; k += blockDim.x)
                reinterpret_cast<uint4 *>(s_keys)[k] = ldg_stream_128(reinterpret_cast<const uint4 *>(g0) + k);
            __syncthreads();
        }
        const uint32_t my = s_keys_addr + (uint32_t)(reinterpret_cast<uintptr_t>(blob) + a - g0);
        for (uint32_t q = 0; q < n_patterns; ++q) {       // warp-uniform loop
            const uint32_t pa = s_off[q], pl = s_off[q + 1] - pa;
            const uint32_t lit = s_lit[q];
            const bool m = live && (staged ? glob_match<true>(s_pat + pa, pl, lit, nullptr, my, len)
                                           : glob_match<false>(s_pat + pa, pl, lit, blob + a, 0u, len));
            const uint32_t votes = __ballot_sync(0xFFFFFFFFu, m);
            if (lane == 0 && votes) atomicAdd(&s_hist[q], (uint32_t)__popc(votes));
        }
        if (staged) __syncthreads();                      // the next batch overwrites the staging area
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n_patterns; i += blockDim.x) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&counts[i], (unsigned long long)c);
    }
}

}  // namespace kas

using namespace kas;

extern "C" {

int kas_glob_count(const uint8_t *d_blob, const uint64_t *d_off, uint64_t n_keys, uint32_t sep_bytes,
                   const uint8_t *d_patterns, const uint32_t *d_pat_off, uint32_t n_patterns,
                   uint32_t pattern_bytes, uint64_t *d_counts, void *stream_v) {
    if (!d_off || !d_pat_off || !d_counts || (pattern_bytes && !d_patterns)) return KAS_ERR_INVALID;
    if (n_patterns == 0 || n_patterns > KAS_MAX_PREFIXES) return KAS_ERR_INVALID;
    if (pattern_bytes > kGlobMaxPatternBytes) return KAS_ERR_UNSUPPORTED;
    if (!aligned(d_off, 8) || !aligned(d_pat_off, 4) || !aligned(d_counts, 8)) return KAS_ERR_ALIGNMENT;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
    KAS_CUDA_TRY(cudaMemsetAsync(d_counts, 0, sizeof(uint64_t) * n_patterns, stream));
    if (n_keys == 0) return KAS_OK;
    const int sms = sm_count();
    if (sms <= 0) return KAS_ERR_CUDA;
    const size_t smem = (((3 * (size_t)n_patterns + 1) * sizeof(uint32_t) + pattern_bytes + 15) & ~(size_t)15) +
                        kGlobStageBytes;
    KAS_CUDA_TRY(cudaFuncSetAttribute(glob_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    uint64_t grid = (n_keys + kGlobThreads - 1) / kGlobThreads;
    if (grid > (uint64_t)sms * 8) grid = (uint64_t)sms * 8;
    glob_count_kernel<<<(unsigned)grid, kGlobThreads, smem, stream>>>(
        d_blob, d_off, n_keys, sep_bytes, d_patterns, d_pat_off, n_patterns,
        reinterpret_cast<unsigned long long *>(d_counts));
    KAS_CUDA_TRY(cudaGetLastError());
    return KAS_OK;
}

}  // extern "C"
