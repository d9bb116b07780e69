// kas_ingest.cu -- key-snapshot ingest (variable-length keys -> fixed-width
// heads), synthetic keyspace generator, and the library's small utilities.
//
// Ingest stands where the reference pulls every key over TCP once per queue
// (autoscaler.py:70-71): the keyspace is taken once, each key's first
// rec_bytes bytes are stored zero-padded and contiguous.
#include "kas_common.cuh"
#include "kas_table.h"

#include <string.h>
#include <algorithm>
#include <atomic>
#include <string>
#include <thread>
#include <vector>

namespace kas {

static thread_local char g_cuda_error[512] = "";

void set_cuda_error(cudaError_t e, const char *where) {
    snprintf(g_cuda_error, sizeof(g_cuda_error), "%s: %s (%s)", where, cudaGetErrorName(e),
             cudaGetErrorString(e));
}

void set_text_error(const char *where, const char *what) {
    snprintf(g_cuda_error, sizeof(g_cuda_error), "%s: %s", where, what);
}

static std::atomic<int> g_reserved_sms{0};
int reserved_sms() { return g_reserved_sms.load(std::memory_order_relaxed); }

int sm_count() {
    static thread_local int cached_dev = -1;
    static thread_local int cached_sms = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (dev != cached_dev) {
        int sms = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
        cached_dev = dev;
        cached_sms = sms;
    }
    return cached_sms;
}

template <typename F>
static void parallel_for(uint64_t n, int threads, F fn) {
    if (threads < 1) threads = 1;
    if ((uint64_t)threads > n) threads = n ? (int)n : 1;
    if (threads == 1) {
        fn(0, n);
        return;
    }
    std::vector<std::thread> pool;
    const uint64_t per = (n + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        const uint64_t a = std::min<uint64_t>(n, per * t), b = std::min<uint64_t>(n, per * (t + 1));
        if (a < b) pool.emplace_back([=] { fn(a, b); });
    }
    for (auto &th : pool) th.join();
}

// One thread per 16-byte chunk of the output.  The chunk's source bytes sit at an
// arbitrary byte address, so the thread reads the (up to five) aligned 32-bit words
// that cover them and funnel-shifts them into place; bytes past the key's end are
// masked to zero and words that lie wholly past it are not read at all.
__global__ void __launch_bounds__(256)
pack_heads_kernel(const uint8_t *__restrict__ blob, const uint64_t *__restrict__ off,
                  uint64_t n_keys, uint32_t sep_bytes, uint32_t chunks, uint4 *__restrict__ heads) {
    const uint64_t total = n_keys * chunks;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t key = i / chunks;
        const uint32_t c = (uint32_t)(i - key * chunks);
        const uint64_t a = __ldg(off + key), b = __ldg(off + key + 1);
        const uint64_t len = (b - a) > sep_bytes ? (b - a) - sep_bytes : 0;
        const uint64_t lo = 16ull * c;
        uint32_t w[4] = {0u, 0u, 0u, 0u};
        if (len > lo) {
            const uint32_t n = (uint32_t)(len - lo < 16 ? len - lo : 16);   // valid bytes, 1..16
            const uintptr_t addr = reinterpret_cast<uintptr_t>(blob) + a + lo;
            const uint32_t *wp = reinterpret_cast<const uint32_t *>(addr & ~(uintptr_t)3);
            const uint32_t mis = (uint32_t)(addr & 3u);                     // bytes to drop from word 0
            const uint32_t need = (mis + n + 3u) >> 2;                      // aligned words touched, 1..5
            uint32_t v[5];
#pragma unroll
            for (int j = 0; j < 5; ++j) v[j] = (uint32_t)j < need ? __ldg(wp + j) : 0u;
            const uint32_t sh = mis * 8u;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t x = __funnelshift_r(v[j], v[j + 1], sh);
                const int rem = (int)n - 4 * j;                             // valid bytes of this word
                const uint32_t m = rem >= 4 ? 0xFFFFFFFFu : (rem <= 0 ? 0u : (0xFFFFFFFFu >> (32 - 8 * rem)));
                w[j] = x & m;
            }
        }
        heads[i] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

static inline uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

static const char kAlphabet[65] = "0123456789abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ-:";

// Build synthetic key i into `out` (spec: DESIGN.md "Synthetic keyspace").
static void synth_key(uint64_t seed, uint64_t i, const uint8_t *names, const uint32_t *name_off,
                      uint32_t n_names, int long_tail, std::string &out) {
    out.clear();
    const uint64_t h0 = mix64(seed ^ mix64(i));
    const uint64_t h1 = mix64(h0);
    const uint32_t cls = (uint32_t)(h0 % 100u);
    const uint32_t j = (uint32_t)((h0 >> 8) % n_names);
    uint32_t sl = (uint32_t)((h0 >> 32) % 25u);
    if (long_tail && (h1 & 0xFFu) < 13u) sl += (uint32_t)((h1 >> 8) % 680u);
    const char *nm = reinterpret_cast<const char *>(names + name_off[j]);
    const uint32_t nl = name_off[j + 1] - name_off[j];
    auto suffix = [&]() {
        uint64_t word = 0;
        for (uint32_t k = 0; k < sl; ++k) {
            if (k % 10 == 0) word = mix64(h1 + k / 10);
            out.push_back(kAlphabet[(word >> (6 * (k % 10))) & 63u]);
        }
    };
    if (cls < 50) { out.append("processing-"); out.append(nm, nl); out.push_back(':'); suffix(); }
    else if (cls < 56) { out.append("processing-"); out.append(nm, nl); out.append("x:"); suffix(); }
    else if (cls < 61) { out.append("processing-"); out.append(nm, nl); }
    else if (cls < 66) { out.append("xprocessing-"); out.append(nm, nl); out.push_back(':'); suffix(); }
    else if (cls < 71) { out.append("processing-"); out.append(nm, nl ? nl - 1 : 0); out.push_back(':'); suffix(); }
    else if (cls < 74) { out.append(nm, nl); }
    else if (cls < 77) { out.append("processing-"); out.append(nm, nl); out.push_back(':'); }
    else if (cls < 80) { out.append("Processing-"); out.append(nm, nl); out.push_back(':'); suffix(); }
    else {
        out.append(nm, nl); out.push_back(':'); suffix(); out.push_back(':');
        static const char hexd[] = "0123456789abcdef";
        for (int k = 0; k < 16; ++k) out.push_back(hexd[(h1 >> (4 * k)) & 15u]);
    }
}

}  // namespace kas

using namespace kas;

extern "C" {

int kas_version(void) { return KAS_VERSION; }

const char *kas_strerror(int status) {
    switch (status) {
        case KAS_OK: return "ok";
        case KAS_ERR_INVALID: return "invalid argument";
        case KAS_ERR_UNSUPPORTED: return "outside the accelerated domain";
        case KAS_ERR_CUDA: return "CUDA call failed";
        case KAS_ERR_ALIGNMENT: return "device pointer is not aligned as required";
        case KAS_ERR_ZERO_DIVISION: return "integer division or modulo by zero";
        case KAS_ERR_GLOB_META: return "queue name contains a glob metacharacter";
        case KAS_ERR_TOO_SMALL: return "output buffer too small";
        case KAS_ERR_STATIC_LIMIT:
            return "more resources than the built-in scratch covers (limit 33554432); "
                   "use the variant that takes a scratch buffer";
    }
    return "unknown status";
}

const char *kas_last_cuda_error(void) { return g_cuda_error; }

int kas_set_reserved_sms(int n) {
    if (n < 0 || n > 64) return KAS_ERR_INVALID;
    g_reserved_sms.store(n, std::memory_order_relaxed);
    return KAS_OK;
}

int kas_device_info(int *sms, int *cc_major, int *cc_minor) {
    int dev = 0;
    KAS_CUDA_TRY(cudaGetDevice(&dev));
    int a = 0, b = 0, c = 0;
    KAS_CUDA_TRY(cudaDeviceGetAttribute(&a, cudaDevAttrMultiProcessorCount, dev));
    KAS_CUDA_TRY(cudaDeviceGetAttribute(&b, cudaDevAttrComputeCapabilityMajor, dev));
    KAS_CUDA_TRY(cudaDeviceGetAttribute(&c, cudaDevAttrComputeCapabilityMinor, dev));
    if (sms) *sms = a;
    if (cc_major) *cc_major = b;
    if (cc_minor) *cc_minor = c;
    return KAS_OK;
}

int kas_pack_heads_host(const uint8_t *blob, const uint64_t *off, uint64_t n_keys,
                        uint32_t sep_bytes, uint32_t rec_bytes, uint8_t *out, int threads) {
    if (!off || !out || (!blob && n_keys && off[n_keys] != off[0])) return KAS_ERR_INVALID;
    if (rec_bytes == 0 || rec_bytes % 16) return KAS_ERR_UNSUPPORTED;
    parallel_for(n_keys, threads, [=](uint64_t a, uint64_t b) {
        for (uint64_t i = a; i < b; ++i) {
            const uint64_t span = off[i + 1] - off[i];
            const uint64_t len = span > sep_bytes ? span - sep_bytes : 0;
            const uint32_t n = (uint32_t)(len < rec_bytes ? len : rec_bytes);
            uint8_t *dst = out + i * rec_bytes;
            memcpy(dst, blob + off[i], n);
            memset(dst + n, 0, rec_bytes - n);
        }
    });
    return KAS_OK;
}

int kas_index_lines_host(const uint8_t *blob, uint64_t blob_len, uint64_t *off, uint64_t off_cap,
                         uint64_t *n_keys) {
    if ((!blob && blob_len) || !n_keys) return KAS_ERR_INVALID;
    // off[i] is the start of line i and off[i+1]-1 its end (the byte before the next
    // start is the '\n'); pack with sep_bytes = 1.  A dump that does not end in '\n' gets
    // a final entry of blob_len+1, as if the newline were there.
    uint64_t n = 0;
    uint64_t start = 0;
    for (uint64_t i = 0; i < blob_len; ++i) {
        if (blob[i] == '\n') {
            if (off) {
                if (n >= off_cap) return KAS_ERR_TOO_SMALL;
                off[n] = start;
            }
            ++n;
            start = i + 1;
        }
    }
    if (start < blob_len) {  // last line without a trailing newline
        if (off) {
            if (n >= off_cap) return KAS_ERR_TOO_SMALL;
            off[n] = start;
        }
        ++n;
        start = blob_len + 1;
    }
    if (off) {
        if (n >= off_cap) return KAS_ERR_TOO_SMALL;
        off[n] = start;
    }
    *n_keys = n;
    return KAS_OK;
}

int kas_pack_heads(const uint8_t *d_blob, const uint64_t *d_off, uint64_t n_keys,
                   uint32_t sep_bytes, uint32_t rec_bytes, uint8_t *d_heads, void *stream) {
    if (!d_off || !d_heads) return KAS_ERR_INVALID;
    if (rec_bytes == 0 || rec_bytes % 16) return KAS_ERR_UNSUPPORTED;
    if (!aligned(d_heads, 16) || !aligned(d_off, 8)) return KAS_ERR_ALIGNMENT;
    if (n_keys == 0) return KAS_OK;
    const int sms = sm_count();
    if (sms <= 0) return KAS_ERR_CUDA;
    const uint32_t chunks = rec_bytes / 16;
    const uint64_t total = n_keys * chunks;
    uint64_t grid = (total + 255) / 256;
    const uint64_t cap = (uint64_t)sms * 8;
    if (grid > cap) grid = cap;
    pack_heads_kernel<<<(unsigned)grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        d_blob, d_off, n_keys, sep_bytes, chunks, reinterpret_cast<uint4 *>(d_heads));
    KAS_CUDA_TRY(cudaGetLastError());
    return KAS_OK;
}

int kas_synth_keys_host(uint64_t seed, uint64_t first_key, uint64_t n_keys, const uint8_t *names,
                        const uint32_t *name_off, uint32_t n_names, int long_tail,
                        uint32_t rec_bytes, uint8_t *out_heads, uint8_t *out_blob,
                        uint64_t blob_cap, uint64_t *out_off, uint64_t *blob_len, int threads) {
    if (!names || !name_off || n_names == 0) return KAS_ERR_INVALID;
    if (out_heads && (rec_bytes == 0 || rec_bytes % 16)) return KAS_ERR_UNSUPPORTED;
    if ((out_blob != nullptr) != (out_off != nullptr)) return KAS_ERR_INVALID;
    if (out_heads) {
        parallel_for(n_keys, threads, [=](uint64_t a, uint64_t b) {
            std::string k;
            for (uint64_t i = a; i < b; ++i) {
                synth_key(seed, first_key + i, names, name_off, n_names, long_tail, k);
                const uint32_t n = (uint32_t)std::min<size_t>(k.size(), rec_bytes);
                uint8_t *dst = out_heads + i * rec_bytes;
                memcpy(dst, k.data(), n);
                memset(dst + n, 0, rec_bytes - n);
            }
        });
    }
    if (out_off || blob_len) {
        // lengths first (parallel), then a serial prefix sum, then the bytes (parallel)
        std::vector<uint64_t> local;
        uint64_t *offs = out_off;
        if (!offs) {
            local.resize(n_keys + 1);
            offs = local.data();
        }
        parallel_for(n_keys, threads, [=](uint64_t a, uint64_t b) {
            std::string k;
            for (uint64_t i = a; i < b; ++i) {
                synth_key(seed, first_key + i, names, name_off, n_names, long_tail, k);
                offs[i + 1] = k.size();
            }
        });
        offs[0] = 0;
        for (uint64_t i = 0; i < n_keys; ++i) offs[i + 1] += offs[i];
        if (blob_len) *blob_len = offs[n_keys];
        if (out_blob) {
            if (offs[n_keys] > blob_cap) return KAS_ERR_TOO_SMALL;
            parallel_for(n_keys, threads, [=](uint64_t a, uint64_t b) {
                std::string k;
                for (uint64_t i = a; i < b; ++i) {
                    synth_key(seed, first_key + i, names, name_off, n_names, long_tail, k);
                    memcpy(out_blob + offs[i], k.data(), k.size());
                }
            });
        }
    }
    return KAS_OK;
}

}  // extern "C"
