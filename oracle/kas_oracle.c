/*
 * kas_oracle.c -- CPU restatement of kiosk-autoscaler's tally/decision path.
 *
 * TEST INFRASTRUCTURE ONLY.  Built into oracle/libkas_oracle.so by
 * oracle/build.py; loaded by tests/, __graft_entry__.smoke() and bench.py's
 * CPU-baseline / --impl reference legs.  Nothing in the product library links
 * or calls it.
 *
 * Citations are to the reference checkout (autoscaler/autoscaler.py unless
 * stated).  The arithmetic follows oracle/oracle.py, which is the readable
 * form; this file exists so that parity can be checked at sizes where Python
 * loops are too slow, and so that the CPU baseline can use every host core.
 *
 * Parity status: decision math and LLEN term pinned by the reference's
 * known-answer test (autoscaler_test.py:87-102) and by tests/golden/ (made by
 * executing the reference's own source, oracle/make_golden.py).  SCAN MATCH
 * term: PARITY UNPINNED upstream (glob runs in redis-server, version
 * unpinned; no reference test creates a processing-* key); the matcher below
 * restates the published Redis glob rules.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define KAS_EXPORT __attribute__((visibility("default")))

/* ---- Redis glob (redis-server util.c stringmatchlen semantics) ---------- */
static int glob_rec(const uint8_t *p, long pl, const uint8_t *s, long sl)
{
    while (pl > 0) {
        uint8_t c = p[0];
        if (c == '*') {
            while (pl > 1 && p[1] == '*') { p++; pl--; }
            if (pl == 1) return 1;
            for (;;) {
                if (glob_rec(p + 1, pl - 1, s, sl)) return 1;
                if (sl == 0) return 0;
                s++; sl--;
            }
        } else if (c == '?') {
            if (sl == 0) return 0;
            s++; sl--; p++; pl--;
            continue;
        } else if (c == '[') {
            int neg, hit = 0;
            if (sl == 0) return 0;
            p++; pl--;
            neg = (pl > 0 && p[0] == '^');
            if (neg) { p++; pl--; }
            for (;;) {
                if (pl >= 2 && p[0] == '\\') {
                    p++; pl--;
                    if (p[0] == s[0]) hit = 1;
                } else if (pl > 0 && p[0] == ']') {
                    break;
                } else if (pl <= 0) {
                    p--; pl++;
                    break;
                } else if (pl >= 3 && p[1] == '-') {
                    uint8_t lo = p[0], hi = p[2];
                    if (lo > hi) { uint8_t t = lo; lo = hi; hi = t; }
                    p += 2; pl -= 2;
                    if (s[0] >= lo && s[0] <= hi) hit = 1;
                } else {
                    if (p[0] == s[0]) hit = 1;
                }
                p++; pl--;
            }
            if (neg) hit = !hit;
            if (!hit) return 0;
            s++; sl--; p++; pl--;
            continue;
        } else if (c == '\\' && pl >= 2) {
            p++; pl--;
            c = p[0];
        }
        if (sl == 0 || s[0] != c) return 0;
        s++; sl--; p++; pl--;
    }
    return sl == 0;
}

KAS_EXPORT int kas_oracle_glob_match(const uint8_t *pat, uint32_t pat_len,
                                     const uint8_t *str, uint32_t str_len)
{
    return glob_rec(pat, (long)pat_len, str, (long)str_len);
}

/* ---- SCAN term of tally_queues (autoscaler.py:69-71) --------------------
 * One full pass over the keyspace PER QUEUE, like the reference: pattern
 * "processing-" + q + ":*" (autoscaler.py:69) applied to every key.
 * Keys are a blob + offsets (key i = blob[off[i] .. off[i+1])).             */
KAS_EXPORT void kas_oracle_match_counts(const uint8_t *blob, const uint64_t *off,
                                        uint64_t n_keys,
                                        const uint8_t *names, const uint32_t *name_off,
                                        uint32_t n_queues, uint64_t *counts,
                                        int threads)
{
    for (uint32_t q = 0; q < n_queues; ++q) {
        uint32_t nl = name_off[q + 1] - name_off[q];
        uint32_t pl = 11 + nl + 2;
        uint8_t *pat = (uint8_t *)malloc(pl);
        memcpy(pat, "processing-", 11);
        memcpy(pat + 11, names + name_off[q], nl);
        pat[11 + nl] = ':';
        pat[11 + nl + 1] = '*';
        uint64_t c = 0;
#pragma omp parallel for reduction(+:c) schedule(static) num_threads(threads > 0 ? threads : 1)
        for (uint64_t i = 0; i < n_keys; ++i) {
            c += (uint64_t)glob_rec(pat, pl, blob + off[i], (long)(off[i + 1] - off[i]));
        }
        counts[q] = c;
        free(pat);
    }
}

/* Same, over fixed-width records (each record scanned as a rec_bytes-long
 * key).  For literal queue names a pattern "literal*" gives the same verdict
 * on a zero-padded/truncated head as on the full key as long as the literal
 * fits in the record and contains no NUL.                                   */
KAS_EXPORT void kas_oracle_match_counts_heads(const uint8_t *heads, uint64_t n_keys,
                                              uint32_t rec_bytes,
                                              const uint8_t *names, const uint32_t *name_off,
                                              uint32_t n_queues, uint64_t *counts,
                                              int threads)
{
    for (uint32_t q = 0; q < n_queues; ++q) {
        uint32_t nl = name_off[q + 1] - name_off[q];
        uint32_t pl = 11 + nl + 2;
        uint8_t *pat = (uint8_t *)malloc(pl);
        memcpy(pat, "processing-", 11);
        memcpy(pat + 11, names + name_off[q], nl);
        pat[11 + nl] = ':';
        pat[11 + nl + 1] = '*';
        uint64_t c = 0;
#pragma omp parallel for reduction(+:c) schedule(static) num_threads(threads > 0 ? threads : 1)
        for (uint64_t i = 0; i < n_keys; ++i) {
            c += (uint64_t)glob_rec(pat, pl, heads + i * rec_bytes, (long)rec_bytes);
        }
        counts[q] = c;
        free(pat);
    }
}

/* ---- Decision (autoscaler.py:197-219) ----------------------------------- */
static inline int64_t floordiv64(int64_t a, int64_t b)
{
    int64_t q = a / b;
    if ((a % b != 0) && ((a < 0) != (b < 0))) q -= 1;   /* Python // */
    return q;
}

static inline int64_t clip64(int64_t d, int64_t mn, int64_t mx, int64_t cur)
{
    if (d > mx) d = mx;                 /* :200 */
    else if (d < mn) d = mn;            /* :202 */
    if (0 < d && d < cur) d = cur;      /* :207 */
    return d;
}

KAS_EXPORT int64_t kas_oracle_clip_pod_count(int64_t desired, int64_t mn, int64_t mx, int64_t cur)
{
    return clip64(desired, mn, mx, cur);
}

/* kpp == 0 is the caller's problem (the reference raises ZeroDivisionError). */
KAS_EXPORT int64_t kas_oracle_get_desired_pods(int64_t tally, int64_t kpp, int64_t mn,
                                               int64_t mx, int64_t cur)
{
    return clip64(floordiv64(tally, kpp), mn, mx, cur);
}

/* ---- Replay over a trace (autoscaler.py:247-260 + :227-237) -------------
 * llen [T][Q] tick-major.  inprog_mode 0: none, 1: [Q], 2: [T][Q].
 * group_off NULL -> every queue is its own resource (R == Q).
 * policies [S][3] = kpp, min, max.  Outputs [R][S].
 * dense (optional) [S][T][R] int32 = replicas after each tick.
 *
 * startup_delay D > 0 is the `only_running` reading of a deployment
 * (autoscaler.py:176-177: current pods = status.available_replicas) with a
 * modelled pod start-up time: a replica is available once it has been part of
 * spec.replicas for D consecutive ticks, a removed replica is gone at once, so
 *     available[t+1] = min(spec[t-D], ..., spec[t])      (spec[<0] = cur0).
 * The controller reads `available`, decides, and patches spec.replicas when
 * desired != available (:227) -- also when spec already holds that value.
 * D == 0 is the plain replay (available == spec).  pod_ticks / dense / final
 * are spec.replicas after the tick.                                          */
#define KAS_ORACLE_MAX_DELAY 64
KAS_EXPORT void kas_oracle_replay_delay(const int32_t *llen, const int32_t *inprog, int inprog_mode,
                                        uint32_t T, uint64_t Q,
                                        const uint64_t *group_off, uint64_t R,
                                        const int32_t *cur0,
                                        const int32_t *policies, uint32_t S,
                                        int64_t *pod_ticks, int64_t *n_patches, int32_t *final_cur,
                                        int32_t *dense, const uint8_t *patch_fail /* [T][R] or NULL */,
                                        uint32_t startup_delay, int threads)
{
    if (startup_delay > KAS_ORACLE_MAX_DELAY) startup_delay = KAS_ORACLE_MAX_DELAY;
#pragma omp parallel for schedule(static) num_threads(threads > 0 ? threads : 1)
    for (uint64_t r = 0; r < R; ++r) {
        uint64_t q0 = group_off ? group_off[r] : r;
        uint64_t q1 = group_off ? group_off[r + 1] : r + 1;
        for (uint32_t s = 0; s < S; ++s) {
            int64_t kpp = policies[3 * s], mn = policies[3 * s + 1], mx = policies[3 * s + 2];
            int64_t spec = cur0 ? cur0[r] : 0;      /* spec.replicas / spec.parallelism */
            int64_t avail = spec;                   /* status.available_replicas */
            int64_t hist[KAS_ORACLE_MAX_DELAY];     /* spec of the last D ticks */
            for (uint32_t i = 0; i < startup_delay; ++i) hist[i] = spec;
            int64_t pt = 0, np = 0;
            for (uint32_t t = 0; t < T; ++t) {
                int64_t cur = avail;                                       /* :252 (with only_running) */
                int64_t sum = 0;
                for (uint64_t q = q0; q < q1; ++q) {
                    int64_t tally = llen[(uint64_t)t * Q + q];
                    if (inprog_mode == 1) tally += inprog[q];
                    else if (inprog_mode == 2) tally += inprog[(uint64_t)t * Q + q];
                    sum += clip64(floordiv64(tally, kpp), mn, mx, cur);   /* :255-257 */
                }
                int64_t desired = clip64(sum, mn, mx, cur);               /* :259 */
                if (desired != cur) {                                      /* :227 */
                    np += 1;                                               /* the patch call is made */
                    /* ... and, if it raises, swallowed: replicas stay (:267-273) */
                    if (!(patch_fail && patch_fail[(uint64_t)t * R + r])) spec = desired;
                }
                avail = spec;
                for (uint32_t i = 0; i < startup_delay; ++i)
                    if (hist[i] < avail) avail = hist[i];
                if (startup_delay) hist[t % startup_delay] = spec;
                pt += spec;
                if (dense) dense[((uint64_t)s * T + t) * R + r] = (int32_t)spec;
            }
            pod_ticks[r * S + s] = pt;
            n_patches[r * S + s] = np;
            final_cur[r * S + s] = (int32_t)spec;
        }
    }
}

/* ---- Time to the first pod after a burst (report on top of the replay) -----
 * Same loop as kas_oracle_replay_delay for ONE policy, with the bookkeeping of
 * oracle.first_pod_latency (see there for the definition).  out [R][5] int64:
 * n_served, n_unserved, wait_sum, wait_max, pending.                         */
KAS_EXPORT void kas_oracle_first_pod(const int32_t *llen, const int32_t *inprog, int inprog_mode,
                                     uint32_t T, uint64_t Q,
                                     const uint64_t *group_off, uint64_t R,
                                     const int32_t *cur0, const int32_t *policy /* kpp, min, max */,
                                     int64_t *out, const uint8_t *patch_fail /* [T][R] or NULL */,
                                     uint32_t startup_delay, int threads)
{
    if (startup_delay > KAS_ORACLE_MAX_DELAY) startup_delay = KAS_ORACLE_MAX_DELAY;
    const int64_t kpp = policy[0], mn = policy[1], mx = policy[2];
#pragma omp parallel for schedule(static) num_threads(threads > 0 ? threads : 1)
    for (uint64_t r = 0; r < R; ++r) {
        uint64_t q0 = group_off ? group_off[r] : r;
        uint64_t q1 = group_off ? group_off[r + 1] : r + 1;
        int64_t spec = cur0 ? cur0[r] : 0, avail = spec;
        int64_t hist[KAS_ORACLE_MAX_DELAY];
        for (uint32_t i = 0; i < startup_delay; ++i) hist[i] = spec;
        int64_t waiting = -1, prev_tally = 0;
        int64_t served = 0, unserved = 0, wsum = 0, wmax = 0;
        for (uint32_t t = 0; t < T; ++t) {
            int64_t cur = avail, sum = 0, total = 0;
            for (uint64_t q = q0; q < q1; ++q) {
                int64_t tally = llen[(uint64_t)t * Q + q];
                if (inprog_mode == 1) tally += inprog[q];
                else if (inprog_mode == 2) tally += inprog[(uint64_t)t * Q + q];
                total += tally;
                sum += clip64(floordiv64(tally, kpp), mn, mx, cur);
            }
            int64_t desired = clip64(sum, mn, mx, cur);
            if (desired != cur && !(patch_fail && patch_fail[(uint64_t)t * R + r])) spec = desired;
            avail = spec;
            for (uint32_t i = 0; i < startup_delay; ++i)
                if (hist[i] < avail) avail = hist[i];
            if (startup_delay) hist[t % startup_delay] = spec;
            if (waiting < 0 && total > 0 && prev_tally <= 0 && cur <= 0) waiting = t;
            if (waiting >= 0) {
                if (avail > 0) {
                    int64_t w = (int64_t)t - waiting;
                    served += 1; wsum += w; if (w > wmax) wmax = w;
                    waiting = -1;
                } else if (total <= 0) {
                    unserved += 1;
                    waiting = -1;
                }
            }
            prev_tally = total;
        }
        out[5 * r + 0] = served;
        out[5 * r + 1] = unserved;
        out[5 * r + 2] = wsum;
        out[5 * r + 3] = wmax;
        out[5 * r + 4] = waiting >= 0 ? (int64_t)T - waiting : 0;
    }
}

KAS_EXPORT void kas_oracle_replay(const int32_t *llen, const int32_t *inprog, int inprog_mode,
                                  uint32_t T, uint64_t Q,
                                  const uint64_t *group_off, uint64_t R,
                                  const int32_t *cur0,
                                  const int32_t *policies, uint32_t S,
                                  int64_t *pod_ticks, int64_t *n_patches, int32_t *final_cur,
                                  int32_t *dense, const uint8_t *patch_fail /* [T][R] or NULL */,
                                  int threads)
{
    kas_oracle_replay_delay(llen, inprog, inprog_mode, T, Q, group_off, R, cur0, policies, S,
                            pod_ticks, n_patches, final_cur, dense, patch_fail, 0u, threads);
}

/* ---- Twin of the library's synthetic trace generator --------------------
 * (restated from the written spec in DESIGN.md, used to rebuild any sampled
 * queue's full trajectory when the trace only exists in HBM).               */
static inline uint64_t mix64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

KAS_EXPORT int32_t kas_oracle_trace_value(uint64_t seed, uint64_t q, uint32_t t)
{
    uint64_t epoch = t >> 6;
    uint64_t he = mix64(mix64(seed ^ (q * 0xD1342543DE82EF95ULL)) + epoch);
    uint64_t ht = mix64(he ^ ((uint64_t)t * 0x2545F4914F6CDD1DULL));
    uint32_t active = (uint32_t)(he % 100u) < 45u;
    uint32_t level = (uint32_t)((he >> 8) % 400u);
    uint32_t jitter = (uint32_t)(ht & 0xFFu);
    if (active) return (int32_t)((level * jitter) >> 7);
    if (jitter < 10u) return (int32_t)((ht >> 8) & 3u);
    return 0;
}

KAS_EXPORT void kas_oracle_trace_fill(uint64_t seed, uint32_t T, uint64_t q_first,
                                      uint64_t n_q, int32_t *out /* [T][n_q] */, int threads)
{
#pragma omp parallel for schedule(static) num_threads(threads > 0 ? threads : 1)
    for (uint32_t t = 0; t < T; ++t)
        for (uint64_t j = 0; j < n_q; ++j)
            out[(uint64_t)t * n_q + j] = kas_oracle_trace_value(seed, q_first + j, t);
}

/* ---- Twin of the library's synthetic keyspace generator ------------------
 * Restated from the written spec, so that the CPU arm
 * of the bench and the tests can make their inputs without loading the product
 * library.  Key i of a keyspace (seed, names, long_tail):
 *   h0 = mix64(seed ^ mix64(i)); h1 = mix64(h0); cls = h0 % 100;
 *   name = names[(h0 >> 8) % n]; sl = (h0 >> 32) % 25
 *          (+ (h1 >> 8) % 680 when long_tail and (h1 & 0xFF) < 13);
 *   suffix char k = alphabet64[(mix64(h1 + k/10) >> 6*(k%10)) & 63].          */
#define KAS_ORACLE_KEY_CAP 4096
static const char kas_alphabet64[65] =
    "0123456789abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ-:";

static uint32_t put_bytes(uint8_t *out, uint32_t at, const void *src, uint32_t n)
{
    if (at + n > KAS_ORACLE_KEY_CAP) n = KAS_ORACLE_KEY_CAP - at;
    memcpy(out + at, src, n);
    return at + n;
}

static uint32_t put_suffix(uint8_t *out, uint32_t at, uint64_t h1, uint32_t sl)
{
    uint64_t word = 0;
    for (uint32_t k = 0; k < sl && at < KAS_ORACLE_KEY_CAP; ++k) {
        if (k % 10 == 0) word = mix64(h1 + k / 10);
        out[at++] = (uint8_t)kas_alphabet64[(word >> (6 * (k % 10))) & 63u];
    }
    return at;
}

/* Writes key i into out (capacity KAS_ORACLE_KEY_CAP) and returns its length. */
static uint32_t synth_key(uint64_t seed, uint64_t i, const uint8_t *names, const uint32_t *name_off,
                          uint32_t n_names, int long_tail, uint8_t *out)
{
    const uint64_t h0 = mix64(seed ^ mix64(i));
    const uint64_t h1 = mix64(h0);
    const uint32_t cls = (uint32_t)(h0 % 100u);
    const uint32_t j = (uint32_t)((h0 >> 8) % n_names);
    uint32_t sl = (uint32_t)((h0 >> 32) % 25u);
    if (long_tail && (h1 & 0xFFu) < 13u) sl += (uint32_t)((h1 >> 8) % 680u);
    const uint8_t *nm = names + name_off[j];
    const uint32_t nl = name_off[j + 1] - name_off[j];
    uint32_t at = 0;
    if (cls < 50) {            /* processing-<name>:<suffix>      counts for <name> */
        at = put_bytes(out, at, "processing-", 11); at = put_bytes(out, at, nm, nl);
        at = put_bytes(out, at, ":", 1); at = put_suffix(out, at, h1, sl);
    } else if (cls < 56) {     /* processing-<name>x:<suffix>     near miss */
        at = put_bytes(out, at, "processing-", 11); at = put_bytes(out, at, nm, nl);
        at = put_bytes(out, at, "x:", 2); at = put_suffix(out, at, h1, sl);
    } else if (cls < 61) {     /* processing-<name>               no colon */
        at = put_bytes(out, at, "processing-", 11); at = put_bytes(out, at, nm, nl);
    } else if (cls < 66) {     /* xprocessing-<name>:<suffix> */
        at = put_bytes(out, at, "xprocessing-", 12); at = put_bytes(out, at, nm, nl);
        at = put_bytes(out, at, ":", 1); at = put_suffix(out, at, h1, sl);
    } else if (cls < 71) {     /* processing-<name minus last byte>:<suffix> */
        at = put_bytes(out, at, "processing-", 11); at = put_bytes(out, at, nm, nl ? nl - 1 : 0);
        at = put_bytes(out, at, ":", 1); at = put_suffix(out, at, h1, sl);
    } else if (cls < 74) {     /* <name>                          the queue's own list key */
        at = put_bytes(out, at, nm, nl);
    } else if (cls < 77) {     /* processing-<name>:              empty suffix, matches */
        at = put_bytes(out, at, "processing-", 11); at = put_bytes(out, at, nm, nl);
        at = put_bytes(out, at, ":", 1);
    } else if (cls < 80) {     /* Processing-<name>:<suffix>      case differs */
        at = put_bytes(out, at, "Processing-", 11); at = put_bytes(out, at, nm, nl);
        at = put_bytes(out, at, ":", 1); at = put_suffix(out, at, h1, sl);
    } else {                   /* <name>:<suffix>:<hex16(h1)>     a job hash */
        static const char hexd[] = "0123456789abcdef";
        at = put_bytes(out, at, nm, nl); at = put_bytes(out, at, ":", 1);
        at = put_suffix(out, at, h1, sl); at = put_bytes(out, at, ":", 1);
        for (int k = 0; k < 16 && at < KAS_ORACLE_KEY_CAP; ++k) out[at++] = (uint8_t)hexd[(h1 >> (4 * k)) & 15u];
    }
    return at;
}

/* Keys first_key .. first_key+n_keys-1 as zero-padded heads of rec_bytes. */
KAS_EXPORT void kas_oracle_synth_heads(uint64_t seed, uint64_t first_key, uint64_t n_keys,
                                       const uint8_t *names, const uint32_t *name_off,
                                       uint32_t n_names, int long_tail, uint32_t rec_bytes,
                                       uint8_t *out_heads, int threads)
{
#pragma omp parallel num_threads(threads > 0 ? threads : 1)
    {
        uint8_t key[KAS_ORACLE_KEY_CAP];
#pragma omp for schedule(static)
        for (uint64_t i = 0; i < n_keys; ++i) {
            uint32_t len = synth_key(seed, first_key + i, names, name_off, n_names, long_tail, key);
            uint32_t n = len < rec_bytes ? len : rec_bytes;
            memcpy(out_heads + i * rec_bytes, key, n);
            memset(out_heads + i * rec_bytes + n, 0, rec_bytes - n);
        }
    }
}

/* The same keys as a blob + offsets.  out_off [n_keys+1] is always written;
 * out_blob may be NULL to size the blob (returns the bytes needed).          */
KAS_EXPORT uint64_t kas_oracle_synth_blob(uint64_t seed, uint64_t first_key, uint64_t n_keys,
                                          const uint8_t *names, const uint32_t *name_off,
                                          uint32_t n_names, int long_tail,
                                          uint8_t *out_blob, uint64_t *out_off, int threads)
{
#pragma omp parallel num_threads(threads > 0 ? threads : 1)
    {
        uint8_t key[KAS_ORACLE_KEY_CAP];
#pragma omp for schedule(static)
        for (uint64_t i = 0; i < n_keys; ++i)
            out_off[i + 1] = synth_key(seed, first_key + i, names, name_off, n_names, long_tail, key);
    }
    out_off[0] = 0;
    for (uint64_t i = 0; i < n_keys; ++i) out_off[i + 1] += out_off[i];
    if (out_blob) {
#pragma omp parallel num_threads(threads > 0 ? threads : 1)
        {
            uint8_t key[KAS_ORACLE_KEY_CAP];
#pragma omp for schedule(static)
            for (uint64_t i = 0; i < n_keys; ++i) {
                uint32_t len = synth_key(seed, first_key + i, names, name_off, n_names, long_tail, key);
                memcpy(out_blob + out_off[i], key, len);
            }
        }
    }
    return out_off[n_keys];
}

KAS_EXPORT int kas_oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
