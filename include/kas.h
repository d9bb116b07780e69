/*
 * kas.h -- C ABI of libkas.so: the B200 (sm_100a) implementation of
 * kiosk-autoscaler's queue tally and replica decision, batched over a packed
 * keyspace snapshot and a queue-depth trace.
 *
 * The reference has no FFI; its boundary is the Python object API
 * (autoscaler/__init__.py:30-32).  This header is what the drop-in Python
 * package (kiosk-autoscaler_b200/autoscaler) binds with ctypes; every entry
 * point names the reference lines whose work it takes over.  Paths are
 * relative to the reference checkout; "autoscaler.py" means
 * autoscaler/autoscaler.py.
 *
 * Conventions
 *   - plain C types, no torch / C++ types; return 0 (KAS_OK) or a negative
 *     kas_status; nothing throws across the boundary.
 *   - the CALLER owns every device buffer ("d_" prefix) and every host buffer;
 *     the library allocates no device memory.  Its only global mutable state is
 *     a per-thread copy of the last CUDA error string, a per-thread cache of the
 *     current device's SM count, the kas_set_reserved_sms setting, the NCCL
 *     communicators of kas_comm_init, and -- for kas_patch_event_offsets_static
 *     alone -- one fixed 64 KB device array per GPU plus the per-device cache of
 *     its address (see that entry point for the sharing rules).
 *   - all device work is enqueued on the given stream and is asynchronous with
 *     respect to the host; `stream` is a cudaStream_t passed as void*.
 *   - integer results are bit-exact with the reference's Python arithmetic
 *     inside the stated domains; outside them the call is rejected, never
 *     approximated.  There is no CPU fallback: without a CUDA device every
 *     entry point that launches work returns KAS_ERR_CUDA.
 */
#ifndef KAS_H_
#define KAS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KAS_VERSION 103 /* 0.1.3: kas_replay_args.d_resume, kas_totals_chain */

#if defined(__GNUC__)
#define KAS_API __attribute__((visibility("default")))
#else
#define KAS_API
#endif

typedef enum kas_status {
    KAS_OK = 0,
    KAS_ERR_INVALID = -1,       /* null pointer, bad size, bad flag            */
    KAS_ERR_UNSUPPORTED = -2,   /* outside the accelerated domain (see call)    */
    KAS_ERR_CUDA = -3,          /* a CUDA call failed; see kas_last_cuda_error  */
    KAS_ERR_ALIGNMENT = -4,     /* device pointer not aligned as required       */
    KAS_ERR_ZERO_DIVISION = -5, /* keys_per_pod == 0 (reference: ZeroDivisionError,
                                   autoscaler.py:217)                           */
    KAS_ERR_GLOB_META = -6,     /* queue name holds * ? [ \ : a live glob in the
                                   reference (autoscaler.py:69), which the
                                   prefix table cannot hold: use kas_glob_count  */
    KAS_ERR_TOO_SMALL = -7,     /* caller's output buffer is too small          */
    KAS_ERR_STATIC_LIMIT = -8   /* more resources than the library's built-in
                                   scratch covers (kas_patch_event_offsets_static) */
} kas_status;

KAS_API int kas_version(void);
KAS_API const char *kas_strerror(int status);
/* Text of the last CUDA error seen by the calling thread ("" if none). */
KAS_API const char *kas_last_cuda_error(void);
/* Leave n SMs (0..64, default 0) out of the persistent grid of kas_match_count, process-wide.
 * The match kernel otherwise holds every SM for its whole run, so a small kernel on another
 * stream -- the all-reduce of the previous step's counts -- can only start in the gap between two
 * launches and delays the next one by its own latency.  With one SM set aside the collective runs
 * beside the match kernel (cost: 1/148 of its throughput).  No counterpart in the reference. */
KAS_API int kas_set_reserved_sms(int n);
/* SM count and compute capability of the current device. */
KAS_API int kas_device_info(int *sm_count, int *cc_major, int *cc_minor);

/* ------------------------------------------------------------------------
 * Key-snapshot ingest: variable-length keys -> fixed-width "heads".
 * Replaces the client side of SCAN iteration (autoscaler.py:70-71, which
 * pulls every key over TCP once per queue): the keyspace is taken once and
 * the first rec_bytes bytes of every key are stored zero-padded, rec_bytes a
 * multiple of 16, records contiguous, so the match kernel streams them.
 * ---------------------------------------------------------------------- */

/* Smallest record width (multiple of 16) that holds "processing-" + name + ":"
 * for a name of max_name_len bytes (autoscaler.py:69). */
KAS_API uint32_t kas_rec_bytes_for(uint32_t max_name_len);

/* Host packer.  Key i is blob[off[i] .. off[i+1]-sep_bytes): sep_bytes is 0 for
 * a plain blob+offsets store and 1 for a newline-separated dump indexed by
 * kas_index_lines_host.  out is n_keys*rec_bytes. */
KAS_API int kas_pack_heads_host(const uint8_t *blob, const uint64_t *off, uint64_t n_keys,
                        uint32_t sep_bytes, uint32_t rec_bytes, uint8_t *out, int threads);

/* Index a newline-separated key dump (`redis-cli --scan` output): off[i] is the
 * start of line i, and one more entry closes the last line (blob_len+1 when the
 * dump does not end in '\n', as if it did), so key i is
 * blob[off[i] .. off[i+1]-1) -- pack with sep_bytes = 1.  *n_keys gets the
 * count.  A trailing '\r' is not stripped (keys are binary-safe bytes).  off
 * may be NULL to count only; off_cap is its capacity in elements
 * (KAS_ERR_TOO_SMALL if n_keys+1 does not fit). */
KAS_API int kas_index_lines_host(const uint8_t *blob, uint64_t blob_len, uint64_t *off,
                         uint64_t off_cap, uint64_t *n_keys);

/* Keys of a Redis RDB dump (the keyspace at rest instead of a SCAN over TCP, autoscaler.py:67-71).
 * Walks the dump once, skipping values by type: key i is out_blob[out_off[i] .. out_off[i+1])
 * (integer-encoded keys as their decimal text, LZF-compressed keys decompressed), out_type[i] its
 * RDB value type, out_db[i] its database, out_list_len[i] the number of elements when the key
 * holds a list -- what LLEN would answer (autoscaler.py:67) -- or UINT64_MAX when a list node's
 * 16-bit counter is saturated; 0 for other types.  Every output pointer may be NULL (call once with
 * NULLs to size the buffers: *n_keys, *blob_len); out_off has n_keys+1 entries (off_cap is its
 * capacity in elements), blob_cap the capacity of out_blob.  RDB versions 1-12; value types 0-5, 7,
 * 9-21.  KAS_ERR_INVALID: not an RDB file, truncated or malformed; KAS_ERR_UNSUPPORTED: an opcode
 * or value type this reader does not know (the dump is never half-read); KAS_ERR_TOO_SMALL. */
KAS_API int kas_rdb_keys_host(const uint8_t *rdb, uint64_t rdb_len, uint8_t *out_blob, uint64_t blob_cap,
                              uint64_t *out_off, uint64_t off_cap, uint8_t *out_type, uint32_t *out_db,
                              uint64_t *out_list_len, uint64_t *n_keys, uint64_t *blob_len);

/* Device packer for a blob already resident in HBM.  d_heads must be 16-byte
 * aligned.  Reads every key's first min(len, rec_bytes) bytes. */
KAS_API int kas_pack_heads(const uint8_t *d_blob, const uint64_t *d_off, uint64_t n_keys,
                   uint32_t sep_bytes, uint32_t rec_bytes, uint8_t *d_heads, void *stream);

/* ------------------------------------------------------------------------
 * Queue-prefix table.  Built on the host from the queue names held by
 * Autoscaler.redis_keys (autoscaler.py:52), copied to the device by the
 * caller, staged in shared memory by the match kernel.
 * Names must be unique, free of NUL and of the glob metacharacters * ? [ \
 * (KAS_ERR_GLOB_META otherwise), at most rec_bytes-12 bytes long, and at most
 * KAS_MAX_PREFIXES of them; the table (which also holds every cut of a name at
 * an inner ':') must fit 200 KB of shared memory, else KAS_ERR_UNSUPPORTED /
 * a size of 0.
 * ---------------------------------------------------------------------- */
#define KAS_MAX_PREFIXES 1024u
#define KAS_MAX_REC_BYTES 64u

KAS_API size_t kas_prefix_table_bytes(uint32_t n_prefix, uint32_t rec_bytes);
/* names: concatenated name bytes; name q is names[name_off[q] .. name_off[q+1]). */
KAS_API int kas_prefix_table_build(const uint8_t *names, const uint32_t *name_off, uint32_t n_prefix,
                           uint32_t rec_bytes, void *host_tbl, size_t tbl_bytes);

/* ------------------------------------------------------------------------
 * K1: prefix match and count.
 * Replaces, for all queues in ONE pass over the keyspace,
 *     scan = redis_client.scan_iter(match='processing-{}:*'.format(q), count=1000)
 *     num_in_progress = len([x for x in scan])          (autoscaler.py:69-71)
 * d_counts[q] = number of records that start with "processing-" + name_q + ":".
 * Every queue is counted independently, so a key counts for several queues
 * when one such prefix is a prefix of another (as in the reference, where
 * every queue runs its own SCAN).  No dedup (autoscaler.py:71).
 * d_heads: n_keys records of rec_bytes, 16-byte aligned.  d_counts: uint64
 * [n_prefix].  kas_match_count overwrites d_counts; kas_match_count_accumulate
 * adds to it (for snapshots streamed through HBM in chunks).
 * ---------------------------------------------------------------------- */
KAS_API int kas_match_count(const void *d_heads, uint64_t n_keys, uint32_t rec_bytes,
                    const void *d_prefix_tbl, uint32_t n_prefix, uint64_t *d_counts,
                    void *stream);
KAS_API int kas_match_count_accumulate(const void *d_heads, uint64_t n_keys, uint32_t rec_bytes,
                               const void *d_prefix_tbl, uint32_t n_prefix, uint64_t *d_counts,
                               void *stream);

/* ------------------------------------------------------------------------
 * General glob match and count, for queue names that hold * ? [ \ .
 * autoscaler.py:69 formats the queue name into the SCAN pattern unescaped, so
 * such a name is a live glob in the reference; this call counts, for each
 * pattern, the keys it matches with the semantics of `SCAN ... MATCH`
 * (autoscaler.py:70-71).  Patterns are whole globs, e.g. "processing-pre*:*".
 * Keys are a blob + offsets in HBM (key i = blob[off[i] .. off[i+1]-sep_bytes)),
 * because a glob may need all bytes of a key.  d_patterns holds the pattern
 * bytes back to back, pattern q = [d_pat_off[q], d_pat_off[q+1]); pattern_bytes
 * = d_pat_off[n_patterns] (at most 16 KB, at most KAS_MAX_PREFIXES patterns).
 * d_counts: uint64 [n_patterns], overwritten.
 * ---------------------------------------------------------------------- */
KAS_API int kas_glob_count(const uint8_t *d_blob, const uint64_t *d_off, uint64_t n_keys,
                           uint32_t sep_bytes, const uint8_t *d_patterns, const uint32_t *d_pat_off,
                           uint32_t n_patterns, uint32_t pattern_bytes, uint64_t *d_counts,
                           void *stream);

/* ------------------------------------------------------------------------
 * Element-wise decision, the scalar API's arithmetic.
 * mode 0: out[i] = clip_pod_count(a[i] // kpp[i], mn[i], mx[i], cur[i])
 *         = Autoscaler.get_desired_pods with redis_keys[key] == a[i]
 *           (autoscaler.py:215-219); Python floor division; any kpp[i] == 0
 *           -> the whole call is rejected with KAS_ERR_ZERO_DIVISION and
 *           d_status (int32[1], device) is set to 1.
 * mode 1: out[i] = Autoscaler.clip_pod_count(a[i], mn[i], mx[i], cur[i])
 *           (autoscaler.py:197-213); d_kpp is ignored and may be NULL.
 * All arrays int64 [n] on the device.  Domain: every int64 value.
 * The zero-divisor check runs on the device; read d_status after syncing.
 * ---------------------------------------------------------------------- */
KAS_API int kas_decide(const int64_t *d_a, const int64_t *d_kpp, const int64_t *d_mn,
               const int64_t *d_mx, const int64_t *d_cur, uint64_t n, int mode,
               int64_t *d_out, int32_t *d_status, void *stream);

/* ------------------------------------------------------------------------
 * K2: fused tally + desired pods + clamp + hysteresis replay over a trace.
 * Replaces T iterations of Autoscaler.scale for R resources and S policies
 * (autoscaler.py:244-260), with a well-behaved cluster standing in for
 * get_current_pods / scale_resource: what tick t patches is what tick t+1
 * reads (autoscaler.py:179,189 <- :232,236), no patch when nothing changes
 * (:227-228).  Per tick and resource r with queues q in
 * [group_off[r], group_off[r+1]):
 *     tally_q  = llen[t][q] + inprog(t,q)                     (:73)
 *     d_q      = clip(tally_q // kpp, min, max, cur)          (:215-219)
 *     desired  = clip(sum_q d_q, min, max, cur)               (:254-260)
 *     n_patches += desired != cur;  cur = desired             (:227-237)
 *     pod_ticks += cur
 * ---------------------------------------------------------------------- */
typedef struct kas_policy {
    int32_t keys_per_pod; /* KEYS_PER_POD, scale.py:92; != 0, may be negative  */
    int32_t min_pods;     /* MIN_PODS,     scale.py:90                          */
    int32_t max_pods;     /* MAX_PODS,     scale.py:91                          */
} kas_policy;

typedef struct kas_totals {
    int64_t pod_ticks;     /* sum over ticks of replicas after the tick         */
    uint32_t n_patches;    /* ticks on which scale_resource would have patched  */
    int32_t final_current; /* replicas after the last tick                      */
} kas_totals;

#define KAS_TRACE_INT32 0
#define KAS_TRACE_UINT16 1 /* depths below 65536: half the bytes of the HBM-bound replay */

typedef struct kas_replay_args {
    const void *d_llen;         /* [T][Q] tick-major LLEN trace (autoscaler.py:67),
                                   int32 or uint16 elements (trace_dtype)          */
    const void *d_inprog;       /* SCAN-count term: NULL, int32 [Q], or [T][Q] of
                                   the trace's element type                        */
    int32_t inprog_mode;        /* 0 none, 1 per queue [Q], 2 per tick [T][Q]      */
    uint32_t n_ticks;           /* T                                               */
    uint64_t n_queues;          /* Q                                               */
    uint64_t n_resources;       /* R (== Q when d_group_off is NULL)               */
    const uint64_t *d_group_off;/* [R+1] CSR offsets into the queue axis, or NULL.
                                   Expected non-decreasing with group_off[R] <= Q
                                   (a resource's queues are one contiguous run).
                                   Offsets that go back or overlap are not an error
                                   as long as every one is <= Q: a range that goes
                                   back is empty, overlapping ranges are each summed;
                                   the kernels that stage runs in shared memory check
                                   this per block and read such blocks element by
                                   element.  An offset beyond Q reads out of bounds. */
    const int32_t *d_cur0;      /* [R] replicas before tick 0, or NULL for zeros   */
    const kas_policy *policies; /* HOST pointer, [S]                               */
    uint32_t n_policies;        /* S                                               */
    kas_totals *d_totals;       /* [R][S] out, or NULL                             */
    int64_t *d_sums;            /* [S][3] out: sum over resources of pod_ticks,
                                   n_patches, final_current (what the box-wide
                                   allreduce carries), or NULL                     */
    int32_t *d_dense;           /* [S][T][R] replicas after each tick, or NULL     */
    int32_t trace_dtype;        /* KAS_TRACE_INT32 (0) or KAS_TRACE_UINT16          */
    const uint8_t *d_patch_fail;/* [T][R] or NULL: non-zero = the patch call of that
                                   tick and resource raises ApiException, which
                                   scale() logs and swallows (autoscaler.py:267-273):
                                   the call counts in n_patches, the replicas stay   */
    uint32_t startup_delay;     /* 0: the controller reads spec.replicas, as scale()
                                   does (autoscaler.py:179,252).  D > 0
                                   (<= KAS_MAX_STARTUP_DELAY): it reads
                                   status.available_replicas, the `only_running`
                                   branch of get_current_pods (autoscaler.py:176-177),
                                   under a pod start-up model: a replica is available
                                   once it has been in the spec for D ticks in a row
                                   and gone as soon as the spec drops it,
                                     available[t+1] = min(spec[t-D..t]), spec[<0] = cur0.
                                   A patch is issued whenever desired != available
                                   (:227), also when the spec already holds the value.
                                   Outputs stay spec.replicas after the tick.  Only
                                   deployments have this reading (a job always reports
                                   spec.parallelism, :185-190): pass 0 for jobs.       */
    const kas_totals *d_resume; /* NULL, or the d_totals [R][S] of an earlier replay of
                                   the SAME resources and policy list: every (resource,
                                   policy) starts from that entry's final_current
                                   instead of d_cur0 (which must then be NULL).  This
                                   is how a trace longer than HBM is replayed in
                                   chunks of ticks: chunk k+1 resumes from chunk k's
                                   totals and kas_totals_chain adds them up.  Not with
                                   startup_delay (the availability window is not part
                                   of the totals); must not alias d_totals.            */
} kas_replay_args;
#define KAS_MAX_STARTUP_DELAY 16u

/* Domain: trace values and cur0 any int32; policies any int32 with
 * keys_per_pod != 0 (KAS_ERR_ZERO_DIVISION otherwise, nothing is launched).
 * d_llen / d_inprog need their element's alignment, d_dense 4 bytes; 16-byte
 * alignment and Q a multiple of 4 (int32) or 8 (uint16) enable the vectorised
 * paths for single-queue resources.                                          */
KAS_API int kas_replay(const kas_replay_args *args, void *stream);

/* Chain the totals of consecutive tick chunks of one replay (checkpoint / resume; the
 * reference keeps no state between ticks but what Kubernetes holds, autoscaler.py:252):
 *   d_acc[i].pod_ticks += d_chunk[i].pod_ticks;  n_patches likewise;
 *   d_acc[i].final_current = d_chunk[i].final_current            for i < n_resources * n_policies,
 * and, when both are given, the same for the [S][3] sums (sums of pod_ticks and n_patches
 * add up, the sum of final_current is the later chunk's). */
KAS_API int kas_totals_chain(kas_totals *d_acc, const kas_totals *d_chunk, uint64_t n_resources,
                             uint32_t n_policies, int64_t *d_sums_acc, const int64_t *d_sums_chunk,
                             void *stream);

/* ------------------------------------------------------------------------
 * Patch-event stream of a replay: the calls Autoscaler.scale_resource would
 * have made (autoscaler.py:221-242).  An event is a tick on which
 * desired != current (:227); the body the reference sends is
 * {'spec': {'replicas': to_pods}} for a deployment (:236) and
 * {'spec': {'parallelism': to_pods}} for a job (:232) -- the resource type
 * only picks the field name, so it stays on the host.
 * Events are grouped by resource, in tick order inside a resource:
 *     events of resource r = d_events[d_event_off[r] .. d_event_off[r+1])
 * Same inputs, domain and policy handling as kas_replay (d_totals / d_sums /
 * d_dense of args are ignored); one policy, args->policies[policy_index].
 *
 *   1. kas_patch_event_offsets  counts per resource and scans: writes
 *      d_event_off[0..R]; d_event_off[R] is the number of events.  Needs
 *      kas_patch_events_scratch_bytes(R) bytes of scratch (0 for R <= 4096).
 *   2. kas_patch_event_emit     writes the events that fit in `capacity`
 *      (every event whose index is below capacity; size the buffer from
 *      d_event_off[R]).
 * ---------------------------------------------------------------------- */
typedef struct kas_patch_event {
    uint32_t tick;      /* tick on which the patch would have been issued   */
    int32_t from_pods;  /* current pods read at that tick (autoscaler.py:252;
                           the available ones when startup_delay > 0)         */
    int32_t to_pods;    /* value patched                                      */
} kas_patch_event;

KAS_API size_t kas_patch_events_scratch_bytes(uint64_t n_resources);
KAS_API int kas_patch_event_offsets(const kas_replay_args *args, uint32_t policy_index,
                                    uint64_t *d_event_off /* [R+1] */, void *d_scratch,
                                    size_t scratch_bytes, void *stream);
KAS_API int kas_patch_event_emit(const kas_replay_args *args, uint32_t policy_index,
                                 const uint64_t *d_event_off, kas_patch_event *d_events,
                                 uint64_t capacity, void *stream);

/* ------------------------------------------------------------------------
 * Time to the first pod after a burst: the third number of a policy report,
 * beside pod-hours (cost) and patch count (churn).  The reference computes
 * none of them; README.md:7 names the purpose ("minimizing GPU costs").  Same
 * inputs, domain and policy handling as kas_patch_event_offsets (one policy,
 * args->policies[policy_index]; patch_fail and startup_delay honoured).
 *
 * Per resource, over the replayed scale loop (autoscaler.py:244-273):
 *   - a burst STARTS on a tick whose tally (sum over the resource's queues of
 *     LLEN + SCAN count, autoscaler.py:73) is > 0 after a tick where it was
 *     <= 0 (before tick 0 it counts as 0), if the controller reads <= 0 pods on
 *     that tick (autoscaler.py:252; the available ones when startup_delay > 0,
 *     :176-177) and no earlier burst is still waiting;
 *   - it is SERVED on the first tick, at or after its start, after which pods
 *     run (the next tick would read > 0): wait = that tick - starting tick;
 *   - it went UNSERVED if the tally is back at <= 0 before that.
 * d_out [R] (or NULL); d_sums int64 [5] (or NULL): sum n_served, sum
 * n_unserved, sum wait_sum, MAX wait_max, sum pending.
 * ---------------------------------------------------------------------- */
typedef struct kas_first_pod {
    uint32_t n_served;   /* bursts that got their first pod                     */
    uint32_t n_unserved; /* bursts whose work left before a pod ran             */
    uint32_t wait_max;   /* longest wait of a served burst, in ticks            */
    uint32_t pending;    /* ticks the last burst has waited when the trace ends */
    uint64_t wait_sum;   /* sum of the waits of served bursts, in ticks         */
} kas_first_pod;

KAS_API int kas_first_pod_latency(const kas_replay_args *args, uint32_t policy_index,
                                  kas_first_pod *d_out /* [R] or NULL */,
                                  int64_t *d_sums /* [5] or NULL */, void *stream);

/* kas_patch_event_offsets without a scratch buffer from the caller.
 * (No counterpart in the reference, which never touches a GPU.)
 *
 * Who needs it: a caller that is not allowed to obtain device memory at the
 * moment it wants the offsets.  Examples are a scaling service that sized all
 * of its HBM at start-up (the trace alone is 80 GB per GPU at the 16M-queue
 * configuration) and runs with a fixed budget afterwards, a plug-in inside a
 * host process whose memory pool belongs to another framework, or a control
 * loop that cannot afford the device-wide synchronisation a cudaMalloc /
 * cudaFree pair brings.  For them the tile sums of the scan live in a fixed
 * 64 KB array inside the library image instead: every GPU that loads libkas
 * holds that array until the process exits.
 *
 * Same result as kas_patch_event_offsets, bit for bit; the caller still passes
 * the trace and d_event_off[R+1].  The call makes no cudaMalloc / cudaFree and
 * uses no allocating library routine.
 *
 * Limits and rules
 *   - at most KAS_STATIC_SCRATCH_MAX_RESOURCES resources; beyond that the call
 *     is refused with KAS_ERR_STATIC_LIMIT and nothing is launched (use
 *     kas_patch_event_offsets with your own scratch).
 *   - the first call on each device looks the array's device address up
 *     (cudaGetSymbolAddress, with that device current) and remembers it; make
 *     that first call before starting any stream capture.
 *   - all calls on a device share the one array, which keeps whatever the
 *     last call left in it.  Nothing has to be cleared -- each call writes every
 *     tile sum before it reads it -- but two calls whose GPU work overlaps
 *     (different streams, or different host threads on the same device) race
 *     on it: keep them on one stream or order them yourself. */
#define KAS_STATIC_SCRATCH_MAX_RESOURCES 33554432ull /* 8192 scan tiles of 4096 */
KAS_API int kas_patch_event_offsets_static(const kas_replay_args *args, uint32_t policy_index,
                                           uint64_t *d_event_off /* [R+1] */, void *stream);

/* ------------------------------------------------------------------------
 * Box-wide totals: the single collective of the path.
 * Keys (kas_match_count) and resources (kas_replay) are split over the GPUs of
 * a box by contiguous range with no exchange on the data path; afterwards one
 * all-reduce(sum) makes every rank hold what ONE Autoscaler would have seen:
 * the uint64 [Q] counts behind redis_keys (autoscaler.py:60-77) and the int64
 * [S][3] sums of the replayed scale loop (autoscaler.py:244-260).  Integer
 * sums do not depend on the order, so the result is the same bits at every
 * world size.  One process per GPU; NCCL is bound at run time (dlopen), and
 * without it these calls return KAS_ERR_UNSUPPORTED.
 *
 *   rank 0:     kas_comm_unique_id(id)        then hand `id` to every rank
 *                                             (any side channel: a file, a
 *                                             TCP store, torch.distributed)
 *   every rank: kas_comm_init(&comm, n_ranks, rank, id)   with its GPU current
 *               kas_allreduce_i64(comm, d_buf, n, stream) in place, async on
 *                                             the stream; uint64 counts are
 *                                             reduced as int64 (same bits)
 *               kas_comm_destroy(comm)
 * The communicators are the library's only long-lived global state.
 * ---------------------------------------------------------------------- */
#define KAS_COMM_ID_BYTES 128
KAS_API int kas_comm_unique_id(uint8_t *id /* [KAS_COMM_ID_BYTES] */);
KAS_API int kas_comm_init(void **comm, int n_ranks, int rank, const uint8_t *id);
KAS_API int kas_allreduce_i64(void *comm, int64_t *d_buf, size_t n, void *stream);
KAS_API int kas_comm_destroy(void *comm);

/* ------------------------------------------------------------------------
 * Synthetic workload generators (no counterpart in the reference, which has
 * no benchmark; used by bench.py and the tests; specified in DESIGN.md so the
 * oracle can rebuild any element independently).
 * ---------------------------------------------------------------------- */
/* d_out[t][j] = f(seed, q_first + j, t), int32 [T][n_q], tick-major. */
KAS_API int kas_synth_trace(uint64_t seed, uint32_t n_ticks, uint64_t q_first, uint64_t n_q,
                    int32_t *d_out, void *stream);

/* Host: write synthetic keys first_key .. first_key+n_keys-1 as packed heads
 * (out_heads, n_keys*rec_bytes, may be NULL) and/or as a blob with offsets
 * (out_blob/out_off, may be NULL; out_off has n_keys+1 entries; blob_cap is
 * the capacity of out_blob; *blob_len gets the bytes needed/written).
 * About half of the keys are "processing-<name>:<suffix>" for one of the
 * given names; the rest are near misses.  long_tail != 0 gives a long-tailed
 * length distribution (keys up to ~700 bytes). */
KAS_API int kas_synth_keys_host(uint64_t seed, uint64_t first_key, uint64_t n_keys,
                        const uint8_t *names, const uint32_t *name_off, uint32_t n_names,
                        int long_tail, uint32_t rec_bytes, uint8_t *out_heads,
                        uint8_t *out_blob, uint64_t blob_cap, uint64_t *out_off,
                        uint64_t *blob_len, int threads);

#ifdef __cplusplus
}
#endif
#endif /* KAS_H_ */
