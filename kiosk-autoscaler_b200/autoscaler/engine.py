"""ctypes binding of libkas.so (include/kas.h) plus the buffer hand-off.

PyTorch is used for exactly three things: allocating device / pinned buffers,
``.data_ptr()``, and the current CUDA stream.  All arithmetic of the tally and
decision path happens in the library's sm_100a kernels; there is no CPU
fallback -- if the library is missing or no CUDA device is present every call
raises.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import numpy as np

KAS_OK = 0
KAS_ERR_INVALID = -1
KAS_ERR_UNSUPPORTED = -2
KAS_ERR_CUDA = -3
KAS_ERR_ALIGNMENT = -4
KAS_ERR_ZERO_DIVISION = -5
KAS_ERR_GLOB_META = -6
KAS_ERR_TOO_SMALL = -7
KAS_ERR_STATIC_LIMIT = -8
TRACE_INT32 = 0
TRACE_UINT16 = 1
STATIC_SCRATCH_MAX_RESOURCES = 33554432
COMM_ID_BYTES = 128

MAX_PREFIXES = 1024
MAX_REC_BYTES = 64
LITERAL = b'processing-'          # autoscaler.py:69 of the reference
GLOB_META = frozenset(b'*?[\\')

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get('KAS_LIB', os.path.join(_PKG_DIR, 'lib', 'libkas.so'))


class KasError(RuntimeError):
    """A libkas call returned a negative status."""

    def __init__(self, status, what, detail=''):
        self.status = status
        msg = '%s failed: %s' % (what, _strerror(status))
        if detail:
            msg += ' [%s]' % detail
        super().__init__(msg)


class GlobQueueNameError(NotImplementedError):
    """A queue name holds a glob metacharacter; in the reference such a name is
    a live glob (autoscaler.py:69) and needs the general glob kernel."""


class Policy(C.Structure):
    _fields_ = [('keys_per_pod', C.c_int32), ('min_pods', C.c_int32), ('max_pods', C.c_int32)]


class Totals(C.Structure):
    _fields_ = [('pod_ticks', C.c_int64), ('n_patches', C.c_uint32), ('final_current', C.c_int32)]


EVENT_DTYPE = np.dtype([('tick', '<u4'), ('from_pods', '<i4'), ('to_pods', '<i4')])
FIRST_POD_DTYPE = np.dtype([('n_served', '<u4'), ('n_unserved', '<u4'), ('wait_max', '<u4'), ('pending', '<u4'),
                            ('wait_sum', '<u8')])
TOTALS_DTYPE = np.dtype([('pod_ticks', '<i8'), ('n_patches', '<u4'), ('final_current', '<i4')])


class ReplayArgs(C.Structure):
    _fields_ = [('d_llen', C.c_void_p), ('d_inprog', C.c_void_p), ('inprog_mode', C.c_int32),
                ('n_ticks', C.c_uint32), ('n_queues', C.c_uint64), ('n_resources', C.c_uint64),
                ('d_group_off', C.c_void_p), ('d_cur0', C.c_void_p),
                ('policies', C.POINTER(Policy)), ('n_policies', C.c_uint32),
                ('d_totals', C.c_void_p), ('d_sums', C.c_void_p), ('d_dense', C.c_void_p),
                ('trace_dtype', C.c_int32), ('d_patch_fail', C.c_void_p),
                ('startup_delay', C.c_uint32), ('d_resume', C.c_void_p)]


MAX_STARTUP_DELAY = 16      # KAS_MAX_STARTUP_DELAY


def _check_delay(startup_delay):
    d = int(startup_delay)
    if not 0 <= d <= MAX_STARTUP_DELAY:
        raise ValueError('startup_delay must be in 0..%d ticks, got %r' % (MAX_STARTUP_DELAY, startup_delay))
    return d


_lib = None
_lib_lock = threading.Lock()

# name -> (restype, argtypes); every symbol include/kas.h declares.
SYMBOLS = {
    'kas_version': (C.c_int, []),
    'kas_strerror': (C.c_char_p, [C.c_int]),
    'kas_last_cuda_error': (C.c_char_p, []),
    'kas_device_info': (C.c_int, [C.POINTER(C.c_int)] * 3),
    'kas_set_reserved_sms': (C.c_int, [C.c_int]),
    'kas_rec_bytes_for': (C.c_uint32, [C.c_uint32]),
    'kas_pack_heads_host': (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32,
                                      C.c_void_p, C.c_int]),
    'kas_index_lines_host': (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                       C.POINTER(C.c_uint64)]),
    'kas_rdb_keys_host': (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                    C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64),
                                    C.POINTER(C.c_uint64)]),
    'kas_pack_heads': (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32,
                                 C.c_void_p, C.c_void_p]),
    'kas_prefix_table_bytes': (C.c_size_t, [C.c_uint32, C.c_uint32]),
    'kas_prefix_table_build': (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32,
                                         C.c_void_p, C.c_size_t]),
    'kas_match_count': (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint32,
                                  C.c_void_p, C.c_void_p]),
    'kas_match_count_accumulate': (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p,
                                             C.c_uint32, C.c_void_p, C.c_void_p]),
    'kas_glob_count': (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p,
                                 C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]),
    'kas_decide': (C.c_int, [C.c_void_p] * 5 + [C.c_uint64, C.c_int, C.c_void_p, C.c_void_p,
                                                C.c_void_p]),
    'kas_replay': (C.c_int, [C.POINTER(ReplayArgs), C.c_void_p]),
    'kas_patch_events_scratch_bytes': (C.c_size_t, [C.c_uint64]),
    'kas_patch_event_offsets': (C.c_int, [C.POINTER(ReplayArgs), C.c_uint32, C.c_void_p, C.c_void_p,
                                          C.c_size_t, C.c_void_p]),
    'kas_patch_event_offsets_static': (C.c_int, [C.POINTER(ReplayArgs), C.c_uint32, C.c_void_p,
                                                 C.c_void_p]),
    'kas_patch_event_emit': (C.c_int, [C.POINTER(ReplayArgs), C.c_uint32, C.c_void_p, C.c_void_p,
                                       C.c_uint64, C.c_void_p]),
    'kas_totals_chain': (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    'kas_first_pod_latency': (C.c_int, [C.POINTER(ReplayArgs), C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    'kas_comm_unique_id': (C.c_int, [C.c_void_p]),
    'kas_comm_init': (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_void_p]),
    'kas_allreduce_i64': (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    'kas_comm_destroy': (C.c_int, [C.c_void_p]),
    'kas_synth_trace': (C.c_int, [C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint64, C.c_void_p,
                                  C.c_void_p]),
    'kas_synth_keys_host': (C.c_int, [C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p,
                                      C.c_uint32, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p,
                                      C.c_uint64, C.c_void_p, C.POINTER(C.c_uint64), C.c_int]),
}


def lib():
    """Load libkas.so once.  Raises if it has not been built: the GPU path is
    the only path."""
    global _lib
    if _lib is None:
        with _lib_lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise ImportError(
                        'libkas.so not found at %s -- build it with '
                        '`python kiosk-autoscaler_b200/build.py`; there is no CPU fallback'
                        % LIB_PATH)
                L = C.CDLL(LIB_PATH)
                for name, (res, args) in SYMBOLS.items():
                    fn = getattr(L, name)
                    fn.restype = res
                    fn.argtypes = args
                _lib = L
    return _lib


def _strerror(status):
    try:
        return lib().kas_strerror(status).decode()
    except Exception:  # pragma: no cover - library missing
        return 'status %d' % status


def check(status, what):
    if status == KAS_OK:
        return
    if status == KAS_ERR_ZERO_DIVISION:
        # same exception type and text as `x // 0` in the reference (autoscaler.py:217)
        raise ZeroDivisionError('integer division or modulo by zero')
    if status == KAS_ERR_GLOB_META:
        raise GlobQueueNameError(
            '%s: queue names containing * ? [ \\ are globs in the reference '
            '(autoscaler.py:69) and are not handled by the prefix kernel' % what)
    detail = lib().kas_last_cuda_error().decode() if status == KAS_ERR_CUDA else ''
    raise KasError(status, what, detail)


def _torch():
    import torch
    return torch


def require_cuda():
    torch = _torch()
    if not torch.cuda.is_available():
        raise RuntimeError('no CUDA device: the tally/decision path runs on the GPU only '
                           '(there is no CPU fallback)')
    return torch


def stream_ptr():
    return C.c_void_p(_torch().cuda.current_stream().cuda_stream)


def ptr(t):
    return C.c_void_p(0 if t is None else t.data_ptr())


def np_ptr(a):
    return C.c_void_p(0 if a is None else a.ctypes.data)


def encode_name(q):
    return q if isinstance(q, (bytes, bytearray)) else str(q).encode('utf-8')


def names_blob(names):
    enc = [bytes(encode_name(q)) for q in names]
    off = np.zeros(len(enc) + 1, dtype=np.uint32)
    if enc:
        off[1:] = np.cumsum([len(e) for e in enc])
    blob = np.frombuffer(b''.join(enc) + b'\0', dtype=np.uint8).copy()
    return blob, off


def rec_bytes_for(names):
    longest = max((len(encode_name(q)) for q in names), default=0)
    return int(lib().kas_rec_bytes_for(longest))


def host_threads():
    return max(1, min(32, os.cpu_count() or 1))


# --------------------------------------------------------------------------
# host-side helpers (no GPU needed)
# --------------------------------------------------------------------------
def build_prefix_table(names, rec_bytes):
    """Host image of the queue-prefix table (uint8 numpy array)."""
    names = list(names)
    if len(names) == 0 or len(names) > MAX_PREFIXES:
        raise ValueError('between 1 and %d queue names are supported, got %d'
                         % (MAX_PREFIXES, len(names)))
    nbytes = lib().kas_prefix_table_bytes(len(names), rec_bytes)
    if nbytes == 0:
        raise KasError(KAS_ERR_UNSUPPORTED, 'kas_prefix_table_bytes')
    blob, off = names_blob(names)
    tbl = np.zeros(nbytes, dtype=np.uint8)
    check(lib().kas_prefix_table_build(np_ptr(blob), np_ptr(off), len(names), rec_bytes,
                                       np_ptr(tbl), nbytes), 'kas_prefix_table_build')
    return tbl


def pack_heads_host(blob, off, rec_bytes, out=None, sep_bytes=0, threads=None):
    """blob+offsets -> uint8 [n][rec_bytes] (numpy, or written into `out`)."""
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    off = np.ascontiguousarray(off, dtype=np.uint64)
    n = len(off) - 1
    if out is None:
        out = np.empty((n, rec_bytes), dtype=np.uint8)
    check(lib().kas_pack_heads_host(np_ptr(blob), np_ptr(off), n, sep_bytes, rec_bytes,
                                    C.c_void_p(out.ctypes.data if isinstance(out, np.ndarray)
                                               else out.data_ptr()),
                                    threads or host_threads()), 'kas_pack_heads_host')
    return out


def index_lines_host(blob):
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    n = C.c_uint64(0)
    check(lib().kas_index_lines_host(np_ptr(blob), blob.size, None, 0, C.byref(n)),
          'kas_index_lines_host')
    off = np.empty(n.value + 1, dtype=np.uint64)
    check(lib().kas_index_lines_host(np_ptr(blob), blob.size, np_ptr(off), off.size, C.byref(n)),
          'kas_index_lines_host')
    return off


RDB_LIST_TYPES = frozenset((1, 10, 14, 18))
RDB_UNKNOWN_LEN = 0xFFFFFFFFFFFFFFFF


def rdb_keys_host(data):
    """Keys of an RDB dump (bytes-like).  Returns dict(blob uint8, off uint64 [n+1], type uint8 [n],
    db uint32 [n], list_len uint64 [n])."""
    rdb = np.frombuffer(data, dtype=np.uint8)
    n, nbytes = C.c_uint64(0), C.c_uint64(0)
    check(lib().kas_rdb_keys_host(np_ptr(rdb), rdb.size, None, 0, None, 0, None, None, None,
                                  C.byref(n), C.byref(nbytes)), 'kas_rdb_keys_host')
    blob = np.empty(max(1, nbytes.value), dtype=np.uint8)
    off = np.empty(n.value + 1, dtype=np.uint64)
    typ = np.empty(max(1, n.value), dtype=np.uint8)
    db = np.empty(max(1, n.value), dtype=np.uint32)
    llen = np.empty(max(1, n.value), dtype=np.uint64)
    check(lib().kas_rdb_keys_host(np_ptr(rdb), rdb.size, np_ptr(blob), blob.size, np_ptr(off), off.size,
                                  np_ptr(typ), np_ptr(db), np_ptr(llen), C.byref(n), C.byref(nbytes)),
          'kas_rdb_keys_host')
    return dict(blob=blob[:nbytes.value], off=off, type=typ[:n.value], db=db[:n.value],
                list_len=llen[:n.value])


def synth_keys_host(seed, first_key, n_keys, names, long_tail=False, rec_bytes=0,
                    want_heads=True, want_blob=False, heads_out=None, threads=None):
    """Synthetic keyspace (spec in DESIGN.md).  Returns (heads|None, blob|None, off|None)."""
    nblob, noff = names_blob(names)
    threads = threads or host_threads()
    heads = None
    if want_heads:
        heads = heads_out if heads_out is not None else np.empty((n_keys, rec_bytes), dtype=np.uint8)
    hp = C.c_void_p(0) if heads is None else C.c_void_p(
        heads.ctypes.data if isinstance(heads, np.ndarray) else heads.data_ptr())
    blob = off = None
    if want_blob:
        need = C.c_uint64(0)
        check(lib().kas_synth_keys_host(seed, first_key, n_keys, np_ptr(nblob), np_ptr(noff),
                                        len(names), int(long_tail), rec_bytes, None, None, 0, None,
                                        C.byref(need), threads), 'kas_synth_keys_host')
        blob = np.empty(max(1, need.value), dtype=np.uint8)
        off = np.empty(n_keys + 1, dtype=np.uint64)
        check(lib().kas_synth_keys_host(seed, first_key, n_keys, np_ptr(nblob), np_ptr(noff),
                                        len(names), int(long_tail), rec_bytes, hp, np_ptr(blob),
                                        blob.size, np_ptr(off), C.byref(need), threads),
              'kas_synth_keys_host')
        blob = blob[:need.value]
    elif want_heads:
        check(lib().kas_synth_keys_host(seed, first_key, n_keys, np_ptr(nblob), np_ptr(noff),
                                        len(names), int(long_tail), rec_bytes, hp, None, 0, None,
                                        None, threads), 'kas_synth_keys_host')
    return heads, blob, off


# --------------------------------------------------------------------------
# device calls
# --------------------------------------------------------------------------
POISON = np.uint64(0xFFFFFFFFFFFFFFFF)


def match_count(d_heads, n_keys, rec_bytes, d_table, n_prefix, d_counts=None, accumulate=False):
    """K1 on device tensors.  Returns the uint64 counts as an int64-typed CUDA
    tensor of length n_prefix (not synchronised)."""
    torch = require_cuda()
    if d_counts is None:
        d_counts = torch.empty(n_prefix, dtype=torch.int64, device=d_table.device)
        accumulate = False
    fn = lib().kas_match_count_accumulate if accumulate else lib().kas_match_count
    check(fn(ptr(d_heads), n_keys, rec_bytes, ptr(d_table), n_prefix, ptr(d_counts), stream_ptr()),
          'kas_match_count')
    return d_counts


def glob_count(d_blob, d_off, n_keys, patterns, sep_bytes=0):
    """General glob match and count on a device-resident blob+offsets keyspace.
    patterns: sequence of bytes globs.  Returns an int64-typed CUDA tensor."""
    torch = require_cuda()
    pats = [bytes(p) for p in patterns]
    poff = np.zeros(len(pats) + 1, dtype=np.uint32)
    poff[1:] = np.cumsum([len(p) for p in pats])
    pblob = np.frombuffer(b''.join(pats) + b'\0', dtype=np.uint8).copy()
    dev = d_off.device
    d_pat = torch.from_numpy(pblob).to(dev)
    d_poff = torch.from_numpy(poff.view(np.int32)).to(dev)
    counts = torch.empty(len(pats), dtype=torch.int64, device=dev)
    check(lib().kas_glob_count(ptr(d_blob), ptr(d_off), n_keys, sep_bytes, ptr(d_pat), ptr(d_poff),
                               len(pats), int(poff[-1]), ptr(counts), stream_ptr()), 'kas_glob_count')
    return counts


def counts_to_host(d_counts):
    """Synchronising read of K1's result; detects the table-mismatch poison."""
    out = d_counts.cpu().numpy().view(np.uint64)
    if out.size and (out == POISON).all():
        raise KasError(KAS_ERR_INVALID, 'kas_match_count',
                       'prefix table image does not match n_prefix/rec_bytes')
    return out


def decide(a, kpp, mn, mx, cur, mode=0):
    """Element-wise get_desired_pods (mode 0) / clip_pod_count (mode 1) on the GPU.
    Arguments are sequences of Python ints of equal length; returns list of int."""
    torch = require_cuda()
    n = len(a)
    cols = [a, kpp if mode == 0 else [1] * n, mn, mx, cur]
    try:
        host = np.array(cols, dtype=np.int64)
    except OverflowError as e:
        raise OverflowError('pod counts and tallies must fit in a signed 64-bit integer') from e
    dev = torch.from_numpy(host).to('cuda', non_blocking=False)
    out = torch.empty(n, dtype=torch.int64, device='cuda')
    status = torch.zeros(1, dtype=torch.int32, device='cuda')
    check(lib().kas_decide(ptr(dev[0]), ptr(dev[1]), ptr(dev[2]), ptr(dev[3]), ptr(dev[4]), n,
                           mode, ptr(out), ptr(status), stream_ptr()), 'kas_decide')
    res = out.cpu()
    if int(status.cpu()[0]) != 0:
        check(KAS_ERR_ZERO_DIVISION, 'kas_decide')
    return [int(v) for v in res.tolist()]


def trace_dtype_of(t):
    """TRACE_INT32 / TRACE_UINT16 for a trace tensor (int16 storage is read as uint16)."""
    torch = _torch()
    if t.dtype == torch.int32:
        return TRACE_INT32
    if t.dtype in (torch.uint16, torch.int16):
        return TRACE_UINT16
    raise TypeError('a trace must be int32 or uint16, got %s' % (t.dtype,))


def replay(d_llen, n_ticks, n_queues, policies, d_inprog=None, inprog_mode=0, d_group_off=None,
           n_resources=None, d_cur0=None, want_totals=True, want_sums=True, want_dense=False,
           d_patch_fail=None, startup_delay=0, d_resume=None, out_totals=None, out_sums=None):
    """K2 on device tensors.  policies: sequence of (kpp, min, max).
    ``startup_delay`` > 0: the controller reads the AVAILABLE replicas of a deployment
    (the reference's ``only_running`` branch) under the start-up model of include/kas.h.
    ``d_resume``: the totals tensor of an earlier replay of the same fleet and policies; every
    (resource, policy) starts from its final pods (chunked replay, see ``totals_chain``).
    ``out_totals`` / ``out_sums``: tensors to write into instead of fresh ones.
    The trace may be int32 or uint16 (a per-tick in-progress trace has the same type).
    Returns dict of CUDA tensors: totals (uint8 view [R][S][16]), sums int64 [S][3],
    dense int32 [S][T][R] (those requested)."""
    torch = require_cuda()
    pol = [(int(k), int(a), int(b)) for (k, a, b) in policies]
    for tup in pol:
        for v in tup:
            if not -(2 ** 31) <= v < 2 ** 31:
                raise OverflowError('policy values must fit in int32, got %r' % (tup,))
    S = len(pol)
    R = n_queues if n_resources is None else n_resources
    parr = (Policy * S)(*[Policy(*p) for p in pol])
    dev = d_llen.device
    out = {}
    totals = (out_totals if out_totals is not None
              else torch.empty((R, S, 16), dtype=torch.uint8, device=dev)) if want_totals else None
    sums = (out_sums if out_sums is not None
            else torch.empty((S, 3), dtype=torch.int64, device=dev)) if want_sums else None
    if d_resume is not None and tuple(d_resume.shape) != (R, S, 16):
        raise ValueError('d_resume must be the totals [R][S] of an earlier replay of the same fleet and policies')
    dense = torch.empty((S, n_ticks, R), dtype=torch.int32, device=dev) if want_dense else None
    dt = trace_dtype_of(d_llen)
    if inprog_mode == 2 and trace_dtype_of(d_inprog) != dt:
        raise TypeError('a per-tick in-progress trace must have the element type of the LLEN trace')
    args = ReplayArgs(ptr(d_llen), ptr(d_inprog), inprog_mode, n_ticks, n_queues, R,
                      ptr(d_group_off), ptr(d_cur0), parr, S, ptr(totals), ptr(sums), ptr(dense), dt,
                      ptr(d_patch_fail), _check_delay(startup_delay), ptr(d_resume))
    check(lib().kas_replay(C.byref(args), stream_ptr()), 'kas_replay')
    if want_totals:
        out['totals'] = totals
    if want_sums:
        out['sums'] = sums
    if want_dense:
        out['dense'] = dense
    return out


def patch_events(d_llen, n_ticks, n_queues, policy, d_inprog=None, inprog_mode=0, d_group_off=None,
                 n_resources=None, d_cur0=None, d_patch_fail=None, startup_delay=0):
    """Patch-event stream of one policy.  Returns (event_off numpy uint64 [R+1],
    events numpy structured [total] with tick/from_pods/to_pods)."""
    torch = require_cuda()
    k, a, b = (int(v) for v in policy)
    for v in (k, a, b):
        if not -(2 ** 31) <= v < 2 ** 31:
            raise OverflowError('policy values must fit in int32, got %r' % (policy,))
    R = n_queues if n_resources is None else n_resources
    parr = (Policy * 1)(Policy(k, a, b))
    dev = d_llen.device
    dt = trace_dtype_of(d_llen)
    if inprog_mode == 2 and trace_dtype_of(d_inprog) != dt:
        raise TypeError('a per-tick in-progress trace must have the element type of the LLEN trace')
    args = ReplayArgs(ptr(d_llen), ptr(d_inprog), inprog_mode, n_ticks, n_queues, R,
                      ptr(d_group_off), ptr(d_cur0), parr, 1, None, None, None, dt, ptr(d_patch_fail),
                      _check_delay(startup_delay), None)
    off = torch.empty(R + 1, dtype=torch.int64, device=dev)
    nscratch = lib().kas_patch_events_scratch_bytes(R)
    scratch = torch.empty(max(nscratch, 8), dtype=torch.uint8, device=dev)
    check(lib().kas_patch_event_offsets(C.byref(args), 0, ptr(off), ptr(scratch), nscratch,
                                        stream_ptr()), 'kas_patch_event_offsets')
    total = int(off[R].item())
    events = torch.empty((max(total, 1), 12), dtype=torch.uint8, device=dev)
    check(lib().kas_patch_event_emit(C.byref(args), 0, ptr(off), ptr(events), total, stream_ptr()),
          'kas_patch_event_emit')
    ev = events.cpu().numpy().reshape(-1)[:total * 12].view(EVENT_DTYPE)
    return off.cpu().numpy().view(np.uint64), ev


def first_pod_latency(d_llen, n_ticks, n_queues, policy, d_inprog=None, inprog_mode=0, d_group_off=None,
                      n_resources=None, d_cur0=None, d_patch_fail=None, startup_delay=0, want_rows=True):
    """Time to the first pod after a burst (kas_first_pod_latency) for one policy.  Returns
    (rows, sums): rows a numpy structured array [R] (FIRST_POD_DTYPE) or None, sums int64 [5] =
    served, unserved, wait_sum, max wait, pending."""
    torch = require_cuda()
    k, a, b = (int(v) for v in policy)
    for v in (k, a, b):
        if not -(2 ** 31) <= v < 2 ** 31:
            raise OverflowError('policy values must fit in int32, got %r' % (policy,))
    R = n_queues if n_resources is None else n_resources
    parr = (Policy * 1)(Policy(k, a, b))
    dev = d_llen.device
    dt = trace_dtype_of(d_llen)
    if inprog_mode == 2 and trace_dtype_of(d_inprog) != dt:
        raise TypeError('a per-tick in-progress trace must have the element type of the LLEN trace')
    args = ReplayArgs(ptr(d_llen), ptr(d_inprog), inprog_mode, n_ticks, n_queues, R,
                      ptr(d_group_off), ptr(d_cur0), parr, 1, None, None, None, dt, ptr(d_patch_fail),
                      _check_delay(startup_delay), None)
    rows = torch.empty((R, FIRST_POD_DTYPE.itemsize), dtype=torch.uint8, device=dev) if want_rows else None
    sums = torch.empty(5, dtype=torch.int64, device=dev)
    check(lib().kas_first_pod_latency(C.byref(args), 0, ptr(rows), ptr(sums), stream_ptr()),
          'kas_first_pod_latency')
    host_rows = rows.cpu().numpy().reshape(-1).view(FIRST_POD_DTYPE) if want_rows else None
    return host_rows, sums.cpu().numpy()


def totals_chain(d_acc, d_chunk, d_sums_acc=None, d_sums_chunk=None):
    """acc += chunk for the totals (and sums) of consecutive tick chunks (kas_totals_chain)."""
    R, S = int(d_acc.shape[0]), int(d_acc.shape[1])
    check(lib().kas_totals_chain(ptr(d_acc), ptr(d_chunk), R, S, ptr(d_sums_acc), ptr(d_sums_chunk),
                                 stream_ptr()), 'kas_totals_chain')


def totals_to_host(d_totals):
    """uint8 [R][S][16] CUDA tensor -> numpy structured array [R][S]."""
    a = d_totals.cpu().numpy()
    return a.reshape(a.shape[0], a.shape[1] * 16).view(TOTALS_DTYPE).reshape(a.shape[0], a.shape[1])


def synth_trace(seed, n_ticks, q_first, n_q, device='cuda', out=None):
    torch = require_cuda()
    if out is None:
        out = torch.empty((n_ticks, n_q), dtype=torch.int32, device=device)
    check(lib().kas_synth_trace(seed, n_ticks, q_first, n_q, ptr(out), stream_ptr()),
          'kas_synth_trace')
    return out


def pack_heads_device(d_blob, d_off, n_keys, rec_bytes, sep_bytes=0):
    torch = require_cuda()
    heads = torch.empty((n_keys, rec_bytes), dtype=torch.uint8, device=d_blob.device)
    check(lib().kas_pack_heads(ptr(d_blob), ptr(d_off), n_keys, sep_bytes, rec_bytes, ptr(heads),
                               stream_ptr()), 'kas_pack_heads')
    return heads


def set_reserved_sms(n):
    """Leave ``n`` SMs out of the match kernel's persistent grid (room for a concurrent collective)."""
    check(lib().kas_set_reserved_sms(int(n)), 'kas_set_reserved_sms')


def device_info():
    require_cuda()
    a, b, c = C.c_int(0), C.c_int(0), C.c_int(0)
    check(lib().kas_device_info(C.byref(a), C.byref(b), C.byref(c)), 'kas_device_info')
    return dict(sm_count=a.value, cc=(b.value, c.value))


# --------------------------------------------------------------------------
# box-wide totals (the single collective of the path)
# --------------------------------------------------------------------------
class Comm(object):
    """One rank's handle on the library's own communicator (kas_comm_* in include/kas.h):
    an NCCL all-reduce of int64 totals issued by libkas on the caller's stream.  Torch is
    not involved; any side channel may carry the 128-byte id from rank 0 to the others
    (``Comm.from_torch_group`` uses an initialised torch.distributed group for that)."""

    def __init__(self, world_size, rank, unique_id):
        require_cuda()
        uid = bytes(unique_id)
        if len(uid) != COMM_ID_BYTES:
            raise ValueError('the communicator id has %d bytes' % COMM_ID_BYTES)
        self.world_size, self.rank = int(world_size), int(rank)
        self._buf = C.create_string_buffer(uid, COMM_ID_BYTES)
        self._comm = C.c_void_p(0)
        check(lib().kas_comm_init(C.byref(self._comm), self.world_size, self.rank, self._buf),
              'kas_comm_init')

    @staticmethod
    def unique_id():
        buf = C.create_string_buffer(COMM_ID_BYTES)
        check(lib().kas_comm_unique_id(buf), 'kas_comm_unique_id')
        return buf.raw

    @classmethod
    def from_torch_group(cls, group=None):
        """Every rank of an initialised torch.distributed group calls this (its GPU current)."""
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        box = [cls.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0,
                                   group=group)
        return cls(world, rank, box[0])

    def allreduce_sum_(self, t):
        """In-place sum over the ranks of an int64 CUDA tensor, on the current stream."""
        torch = _torch()
        if t.dtype != torch.int64 or not t.is_cuda or not t.is_contiguous():
            raise TypeError('kas_allreduce_i64 takes a contiguous int64 CUDA tensor')
        check(lib().kas_allreduce_i64(self._comm, ptr(t), t.numel(), stream_ptr()), 'kas_allreduce_i64')
        return t

    def destroy(self):
        if self._comm:
            check(lib().kas_comm_destroy(self._comm), 'kas_comm_destroy')
            self._comm = C.c_void_p(0)
