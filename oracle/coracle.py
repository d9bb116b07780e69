"""ctypes front for oracle/libkas_oracle.so -- TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build as _build

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = _build.OUT
        if not os.path.exists(path) or (
                os.path.exists(_build.SRC)
                and os.path.getmtime(path) < os.path.getmtime(_build.SRC)):
            _build.build()
        L = C.CDLL(path)
        u8p, u32p, u64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)
        i32p, i64p = C.POINTER(C.c_int32), C.POINTER(C.c_int64)
        L.kas_oracle_glob_match.argtypes = [C.c_char_p, C.c_uint32, C.c_char_p, C.c_uint32]
        L.kas_oracle_glob_match.restype = C.c_int
        L.kas_oracle_match_counts.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p,
                                              C.c_void_p, C.c_uint32, C.c_void_p, C.c_int]
        L.kas_oracle_match_counts.restype = None
        L.kas_oracle_match_counts_heads.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p,
                                                    C.c_void_p, C.c_uint32, C.c_void_p, C.c_int]
        L.kas_oracle_match_counts_heads.restype = None
        for f in (L.kas_oracle_clip_pod_count,):
            f.argtypes = [C.c_int64] * 4
            f.restype = C.c_int64
        L.kas_oracle_get_desired_pods.argtypes = [C.c_int64] * 5
        L.kas_oracle_get_desired_pods.restype = C.c_int64
        L.kas_oracle_replay.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint32, C.c_uint64,
                                        C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32,
                                        C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.kas_oracle_replay.restype = None
        L.kas_oracle_replay_delay.argtypes = L.kas_oracle_replay.argtypes[:-1] + [C.c_uint32, C.c_int]
        L.kas_oracle_replay_delay.restype = None
        L.kas_oracle_first_pod.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint32, C.c_uint64, C.c_void_p,
                                           C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                           C.c_uint32, C.c_int]
        L.kas_oracle_first_pod.restype = None
        L.kas_oracle_trace_value.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32]
        L.kas_oracle_trace_value.restype = C.c_int32
        L.kas_oracle_trace_fill.argtypes = [C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint64,
                                            C.c_void_p, C.c_int]
        L.kas_oracle_trace_fill.restype = None
        L.kas_oracle_max_threads.restype = C.c_int
        L.kas_oracle_synth_heads.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p,
                                             C.c_uint32, C.c_int, C.c_uint32, C.c_void_p, C.c_int]
        L.kas_oracle_synth_heads.restype = None
        L.kas_oracle_synth_blob.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p,
                                            C.c_uint32, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.kas_oracle_synth_blob.restype = C.c_uint64
        _lib = L
    return _lib


def max_threads():
    """Every core this process may run on.  OMP_NUM_THREADS is deliberately not
    consulted (torchrun sets it to 1): the CPU baseline is meant to use the host."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def _names_blob(queues):
    enc = [q if isinstance(q, bytes) else q.encode('utf-8') for q in queues]
    off = np.zeros(len(enc) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(e) for e in enc])
    blob = np.frombuffer(b''.join(enc) + b'\0', dtype=np.uint8).copy()
    return blob, off


def keys_to_blob(keys):
    off = np.zeros(len(keys) + 1, dtype=np.uint64)
    off[1:] = np.cumsum([len(k) for k in keys], dtype=np.uint64)
    blob = np.frombuffer(b''.join(keys) + b'\0', dtype=np.uint8).copy()
    return blob, off


def glob_match(pat: bytes, s: bytes) -> bool:
    return bool(lib().kas_oracle_glob_match(pat, len(pat), s, len(s)))


def match_counts(queues, blob, off, threads=None):
    """SCAN term per queue over a blob+offsets keyspace (Q passes, like the reference)."""
    names, noff = _names_blob(queues)
    counts = np.zeros(len(queues), dtype=np.uint64)
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    off = np.ascontiguousarray(off, dtype=np.uint64)
    lib().kas_oracle_match_counts(blob.ctypes.data, off.ctypes.data, len(off) - 1,
                                  names.ctypes.data, noff.ctypes.data, len(queues),
                                  counts.ctypes.data, threads or max_threads())
    return counts


def match_counts_heads(queues, heads, threads=None):
    names, noff = _names_blob(queues)
    counts = np.zeros(len(queues), dtype=np.uint64)
    heads = np.ascontiguousarray(heads, dtype=np.uint8)
    n, w = heads.shape
    lib().kas_oracle_match_counts_heads(heads.ctypes.data, n, w, names.ctypes.data,
                                        noff.ctypes.data, len(queues), counts.ctypes.data,
                                        threads or max_threads())
    return counts


MAX_STARTUP_DELAY = 64


def replay(llen, inprog, group_off, cur0, policies, want_dense=False, threads=None, patch_fail=None,
           startup_delay=0):
    if not 0 <= int(startup_delay) <= MAX_STARTUP_DELAY:
        raise ValueError('startup_delay out of range')
    llen = np.ascontiguousarray(llen, dtype=np.int32)
    T, Q = llen.shape
    mode = 0
    ip = None
    if inprog is not None:
        ip = np.ascontiguousarray(inprog, dtype=np.int32)
        mode = 1 if ip.ndim == 1 else 2
    go = None
    R = Q
    if group_off is not None:
        go = np.ascontiguousarray(group_off, dtype=np.uint64)
        R = len(go) - 1
    c0 = None if cur0 is None else np.ascontiguousarray(cur0, dtype=np.int32)
    pol = np.ascontiguousarray(policies, dtype=np.int32).reshape(-1, 3)
    if (pol[:, 0] == 0).any():
        raise ZeroDivisionError('integer division or modulo by zero')
    S = pol.shape[0]
    pod_ticks = np.zeros((R, S), dtype=np.int64)
    n_patches = np.zeros((R, S), dtype=np.int64)
    final = np.zeros((R, S), dtype=np.int32)
    dense = np.zeros((S, T, R), dtype=np.int32) if want_dense else None
    pf = None if patch_fail is None else np.ascontiguousarray(patch_fail, dtype=np.uint8).reshape(T, R)
    lib().kas_oracle_replay_delay(llen.ctypes.data, None if ip is None else ip.ctypes.data, mode, T, Q,
                                  None if go is None else go.ctypes.data, R,
                                  None if c0 is None else c0.ctypes.data, pol.ctypes.data, S,
                                  pod_ticks.ctypes.data, n_patches.ctypes.data, final.ctypes.data,
                                  None if dense is None else dense.ctypes.data,
                                  None if pf is None else pf.ctypes.data,
                                  int(startup_delay), threads or max_threads())
    out = dict(pod_ticks=pod_ticks, n_patches=n_patches, final=final)
    if want_dense:
        out['dense'] = dense
    return out


FIRST_POD_FIELDS = ('n_served', 'n_unserved', 'wait_sum', 'wait_max', 'pending')


def first_pod(llen, inprog, group_off, cur0, policy, threads=None, patch_fail=None, startup_delay=0):
    """Time to the first pod after a burst per resource (oracle.first_pod_latency over a fleet):
    dict of int64 arrays [R] keyed by FIRST_POD_FIELDS."""
    llen = np.ascontiguousarray(llen, dtype=np.int32)
    T, Q = llen.shape
    mode, ip = 0, None
    if inprog is not None:
        ip = np.ascontiguousarray(inprog, dtype=np.int32)
        mode = 1 if ip.ndim == 1 else 2
    go, R = None, Q
    if group_off is not None:
        go = np.ascontiguousarray(group_off, dtype=np.uint64)
        R = len(go) - 1
    c0 = None if cur0 is None else np.ascontiguousarray(cur0, dtype=np.int32)
    pol = np.ascontiguousarray(policy, dtype=np.int32).reshape(3)
    if pol[0] == 0:
        raise ZeroDivisionError('integer division or modulo by zero')
    out = np.zeros((R, 5), dtype=np.int64)
    pf = None if patch_fail is None else np.ascontiguousarray(patch_fail, dtype=np.uint8).reshape(T, R)
    lib().kas_oracle_first_pod(llen.ctypes.data, None if ip is None else ip.ctypes.data, mode, T, Q,
                               None if go is None else go.ctypes.data, R,
                               None if c0 is None else c0.ctypes.data, pol.ctypes.data, out.ctypes.data,
                               None if pf is None else pf.ctypes.data, int(startup_delay),
                               threads or max_threads())
    return {k: out[:, i].copy() for i, k in enumerate(FIRST_POD_FIELDS)}


def trace_fill(seed, T, q_first, n_q, threads=None):
    out = np.empty((T, n_q), dtype=np.int32)
    lib().kas_oracle_trace_fill(seed, T, q_first, n_q, out.ctypes.data, threads or max_threads())
    return out


def rec_bytes_for(queues):
    """Record width the path uses for these names: "processing-" + name + ":" rounded up to 16."""
    longest = max((len(q if isinstance(q, bytes) else q.encode('utf-8')) for q in queues), default=0)
    return 16 * ((11 + longest + 1 + 15) // 16)


def synth_heads(seed, first_key, n_keys, queues, rec_bytes, long_tail=False, threads=None):
    """The synthetic keyspace of DESIGN.md section 5 as packed heads, made by the oracle's own
    generator (no product code involved)."""
    names, noff = _names_blob(queues)
    out = np.empty((n_keys, rec_bytes), dtype=np.uint8)
    lib().kas_oracle_synth_heads(seed, first_key, n_keys, names.ctypes.data, noff.ctypes.data,
                                 len(queues), int(long_tail), rec_bytes, out.ctypes.data,
                                 threads or max_threads())
    return out


def synth_blob(seed, first_key, n_keys, queues, long_tail=False, threads=None):
    """The same keys as (blob uint8, offsets uint64 [n+1])."""
    names, noff = _names_blob(queues)
    off = np.empty(n_keys + 1, dtype=np.uint64)
    t = threads or max_threads()
    need = lib().kas_oracle_synth_blob(seed, first_key, n_keys, names.ctypes.data, noff.ctypes.data,
                                       len(queues), int(long_tail), None, off.ctypes.data, t)
    blob = np.empty(max(1, need), dtype=np.uint8)
    lib().kas_oracle_synth_blob(seed, first_key, n_keys, names.ctypes.data, noff.ctypes.data,
                                len(queues), int(long_tail), blob.ctypes.data, off.ctypes.data, t)
    return blob[:need], off
