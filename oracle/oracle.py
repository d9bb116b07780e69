"""CPU oracle for the queue-tally / replica-decision path -- TEST INFRASTRUCTURE ONLY.

This module is a plain-Python (and small-numpy) restatement of the one path of
vanvalenlab/kiosk-autoscaler that this repository accelerates.  It exists so
that the CUDA path can be checked bit for bit.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl
reference`` legs may import it; nothing under ``kiosk-autoscaler_b200/`` does.

Every function cites the reference lines it restates (paths relative to the
reference checkout, ``autoscaler/autoscaler.py`` unless stated).

Parity status
-------------
* Decision arithmetic (``clip_pod_count``, ``get_desired_pods``, the sum and
  second clip of ``scale``): PINNED.  Checked against the reference's own
  known-answer test (``autoscaler/autoscaler_test.py:87-102``) and against
  outputs produced by executing the reference's source in this container
  (``oracle/make_golden.py`` -> ``tests/golden/*.json``).
* ``LLEN`` term of the tally: PINNED the same way.
* ``SCAN MATCH`` term of the tally: PARITY UNPINNED upstream.  The glob match
  runs inside redis-server (``stringmatchlen`` in redis' ``util.c``; server
  version is not pinned anywhere in the reference, client is ``redis~=3.5.3``,
  test double ``fakeredis`` unpinned) and no reference test ever creates a
  ``processing-*`` key (``autoscaler_test.py:184-210``).  ``redis_glob_match``
  below restates the published glob rules (``*``, ``?``, ``[...]``, ``\\``);
  the golden tallies were produced by the reference's ``tally_queues`` driven
  by an in-memory stub whose ``scan_iter`` uses that same matcher.
"""
from __future__ import annotations

from typing import Callable, Dict, Iterable, List, Optional, Sequence, Tuple

PROCESSING_FMT = 'processing-{}:*'   # autoscaler.py:69


# --------------------------------------------------------------------------
# Redis glob (third-party: redis-server util.c stringmatchlen, unpinned)
# --------------------------------------------------------------------------
def redis_glob_match(pattern: bytes, string: bytes) -> bool:
    """Case-sensitive Redis glob: ``*`` any run (incl. empty), ``?`` one byte,
    ``[abc]`` / ``[a-z]`` / ``[^x]`` classes, ``\\x`` literal x.

    Restates the documented behaviour of ``SCAN ... MATCH`` (call site
    autoscaler.py:70).  Quirks kept from the server's matcher: inside a class
    a backslash escapes the next byte; a range may be written backwards
    (``[z-a]``); an unterminated class ends at the end of the pattern; a
    trailing lone backslash matches a literal backslash.
    """
    return _glob(pattern, 0, string, 0)


def _glob(p: bytes, pi: int, s: bytes, si: int) -> bool:
    plen, slen = len(p), len(s)
    while pi < plen:
        c = p[pi]
        if c == 0x2A:  # '*'
            while pi + 1 < plen and p[pi + 1] == 0x2A:
                pi += 1
            if pi + 1 == plen:
                return True
            while si <= slen:
                if _glob(p, pi + 1, s, si):
                    return True
                si += 1
            return False
        if c == 0x3F:  # '?'
            if si >= slen:
                return False
            si += 1
            pi += 1
            continue
        if c == 0x5B:  # '['
            if si >= slen:
                return False
            pi += 1
            negate = pi < plen and p[pi] == 0x5E
            if negate:
                pi += 1
            matched = False
            while True:
                if pi < plen and p[pi] == 0x5C and pi + 1 < plen:
                    pi += 1
                    if p[pi] == s[si]:
                        matched = True
                elif pi < plen and p[pi] == 0x5D:
                    break
                elif pi >= plen:
                    pi -= 1
                    break
                elif pi + 2 < plen and p[pi + 1] == 0x2D:
                    lo, hi = p[pi], p[pi + 2]
                    if lo > hi:
                        lo, hi = hi, lo
                    pi += 2
                    if lo <= s[si] <= hi:
                        matched = True
                else:
                    if p[pi] == s[si]:
                        matched = True
                pi += 1
            if negate:
                matched = not matched
            if not matched:
                return False
            si += 1
            pi += 1
            continue
        if c == 0x5C and pi + 1 < plen:  # '\\'
            pi += 1
            c = p[pi]
        if si >= slen or s[si] != c:
            return False
        si += 1
        pi += 1
    return si == slen


GLOB_META = b'*?[\\'


def has_glob_meta(name: bytes) -> bool:
    return any(ch in GLOB_META for ch in name)


# --------------------------------------------------------------------------
# Tally  (autoscaler.py:47-58, 60-77)
# --------------------------------------------------------------------------
def parse_queues(queues: str = 'predict', queue_delim: str = ',') -> Dict[str, int]:
    """autoscaler.py:52 -- ``{q: 0 for q in queues.split(queue_delim)}``."""
    return {q: 0 for q in queues.split(queue_delim)}


def _as_bytes(x) -> bytes:
    return x if isinstance(x, (bytes, bytearray)) else str(x).encode('utf-8')


def tally_queues(redis_keys: Dict[str, int],
                 llen: Callable[[str], int],
                 keyspace: Sequence[bytes]) -> Dict[str, int]:
    """autoscaler.py:60-77.  For every queue independently: ``LLEN q`` plus the
    number of keys of the whole keyspace matching ``processing-<q>:*``
    (one full scan per queue, no dedup, autoscaler.py:69-71)."""
    for q in redis_keys:
        num_items = llen(q)                                   # :67
        pattern = _as_bytes(PROCESSING_FMT.format(q))         # :69
        num_in_progress = 0
        for k in keyspace:                                    # :70-71
            if redis_glob_match(pattern, k):
                num_in_progress += 1
        redis_keys[q] = num_items + num_in_progress           # :73
    return redis_keys


def match_counts(queues: Sequence[str], keyspace: Sequence[bytes]) -> List[int]:
    """The SCAN term alone, one count per queue (autoscaler.py:69-71)."""
    out = []
    for q in queues:
        pattern = _as_bytes(PROCESSING_FMT.format(q))
        out.append(sum(1 for k in keyspace if redis_glob_match(pattern, k)))
    return out


# --------------------------------------------------------------------------
# Decision  (autoscaler.py:197-219)
# --------------------------------------------------------------------------
def clip_pod_count(desired_pods: int, min_pods: int, max_pods: int,
                   current_pods: int) -> int:
    """autoscaler.py:197-213.  if/elif order kept: with min > max the upper
    bound wins for large values and the lower bound for small ones."""
    if desired_pods > max_pods:          # :200
        desired_pods = max_pods
    elif desired_pods < min_pods:        # :202
        desired_pods = min_pods
    if 0 < desired_pods < current_pods:  # :207
        desired_pods = current_pods
    return desired_pods


def get_desired_pods(tally: int, keys_per_pod: int, min_pods: int,
                     max_pods: int, current_pods: int) -> int:
    """autoscaler.py:215-219.  Python floor division; ``keys_per_pod == 0``
    raises ZeroDivisionError exactly as the reference does."""
    desired_pods = tally // keys_per_pod
    return clip_pod_count(desired_pods, min_pods, max_pods, current_pods)


def scale_step(tallies: Iterable[int], keys_per_pod: int, min_pods: int,
               max_pods: int, current_pods: int) -> int:
    """autoscaler.py:254-260: sum of per-queue decisions, clipped again."""
    desired = 0
    for t in tallies:
        desired += get_desired_pods(t, keys_per_pod, min_pods, max_pods,
                                    current_pods)
    return clip_pod_count(desired, min_pods, max_pods, current_pods)


def replay_resource(llen_rows: Sequence[Sequence[int]],
                    inprog_rows: Optional[Sequence[Sequence[int]]],
                    keys_per_pod: int, min_pods: int, max_pods: int,
                    cur0: int = 0,
                    patch_fails: Optional[Sequence[bool]] = None,
                    startup_delay: int = 0) -> Tuple[List[int], List[int]]:
    """Replay one resource over T ticks with a well-behaved Kubernetes mock:
    what a tick patches (autoscaler.py:232,236) is what the next tick reads
    (autoscaler.py:179,189), and no patch is issued when desired == current
    (autoscaler.py:227-228).

    ``llen_rows[t][j]`` is LLEN of the resource's j-th queue at tick t;
    ``inprog_rows[t][j]`` the SCAN count (None -> 0).  ``patch_fails[t]`` true
    means the patch call of tick t raises ApiException: ``scale`` swallows it
    (autoscaler.py:267-273) and the resource keeps its replicas; the attempt is
    still listed.  Returns (replicas after each tick, list of values the
    patch calls carried, in order).

    ``startup_delay`` D > 0 models the ``only_running`` reading of a deployment
    (autoscaler.py:176-177, ``status.available_replicas``): a replica counts as
    available once it has been part of ``spec.replicas`` for D consecutive ticks
    and is gone as soon as the spec drops it, i.e. ``available[t+1] =
    min(spec[t-D..t])`` with ``spec[<0] = cur0``.  The controller reads
    ``available`` (autoscaler.py:252 with the flag set), and patches whenever
    desired != available (autoscaler.py:227), also when the spec already holds
    that value.  The returned list is ``spec.replicas`` after each tick.
    """
    spec = cur0
    avail = cur0
    hist = [cur0] * startup_delay          # spec of the last D ticks
    after, patches = [], []
    for t, row in enumerate(llen_rows):
        if inprog_rows is None:
            tallies = list(row)
        else:
            tallies = [a + b for a, b in zip(row, inprog_rows[t])]
        cur = avail
        desired = scale_step(tallies, keys_per_pod, min_pods, max_pods, cur)
        if desired != cur:
            patches.append(desired)
            if not (patch_fails is not None and patch_fails[t]):
                spec = desired
        avail = min([spec] + hist)
        if startup_delay:
            hist[t % startup_delay] = spec
        after.append(spec)
    return after, patches


def first_pod_latency(llen_rows: Sequence[Sequence[int]],
                      inprog_rows: Optional[Sequence[Sequence[int]]],
                      keys_per_pod: int, min_pods: int, max_pods: int,
                      cur0: int = 0,
                      patch_fails: Optional[Sequence[bool]] = None,
                      startup_delay: int = 0) -> dict:
    """Time to the first pod after a burst, for one resource under one policy (a report on
    top of the replayed ``scale`` loop, autoscaler.py:244-273; the reference itself never
    computes it -- README.md:7 states the purpose it serves, "minimizing GPU costs").

    A burst starts on a tick whose tally (sum over the resource's queues of LLEN + SCAN count,
    autoscaler.py:73) is positive after a tick where it was not (before the first tick it counts
    as 0), while the controller reads 0 pods on that tick (autoscaler.py:252; the AVAILABLE
    pods under ``startup_delay``, autoscaler.py:176-177) and no earlier burst is still waiting.
    It is SERVED on the first tick at or after its start after which pods are running (what the
    next tick will read is positive); its wait is that tick minus the starting tick (0 when the
    same tick's patch brings pods up).  If the tally returns to 0 first, the burst went UNSERVED.
    Returns n_served, n_unserved, wait_sum, wait_max and pending (ticks the last burst has
    waited when the trace ends, 0 if none is waiting)."""
    spec = cur0
    avail = cur0
    hist = [cur0] * startup_delay
    waiting = -1
    prev_tally = 0
    out = dict(n_served=0, n_unserved=0, wait_sum=0, wait_max=0, pending=0)
    T = 0
    for t, row in enumerate(llen_rows):
        T = t + 1
        if inprog_rows is None:
            tallies = list(row)
        else:
            tallies = [a + b for a, b in zip(row, inprog_rows[t])]
        tally = sum(tallies)
        cur = avail
        desired = scale_step(tallies, keys_per_pod, min_pods, max_pods, cur)
        if desired != cur and not (patch_fails is not None and patch_fails[t]):
            spec = desired
        avail = min([spec] + hist)
        if startup_delay:
            hist[t % startup_delay] = spec
        if waiting < 0 and tally > 0 and prev_tally <= 0 and cur <= 0:
            waiting = t
        if waiting >= 0:
            if avail > 0:
                w = t - waiting
                out['n_served'] += 1
                out['wait_sum'] += w
                out['wait_max'] = max(out['wait_max'], w)
                waiting = -1
            elif tally <= 0:
                out['n_unserved'] += 1
                waiting = -1
        prev_tally = tally
    if waiting >= 0:
        out['pending'] = T - waiting
    return out


def replay_totals(after: Sequence[int], cur0: int) -> Tuple[int, int, int]:
    """Reductions the replay kernel reports per (resource, policy):
    pod_ticks = sum of replicas after each tick, n_patches = number of ticks on
    which the value changed, final_current = replicas after the last tick."""
    pod_ticks = 0
    n_patches = 0
    cur = cur0
    for a in after:
        pod_ticks += a
        n_patches += int(a != cur)
        cur = a
    return pod_ticks, n_patches, cur


# --------------------------------------------------------------------------
# numpy forms (same arithmetic, vectorised over resources; loop over ticks)
# --------------------------------------------------------------------------
def np_clip(d, mn, mx, cur):
    """Vector form of clip_pod_count; ``d`` and ``cur`` int64 arrays."""
    import numpy as np
    d = np.where(d > mx, mx, np.where(d < mn, mn, d))
    return np.where((d > 0) & (d < cur), cur, d)


def np_replay(llen, inprog, group_off, cur0, policies, want_dense=False):
    """Vectorised replay.

    llen      int [T][Q] tick-major
    inprog    None | int [Q] | int [T][Q]
    group_off None (every queue its own resource) | int [R+1] CSR offsets
    cur0      None | int [R]
    policies  sequence of (kpp, min, max)

    Returns dict(pod_ticks[R][S] int64, n_patches[R][S] int64,
                 final[R][S] int64, dense[S][T][R] int64 if want_dense).
    """
    import numpy as np
    llen = np.asarray(llen, dtype=np.int64)
    T, Q = llen.shape
    if group_off is None:
        group_off = np.arange(Q + 1, dtype=np.int64)
    group_off = np.asarray(group_off, dtype=np.int64)
    R = len(group_off) - 1
    S = len(policies)
    qres = np.repeat(np.arange(R), np.diff(group_off))
    if cur0 is None:
        cur0 = np.zeros(R, dtype=np.int64)
    cur0 = np.asarray(cur0, dtype=np.int64)
    if inprog is not None:
        inprog = np.asarray(inprog, dtype=np.int64)
    pod_ticks = np.zeros((R, S), dtype=np.int64)
    n_patches = np.zeros((R, S), dtype=np.int64)
    final = np.zeros((R, S), dtype=np.int64)
    dense = np.zeros((S, T, R), dtype=np.int64) if want_dense else None
    for s, (kpp, mn, mx) in enumerate(policies):
        if kpp == 0:
            raise ZeroDivisionError('integer division or modulo by zero')
        cur = cur0.copy()
        for t in range(T):
            tally = llen[t]
            if inprog is not None:
                tally = tally + (inprog if inprog.ndim == 1 else inprog[t])
            d = np.floor_divide(tally, kpp)           # Python // semantics
            d = np_clip(d, mn, mx, cur[qres])
            tot = np.zeros(R, dtype=np.int64)
            np.add.at(tot, qres, d)
            tot = np_clip(tot, mn, mx, cur)
            n_patches[:, s] += (tot != cur)
            cur = tot
            pod_ticks[:, s] += cur
            if want_dense:
                dense[s, t] = cur
        final[:, s] = cur
    out = dict(pod_ticks=pod_ticks, n_patches=n_patches, final=final)
    if want_dense:
        out['dense'] = dense
    return out


# --------------------------------------------------------------------------
# Fixed-width head packing (what "ingest" does before the match kernel)
# --------------------------------------------------------------------------
PREFIX_LITERAL = b'processing-'


def rec_bytes_for(queues: Sequence[str]) -> int:
    """Smallest multiple of 16 that holds ``processing-<q>:`` for every q."""
    longest = max((len(_as_bytes(q)) for q in queues), default=0)
    need = len(PREFIX_LITERAL) + longest + 1
    return 16 * ((need + 15) // 16)


def pack_heads(keyspace: Sequence[bytes], rec_bytes: int):
    """First ``rec_bytes`` bytes of every key, zero padded; uint8 [N][W]."""
    import numpy as np
    out = np.zeros((len(keyspace), rec_bytes), dtype=np.uint8)
    for i, k in enumerate(keyspace):
        h = k[:rec_bytes]
        out[i, :len(h)] = np.frombuffer(h, dtype=np.uint8)
    return out
