"""Generate tests/golden/* by EXECUTING THE REFERENCE'S OWN SOURCE.

TEST INFRASTRUCTURE.  Runs only where /root/reference exists (the build
container).  The reference's autoscaler/autoscaler.py is loaded unmodified
(oracle/refload.py) and driven by an in-memory Redis stub and a stateful
Kubernetes double; its outputs are written as small fixtures that travel to
the GPU box, where /root/reference does not exist.

    python oracle/make_golden.py            # rewrites tests/golden/

Fixtures
  known_answers.json   the reference test's own vectors (autoscaler_test.py:87-102,
                       155-182) re-derived by running the reference
  decisions.json       get_desired_pods / clip_pod_count on edge + random tuples
  scale_cases.json     scale() trajectories (multi-queue, job/deployment, quirks)
  tally_cases.json     tally_queues() on adversarial keyspaces
  c1.npz               BASELINE config 1: QUEUES='predict,track', 10k keys,
                       10k-tick trace, policy sweep
"""
from __future__ import annotations

import json
import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import refload  # noqa: E402

GOLD = os.path.join(ROOT, 'tests', 'golden')


def _scaler(queues='queue', delim=',', redis=None, kube=None):
    return refload.make_reference_scaler(redis or refload.StubRedis(), queues, delim, kube)


def known_answers():
    s = _scaler('queue')
    s.redis_keys['queue'] = 10
    gdp = []
    for args in [(2, 0, 2, 1), (5, 9, 10, 0), (3, 0, 5, 1), (10, 0, 5, 3)]:
        gdp.append(dict(args=list(args), tally=10,
                        result=s.get_desired_pods('queue', *args)))
    assert [g['result'] for g in gdp] == [2, 9, 3, 3], gdp   # autoscaler_test.py:92-102
    out = dict(get_desired_pods=gdp)
    # kpp == 0 -> ZeroDivisionError (reference raises; record the type name)
    try:
        s.get_desired_pods('queue', 0, 0, 10, 0)
        out['kpp_zero'] = 'no error'
    except ZeroDivisionError as e:
        out['kpp_zero'] = type(e).__name__
    try:
        s.get_desired_pods('nope', 1, 0, 10, 0)
        out['unknown_key'] = 'no error'
    except KeyError as e:
        out['unknown_key'] = type(e).__name__
    # __init__ parsing (autoscaler.py:52)
    out['init'] = []
    for queues, delim in [('predict', ','), ('predict,track', ','), ('a,a,b', ','),
                          (' a, b', ','), ('', ','), ('a|b|c', '|'), ('a,,b', ',')]:
        out['init'].append(dict(queues=queues, delim=delim,
                                keys=list(_scaler(queues, delim).redis_keys.keys())))
    return out


def decisions():
    s = _scaler('queue')
    rng = random.Random(1234)
    edge = [-(2 ** 31), -(2 ** 31) + 1, -7, -3, -2, -1, 0, 1, 2, 3, 5, 7, 10, 64,
            1000, 2 ** 31 - 2, 2 ** 31 - 1]
    rows = []

    def add(tally, kpp, mn, mx, cur):
        s.redis_keys['queue'] = tally
        rows.append([tally, kpp, mn, mx, cur,
                     s.get_desired_pods('queue', kpp, mn, mx, cur)])

    for tally in [0, 1, 9, 10, 11, 2 ** 31 - 1, 2 ** 32 - 2, -1, -10, -(2 ** 31)]:
        for kpp in [1, 2, 3, 10, 2 ** 31 - 1, -1, -3, -(2 ** 31)]:
            for mn, mx in [(0, 1), (0, 5), (1, 4), (9, 5), (-2, 3), (0, 2 ** 31 - 1),
                           (-(2 ** 31), 2 ** 31 - 1), (3, 3)]:
                for cur in [0, 1, 4, 7, -1, 2 ** 31 - 1]:
                    add(tally, kpp, mn, mx, cur)
    for _ in range(3000):
        pick = lambda: rng.choice(edge) if rng.random() < 0.3 else rng.randint(-50, 50)  # noqa: E731
        kpp = 0
        while kpp == 0:
            kpp = pick()
        tally = rng.choice([rng.randint(0, 200), rng.randint(0, 2 ** 32 - 2), pick()])
        add(tally, kpp, pick(), pick(), pick())
    clip = []
    for _ in range(1500):
        a = [rng.choice(edge) if rng.random() < 0.3 else rng.randint(-20, 20) for _ in range(4)]
        clip.append(a + [s.clip_pod_count(*a)])
    clip.append([7, 9, 5, 0, s.clip_pod_count(7, 9, 5, 0)])
    clip.append([3, 9, 5, 0, s.clip_pod_count(3, 9, 5, 0)])
    return dict(get_desired_pods=rows, clip_pod_count=clip)


def _run_scale(queues, depths, keyspace, kpp, mn, mx, cur0, rtype, delim=',', fails=None):
    """depths: list over ticks of list over queues (LLEN values); fails[t] true makes the
    patch call of tick t raise ApiException (swallowed by scale(), autoscaler.py:267-273)."""
    names = queues.split(delim)
    redis = refload.StubRedis(keyspace=[k.encode('latin-1') for k in keyspace])
    kube = refload.StatefulKube('r', rtype, cur0)
    s = _scaler(queues, delim, redis, kube)
    after, tallies = [], []
    for t, row in enumerate(depths):
        redis.lists = dict(zip(names, row))
        kube.fail_next_patch = bool(fails and fails[t])
        s.scale('ns', rtype, 'r', min_pods=mn, max_pods=mx, keys_per_pod=kpp)
        after.append(int(kube.value))
        tallies.append([s.redis_keys[q] for q in s.redis_keys])
    return dict(queues=queues, delim=delim, depths=depths, keyspace=keyspace,
                kpp=kpp, min=mn, max=mx, cur0=cur0, type=rtype, fails=fails,
                after=after, patches=[int(p) for p in kube.patches], tallies=tallies)


def scale_cases():
    cases = []
    # SURVEY appendix B (re-derived here by running the reference)
    cases.append(_run_scale('a', [[0], [3], [7], [2], [1], [0], [4]], [], 2, 0, 3, 0, 'deployment'))
    cases.append(_run_scale('a,b', [[1, 1]], [], 1, 0, 5, 3, 'deployment'))
    cases.append(_run_scale('a,b', [[0, 0]], [], 1, 1, 5, 0, 'deployment'))
    cases.append(_run_scale('a', [[9], [0], [9]], [], 1, 0, 4, 7, 'deployment'))
    cases.append(_run_scale('a,b', [[2, 3], [0, 0]], [], 1, 0, 10, 0, 'job'))
    # static in-progress keys add to every tick's tally
    ks = ['processing-a:h1', 'processing-a:h2', 'processing-b:', 'processing-ab:h', 'a', 'b',
          'processing-a', 'xprocessing-a:h']
    cases.append(_run_scale('a,b', [[0, 0], [1, 0], [0, 0], [5, 5]], ks, 2, 0, 4, 0, 'deployment'))
    # duplicate queue names collapse (autoscaler.py:52)
    cases.append(_run_scale('a,a,b', [[2, 2], [0, 0]], [], 1, 0, 10, 0, 'deployment'))
    rng = random.Random(99)
    for i in range(60):
        nq = rng.randint(1, 5)
        names = ['q%d' % j for j in range(nq)]
        T = rng.randint(1, 40)
        depths = [[rng.choice([0, 0, 0, rng.randint(0, 30)]) for _ in names] for _ in range(T)]
        kpp = rng.choice([1, 1, 2, 3, 5, 10])
        mn = rng.choice([0, 0, 0, 1, 2])
        mx = rng.choice([1, 2, 4, 8, 16])
        cur0 = rng.choice([0, 0, 1, 3, 9, 20])
        rtype = rng.choice(['deployment', 'job'])
        ks = []
        for q in names:
            ks += ['processing-%s:h%d' % (q, k) for k in range(rng.choice([0, 0, 1, 3]))]
        cases.append(_run_scale(','.join(names), depths, ks, kpp, mn, mx, cur0, rtype))
    # patch calls that raise ApiException: scale() logs and carries on, replicas stay
    cases.append(_run_scale('a', [[3], [3], [0], [5], [5]], [], 1, 0, 9, 0, 'deployment',
                            fails=[1, 0, 1, 0, 1]))
    for i in range(25):
        nq = rng.randint(1, 3)
        names = ['f%d' % j for j in range(nq)]
        T = rng.randint(2, 40)
        depths = [[rng.choice([0, 0, rng.randint(0, 20)]) for _ in names] for _ in range(T)]
        fails = [int(rng.random() < 0.35) for _ in range(T)]
        cases.append(_run_scale(','.join(names), depths, [], rng.choice([1, 2, 3]), rng.choice([0, 1]),
                                rng.choice([2, 5, 9]), rng.choice([0, 2, 12]),
                                rng.choice(['deployment', 'job']), fails=fails))
    return cases


ADVERSARIAL_KEYS = [
    'processing-predict:host-1', 'processing-predict:', 'processing-predict',
    'processing-predictx:host', 'processing-predic:host', 'xprocessing-predict:host',
    'Processing-predict:host', 'processing-predict:zstack:host', 'processing-predict:zstack',
    'processing-predict:zstack:', 'processing-track:a', 'processing-tracker:a',
    'processing-track:' + 'x' * 600, 'processing-track', 'predict', 'track', 'tracker',
    'predict:file.tiff:0f3c', 'processing-', 'processing', '', 'p',
    'processing-track:\x00', 'processing-track\x00:', 'processing-track:\xff\xfe',
    'processing-tr\xe9:z', 'processing-::', 'processing-:', 'processing-a:b:c:d:e:f',
    'processing-train:1', 'processing-segment-0123456789:abc',
]


def tally_cases():
    cases = []

    def run(queues, keyspace, lists, delim=','):
        redis = refload.StubRedis(lists=lists, keyspace=[k.encode('latin-1') for k in keyspace])
        s = _scaler(queues, delim, redis)
        s.tally_queues()
        cases.append(dict(queues=queues, delim=delim, keyspace=keyspace, lists=lists,
                          redis_keys=dict(s.redis_keys)))

    run('predict,track', ADVERSARIAL_KEYS, {'predict': 3, 'track': 0})
    run('predict,predict:zstack,track,tracker,train', ADVERSARIAL_KEYS,
        {'predict': 1, 'predict:zstack': 2, 'tracker': 7})
    run('', ADVERSARIAL_KEYS, {})
    run('tr\xe9,a,a:b,a:b:c,segment-0123456789', ADVERSARIAL_KEYS, {'a': 5})
    run('predict,track', [], {'predict': 9})
    # reference test shape (autoscaler_test.py:184-210): LLEN only
    run('queue', ['queue'], {'queue': 7})
    run('predict,track,train', ['predict', 'track', 'train'], {'predict': 4, 'track': 4, 'train': 4})
    # glob metacharacters in a queue name are live globs in the reference (autoscaler.py:69)
    run('pre*,tra?k,[pt]r*,a\\*b', ADVERSARIAL_KEYS + ['processing-a*b:1', 'processing-axb:1'], {})
    rng = random.Random(7)
    alphabet = 'abc:-x'
    for _ in range(25):
        names = sorted({''.join(rng.choice('abc:') for _ in range(rng.randint(1, 4)))
                        for _ in range(rng.randint(1, 6))})   # sorted: set order depends on the hash seed
        keys = []
        for _ in range(rng.randint(0, 120)):
            base = rng.choice(['processing-', 'processing-', 'processing', 'xprocessing-', ''])
            keys.append(base + ''.join(rng.choice(alphabet) for _ in range(rng.randint(0, 8))))
        lists = {q: rng.randint(0, 5) for q in names if rng.random() < 0.5}
        run(','.join(names) if ',' not in ''.join(names) else '|'.join(names), keys, lists,
            delim=',' if ',' not in ''.join(names) else '|')
    return cases


def c1():
    """BASELINE config 1: QUEUES='predict,track', 10k keys, 10k ticks."""
    from oracle import coracle
    rng = random.Random(0)
    hexd = '0123456789abcdef'

    def host():
        return 'consumer-%s-%s' % (''.join(rng.choice(hexd) for _ in range(10)),
                                   ''.join(rng.choice('bcdfghjklmnpqrstvwxz2456789') for _ in range(5)))

    keys = []
    for _ in range(10000):
        u = rng.random()
        if u < 0.30:
            keys.append('processing-predict:' + host())
        elif u < 0.50:
            keys.append('processing-track:' + host())
        elif u < 0.52:
            keys.append('processing-track:')
        elif u < 0.60:
            keys.append('processing-predictx:' + host())
        elif u < 0.65:
            keys.append('processing-predict')
        elif u < 0.70:
            keys.append('xprocessing-predict:' + host())
        else:
            keys.append('%s:%s.tiff:%s' % (rng.choice(['predict', 'track']),
                                           ''.join(rng.choice(hexd) for _ in range(8)),
                                           ''.join(rng.choice(hexd) for _ in range(32))))
    keys += ['predict', 'track']
    T = 10000
    llen = coracle.trace_fill(0, T, 0, 2)          # [T][2] bursty, >= 30 % zero ticks
    names = ['predict', 'track']

    class TickStub(refload.StubRedis):
        """The snapshot's match result per pattern is computed once with the stub's
        matcher; at tick t only the first ``alive[pattern]`` matching keys exist
        (in-progress keys come and go), and the reference still iterates and
        counts every yielded key itself (autoscaler.py:71)."""
        _cache = {}
        alive = None

        def scan_iter(self, match=None, count=None):
            if match not in self._cache:
                self._cache[match] = list(refload.StubRedis.scan_iter(self, match, count))
            hits = self._cache[match]
            if self.alive is None:
                return iter(hits)
            return iter(hits[:self.alive[match]])

    redis = TickStub(keyspace=[k.encode('latin-1') for k in keys])
    s0 = _scaler('predict,track', ',', redis)
    s0.tally_queues()
    snapshot_counts = [s0.redis_keys[q] for q in names]   # LLEN is 0 here -> pure SCAN counts
    # per-tick in-progress counts: a slice of the snapshot proportional to the backlog
    inprog = np.minimum(np.array(snapshot_counts, dtype=np.int64)[None, :], llen.astype(np.int64) // 8)

    policies = [(1, 0, 1)]                           # scale.py:90-92 defaults
    for kpp in (1, 2, 5, 10, 50):
        for mn in (0, 1):
            for mx in (1, 4, 16, 64):
                policies.append((kpp, mn, mx))
    dense_idx = [0, 7, 23]
    pod_ticks, n_patches, final, dense = [], [], [], {}
    tallies_default = None
    for si, (kpp, mn, mx) in enumerate(policies):
        kube = refload.StatefulKube('r', 'deployment', 0)
        s = _scaler('predict,track', ',', redis, kube)
        after = np.zeros(T, dtype=np.int32)
        tl = np.zeros((T, 2), dtype=np.int64)
        for t in range(T):
            redis.lists = {'predict': int(llen[t, 0]), 'track': int(llen[t, 1])}
            redis.alive = {'processing-predict:*': int(inprog[t, 0]),
                           'processing-track:*': int(inprog[t, 1])}
            s.scale('ns', 'deployment', 'r', min_pods=mn, max_pods=mx, keys_per_pod=kpp)
            after[t] = kube.value
            if si == 0:
                tl[t] = [s.redis_keys['predict'], s.redis_keys['track']]
        if si == 0:
            tallies_default = tl
        pod_ticks.append(int(after.astype(np.int64).sum()))
        n_patches.append(len(kube.patches))
        final.append(int(kube.value))
        if si in dense_idx:
            dense[si] = after
        print('c1 policy %d/%d' % (si + 1, len(policies)), flush=True)
    blob = '\n'.join(keys).encode('latin-1')
    np.savez_compressed(
        os.path.join(GOLD, 'c1.npz'),
        keys_blob=np.frombuffer(blob, dtype=np.uint8), llen=llen,
        snapshot_counts=np.array(snapshot_counts, dtype=np.int64),
        inprog=inprog.astype(np.int32),
        policies=np.array(policies, dtype=np.int32),
        pod_ticks=np.array(pod_ticks, dtype=np.int64),
        n_patches=np.array(n_patches, dtype=np.int64),
        final=np.array(final, dtype=np.int32),
        dense_idx=np.array(dense_idx, dtype=np.int32),
        dense=np.stack([dense[i] for i in dense_idx]),
        tallies_default=tallies_default)


def _run_only_running(queues, depths, kpp, mn, mx, cur0, rtype, delay, fails=None):
    """scale() driven with the reference's only_running branch (autoscaler.py:176-177) against
    a Kubernetes double whose available_replicas lag spec.replicas by `delay` ticks
    (refload.StatefulKube).  For a job the reference ignores the flag (autoscaler.py:185-190)."""
    names = queues.split(',')
    redis = refload.StubRedis()
    kube = refload.StatefulKube('r', rtype, cur0, startup_delay=delay)
    s = refload.make_reference_scaler(redis, queues, ',', kube, only_running=True)
    after, seen = [], []
    for t, row in enumerate(depths):
        redis.lists = dict(zip(names, row))
        kube.fail_next_patch = bool(fails and fails[t])
        seen.append(s.get_current_pods('ns', rtype, 'r'))
        s.scale('ns', rtype, 'r', min_pods=mn, max_pods=mx, keys_per_pod=kpp)
        kube.end_tick()
        after.append(int(kube.value))
    return dict(queues=queues, depths=depths, kpp=kpp, min=mn, max=mx, cur0=cur0, type=rtype,
                delay=delay, fails=fails, after=after, seen=seen,
                patches=[int(p) for p in kube.patches])


def only_running_cases():
    import random
    rnd = random.Random(20261107)
    cases = []
    # a burst, a drain and a second burst: the start-up lag lets the target overshoot nothing
    # but delays the hold ("never scale down to a non-zero count" compares with AVAILABLE pods)
    ramp = [[0], [3], [6], [6], [6], [2], [2], [0], [0], [5], [5], [5], [1], [0], [0], [0]]
    for delay in (0, 1, 2, 3, 5):
        cases.append(_run_only_running('q', ramp, 1, 0, 8, 0, 'deployment', delay))
        cases.append(_run_only_running('q', ramp, 2, 1, 4, 3, 'deployment', delay))
        cases.append(_run_only_running('q', ramp, 1, 0, 8, 0, 'job', delay))
    for _ in range(60):
        nq = rnd.choice([1, 1, 2, 3])
        T = rnd.randint(5, 40)
        depths = []
        level = 0
        for t in range(T):
            if rnd.random() < 0.25:
                level = rnd.choice([0, 0, 1, 4, 9, 30])
            depths.append([max(0, level + rnd.randint(-2, 2)) if level else 0 for _ in range(nq)])
        fails = [rnd.random() < 0.15 for _ in range(T)] if rnd.random() < 0.3 else None
        cases.append(_run_only_running(','.join('q%d' % i for i in range(nq)), depths,
                                       rnd.choice([1, 1, 2, 5]), rnd.choice([0, 0, 1]),
                                       rnd.choice([1, 4, 16]), rnd.choice([0, 0, 2, 20]),
                                       rnd.choice(['deployment', 'deployment', 'job']),
                                       rnd.choice([0, 1, 2, 3, 4, 7]), fails))
    return cases


def _first_pod_stats(tally, seen, running):
    """The written definition (include/kas.h, kas_first_pod_latency) applied to what the
    reference did: tally[t] = sum of redis_keys after tally_queues, seen[t] = what scale() read
    through get_current_pods on tick t, running[t] = what the same call returns after tick t."""
    out = dict(n_served=0, n_unserved=0, wait_sum=0, wait_max=0, pending=0)
    waiting, prev = -1, 0
    for t in range(len(tally)):
        if waiting < 0 and tally[t] > 0 and prev <= 0 and seen[t] <= 0:
            waiting = t
        if waiting >= 0:
            if running[t] > 0:
                w = t - waiting
                out['n_served'] += 1
                out['wait_sum'] += w
                out['wait_max'] = max(out['wait_max'], w)
                waiting = -1
            elif tally[t] <= 0:
                out['n_unserved'] += 1
                waiting = -1
        prev = tally[t]
    if waiting >= 0:
        out['pending'] = len(tally) - waiting
    return out


def _run_first_pod(queues, depths, kpp, mn, mx, cur0, rtype, delay, fails=None):
    """scale() of the reference over a trace, reading pods through its own get_current_pods
    (only_running=True when a start-up delay is modelled, autoscaler.py:176-177), with the reads
    before and after every tick and the tally it computed recorded."""
    names = queues.split(',')
    redis = refload.StubRedis()
    kube = refload.StatefulKube('r', rtype, cur0, startup_delay=delay)
    s = refload.make_reference_scaler(redis, queues, ',', kube, only_running=delay > 0)
    seen, running, tally = [], [], []
    for t, row in enumerate(depths):
        redis.lists = dict(zip(names, row))
        kube.fail_next_patch = bool(fails and fails[t])
        seen.append(s.get_current_pods('ns', rtype, 'r'))
        s.scale('ns', rtype, 'r', min_pods=mn, max_pods=mx, keys_per_pod=kpp)
        kube.end_tick()
        running.append(s.get_current_pods('ns', rtype, 'r'))
        tally.append(sum(s.redis_keys.values()))
    return dict(queues=queues, depths=depths, kpp=kpp, min=mn, max=mx, cur0=cur0, type=rtype,
                delay=delay if rtype == 'deployment' else 0, fails=fails, seen=seen, running=running,
                stats=_first_pod_stats(tally, seen, running))


def first_pod_cases():
    import random
    rnd = random.Random(20261108)
    cases = []
    bursts = [[0], [0], [3], [6], [0], [0], [1], [1], [9], [0], [2], [0], [0], [0], [7], [7]]
    for delay in (0, 1, 3):
        cases.append(_run_first_pod('q', bursts, 1, 0, 8, 0, 'deployment', delay))
        cases.append(_run_first_pod('q', bursts, 4, 0, 8, 0, 'deployment', delay))    # small bursts stay below one pod
        cases.append(_run_first_pod('q', bursts, 1, 1, 4, 0, 'deployment', delay))    # min_pods keeps one up
        cases.append(_run_first_pod('q', bursts, 2, 0, 8, 0, 'job', delay))
    for _ in range(70):
        nq = rnd.choice([1, 1, 2, 3])
        T = rnd.randint(8, 60)
        depths, level = [], 0
        for t in range(T):
            if rnd.random() < 0.3:
                level = rnd.choice([0, 0, 0, 1, 2, 5, 12])
            depths.append([max(0, level + rnd.randint(-1, 1)) if level else 0 for _ in range(nq)])
        fails = [rnd.random() < 0.2 for _ in range(T)] if rnd.random() < 0.3 else None
        cases.append(_run_first_pod(','.join('q%d' % i for i in range(nq)), depths,
                                    rnd.choice([1, 2, 3, 6]), rnd.choice([0, 0, 0, 1]),
                                    rnd.choice([1, 4, 16]), rnd.choice([0, 0, 0, 3]),
                                    rnd.choice(['deployment', 'deployment', 'job']),
                                    rnd.choice([0, 0, 1, 2, 4]), fails))
    return cases


def main():
    assert refload.available(), 'needs /root/reference'
    import logging
    logging.disable(logging.CRITICAL)
    os.makedirs(GOLD, exist_ok=True)
    only = set(sys.argv[1:])          # fixture names to (re)write; none given: all of them
    for name, fn in [('known_answers', known_answers), ('decisions', decisions),
                     ('scale_cases', scale_cases), ('tally_cases', tally_cases),
                     ('only_running_cases', only_running_cases),
                     ('first_pod_cases', first_pod_cases)]:
        if only and name not in only:
            continue
        with open(os.path.join(GOLD, name + '.json'), 'w') as f:
            json.dump(fn(), f, separators=(',', ':'))
        print('wrote', name)
    if not only or 'c1' in only:
        c1()
        print('wrote c1')


if __name__ == '__main__':
    main()
