"""Where the reference checkout is present (the build container), drive its
UNMODIFIED source with fresh random inputs and compare with the oracle.  On the
GPU box /root/reference does not exist and these tests skip; the committed
fixtures (tests/golden) carry the same evidence there."""
import random

import pytest

from oracle import oracle, refload

pytestmark = pytest.mark.skipif(not refload.available(), reason='reference checkout not present')


def test_reference_decisions_random():
    s = refload.make_reference_scaler(refload.StubRedis(), 'q')
    rng = random.Random(2024)
    for _ in range(5000):
        tally = rng.choice([rng.randint(0, 300), rng.randint(-40, 40), rng.randint(0, 2 ** 32)])
        kpp = rng.choice([1, 2, 3, 7, 10, 50, -1, -4, 2 ** 31 - 1])
        mn, mx, cur = rng.randint(-3, 12), rng.randint(-3, 70), rng.randint(-2, 80)
        s.redis_keys['q'] = tally
        assert s.get_desired_pods('q', kpp, mn, mx, cur) == oracle.get_desired_pods(tally, kpp, mn, mx, cur)
        assert s.clip_pod_count(tally, mn, mx, cur) == oracle.clip_pod_count(tally, mn, mx, cur)


def test_reference_scale_trajectories_random():
    rng = random.Random(77)
    for _ in range(40):
        nq = rng.randint(1, 4)
        names = ['n%d' % i for i in range(nq)]
        keys = []
        for q in names:
            keys += [('processing-%s:%d' % (q, i)).encode() for i in range(rng.randint(0, 4))]
        keys += [b'processing-n0', b'processing-n0x:1', b'n0']
        kube = refload.StatefulKube('r', rng.choice(['deployment', 'job']), rng.randint(0, 6))
        redis = refload.StubRedis(keyspace=keys)
        s = refload.make_reference_scaler(redis, ','.join(names), ',', kube)
        kpp, mn, mx = rng.choice([1, 2, 5]), rng.choice([0, 0, 1]), rng.choice([1, 3, 10])
        cur0 = kube.value
        rows, after = [], []
        for _t in range(rng.randint(1, 30)):
            row = [rng.choice([0, 0, rng.randint(0, 25)]) for _ in names]
            rows.append(row)
            redis.lists = dict(zip(names, row))
            s.scale('ns', kube.resource_type, 'r', min_pods=mn, max_pods=mx, keys_per_pod=kpp)
            after.append(int(kube.value))
        inprog = oracle.match_counts(names, keys)
        got, patches = oracle.replay_resource(rows, [inprog] * len(rows), kpp, mn, mx, cur0)
        assert got == after and patches == [int(p) for p in kube.patches]


def test_reference_only_running_trajectories_random():
    """scale() read through the reference's own only_running branch (autoscaler.py:176-177)
    against a Kubernetes double whose available_replicas lag the spec: the oracle's
    start-up-delay replay follows it tick by tick (jobs ignore the flag, :185-190)."""
    rng = random.Random(4242)
    for _ in range(60):
        nq = rng.randint(1, 3)
        names = ['n%d' % i for i in range(nq)]
        rtype = rng.choice(['deployment', 'deployment', 'job'])
        delay = rng.choice([0, 1, 2, 4, 6])
        cur0 = rng.choice([0, 0, 3, 9])
        kube = refload.StatefulKube('r', rtype, cur0, startup_delay=delay)
        redis = refload.StubRedis()
        s = refload.make_reference_scaler(redis, ','.join(names), ',', kube, only_running=True)
        kpp, mn, mx = rng.choice([1, 2, 5]), rng.choice([0, 0, 1]), rng.choice([1, 3, 10])
        rows, after, fails = [], [], []
        level = 0
        for t in range(rng.randint(5, 50)):
            if rng.random() < 0.3:
                level = rng.choice([0, 0, 2, 7, 25])
            row = [max(0, level + rng.randint(-1, 1)) if level else 0 for _ in names]
            rows.append(row)
            redis.lists = dict(zip(names, row))
            fails.append(rng.random() < 0.1)
            kube.fail_next_patch = fails[-1]
            s.scale('ns', rtype, 'r', min_pods=mn, max_pods=mx, keys_per_pod=kpp)
            kube.end_tick()
            after.append(int(kube.value))
        got, patches = oracle.replay_resource(rows, None, kpp, mn, mx, cur0, fails,
                                              startup_delay=delay if rtype == 'deployment' else 0)
        assert got == after
        assert patches == [int(p) for p in kube.patches]
