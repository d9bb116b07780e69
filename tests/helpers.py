"""Shared test helpers: golden fixtures, seeded inputs, doubles."""
import json
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def golden_json(name):
    with open(os.path.join(GOLD, name + '.json')) as f:
        return json.load(f)


def golden_c1():
    d = np.load(os.path.join(GOLD, 'c1.npz'))
    out = {k: d[k] for k in d.files}
    out['keys'] = bytes(out['keys_blob']).split(b'\n')
    return out


def latin(keys):
    return [k.encode('latin-1') for k in keys]


C2_NAMES = None


def c2_queue_names(n=64):
    """64 queue names, 3-20 bytes, with nested and near-miss pairs."""
    base = ['predict', 'predict:zstack', 'track', 'tracker', 'train', 'segment', 'seg',
            'segment:nuclear', 'segment:cyto', 'spot', 'spots', 'caliban', 'polaris', 'mesmer',
            'mesmer:whole-cell', 'a:b', 'a:b:c', 'abc']
    i = 0
    while len(base) < n:
        base.append('queue-%02d-%s' % (i, 'x' * (i % 9)))
        i += 1
    return base[:n]


class Bunch(object):
    def __init__(self, **kw):
        self.__dict__.update(kw)


class ApiException(Exception):
    pass


class DummyKubernetes(object):
    """Same shape as the reference's test double (autoscaler_test.py:51-81):
    string-typed counts, two resources pod1 / pod2."""

    def list_namespaced_deployment(self, *_, **__):
        return Bunch(items=[
            Bunch(spec=Bunch(replicas='4'), metadata=Bunch(name='pod1'),
                  status=Bunch(available_replicas=None)),
            Bunch(spec=Bunch(replicas='8'), metadata=Bunch(name='pod2'),
                  status=Bunch(available_replicas='8'))])

    def list_namespaced_job(self, *_, **__):
        return Bunch(items=[
            Bunch(spec=Bunch(completions='1', parallelism='1'), metadata=Bunch(name='pod1')),
            Bunch(spec=Bunch(completions='2', parallelism='2'), metadata=Bunch(name='pod2'))])

    def patch_namespaced_deployment(self, *_, **__):
        return Bunch(items=[Bunch(spec=Bunch(replicas='4'), metadata=Bunch(name='pod'))])

    def patch_namespaced_job(self, *_, **__):
        return Bunch(items=[Bunch(spec=Bunch(completions='0', parallelism='0'),
                                  metadata=Bunch(name='pod'))])


class StatefulKube(object):
    """A patch stores what the next list returns."""

    def __init__(self, name, replicas=0):
        self.name, self.value, self.patches = name, replicas, []

    def _items(self):
        return Bunch(items=[Bunch(spec=Bunch(replicas=str(self.value), parallelism=str(self.value)),
                                  metadata=Bunch(name=self.name),
                                  status=Bunch(available_replicas=None))])

    list_namespaced_deployment = lambda self, ns: self._items()   # noqa: E731
    list_namespaced_job = lambda self, ns: self._items()          # noqa: E731

    def patch_namespaced_deployment(self, name, ns, body):
        self.value = body['spec']['replicas']
        self.patches.append(self.value)

    def patch_namespaced_job(self, name, ns, body):
        self.value = body['spec']['parallelism']
        self.patches.append(self.value)


class ListRedis(object):
    """Minimal live-server stand-in: llen + scan_iter over a key list, using
    the oracle's glob (test infrastructure)."""

    def __init__(self, lists=None, keyspace=()):
        self.lists = dict(lists or {})
        self.keyspace = list(keyspace)
        self.scans = []

    def llen(self, name):
        return int(self.lists.get(name, 0))

    def scan_iter(self, match=None, count=None):
        from oracle.oracle import redis_glob_match
        self.scans.append(match)
        pat = match.encode('utf-8') if isinstance(match, str) else match
        return (k for k in self.keyspace if redis_glob_match(pat, k))


def random_fleet(rng, T, Q, grouped=False, neg=False):
    llen = rng.integers(0, 40, size=(T, Q)).astype(np.int32)
    llen[rng.random((T, Q)) < 0.5] = 0
    if neg:
        llen[rng.random((T, Q)) < 0.05] = -rng.integers(1, 50)
    group_off = None
    if grouped:
        sizes = []
        left = Q
        while left > 0:
            s = int(min(left, rng.integers(1, 9)))
            sizes.append(s)
            left -= s
        group_off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
    return llen, group_off
