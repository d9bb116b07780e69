"""Load the UNMODIFIED reference sources from /root/reference -- TEST INFRASTRUCTURE.

The reference is pure Python but imports ``kubernetes`` and ``redis`` at module
top (autoscaler/autoscaler.py:34, autoscaler/redis.py:35); neither is installed
here and there is no network.  The hot path touches none of their symbols, so
the two source files are executed as they lie on disk behind minimal stub
modules.  The files are loaded under private module names so they never shadow
this repository's own ``autoscaler`` package.

Only usable where /root/reference exists (the build container); the GPU box
uses the fixtures this produces (tests/golden/, see make_golden.py).
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get('KAS_REFERENCE_ROOT', '/root/reference')


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, 'autoscaler', 'autoscaler.py'))


class ApiException(Exception):
    """Stand-in for kubernetes.client.rest.ApiException."""


def _install_stubs():
    if 'kubernetes' not in sys.modules:
        k = types.ModuleType('kubernetes')
        k.client = types.ModuleType('kubernetes.client')
        k.client.rest = types.ModuleType('kubernetes.client.rest')
        k.client.rest.ApiException = ApiException
        k.config = types.ModuleType('kubernetes.config')
        k.config.load_incluster_config = lambda: None
        k.client.AppsV1Api = object
        k.client.BatchV1Api = object
        sys.modules['kubernetes'] = k
        sys.modules['kubernetes.client'] = k.client
        sys.modules['kubernetes.client.rest'] = k.client.rest
        sys.modules['kubernetes.config'] = k.config
    if 'redis' not in sys.modules:
        r = types.ModuleType('redis')
        r.exceptions = types.ModuleType('redis.exceptions')

        class ConnectionError(Exception):  # noqa: A001 - mirrors redis-py's name
            pass

        class ResponseError(Exception):
            pass

        r.exceptions.ConnectionError = ConnectionError
        r.exceptions.ResponseError = ResponseError
        r.StrictRedis = object
        sys.modules['redis'] = r
        sys.modules['redis.exceptions'] = r.exceptions


def _load(name, relpath):
    path = os.path.join(REFERENCE_ROOT, relpath)
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


_cache = {}


def reference_autoscaler_module():
    """The reference's autoscaler/autoscaler.py, executed from where it lies."""
    if 'a' not in _cache:
        _install_stubs()
        _cache['a'] = _load('_kas_reference_autoscaler', 'autoscaler/autoscaler.py')
    return _cache['a']


def reference_redis_module():
    """The reference's autoscaler/redis.py, executed from where it lies."""
    if 'r' not in _cache:
        _install_stubs()
        _cache['r'] = _load('_kas_reference_redis', 'autoscaler/redis.py')
    return _cache['r']


# --------------------------------------------------------------------------
# Doubles the reference is driven with
# --------------------------------------------------------------------------
class StubRedis(object):
    """In-memory Redis exposing what autoscaler.py:67,70 call.

    ``lists`` maps list-key -> length; ``keyspace`` is every key of the
    database as bytes (the list keys themselves are part of it, like in a real
    server).  ``scan_iter`` applies the Redis glob to every key, yielding each
    matching key once (a real SCAN may repeat keys; the reference does not
    dedup, autoscaler.py:71).
    """

    def __init__(self, lists=None, keyspace=()):
        self.lists = dict(lists or {})
        self.keyspace = list(keyspace)

    def llen(self, name):
        return int(self.lists.get(name, 0))

    def scan_iter(self, match=None, count=None):
        from oracle.oracle import redis_glob_match
        pat = match.encode('utf-8') if isinstance(match, str) else match
        for k in self.keyspace:
            if pat is None or redis_glob_match(pat, k):
                yield k


class _Bunch(object):
    def __init__(self, **kw):
        self.__dict__.update(kw)


class StatefulKube(object):
    """Kubernetes double: a patch stores what the next list returns
    (autoscaler.py:232,236 -> :179,189).  Counts are returned as strings the
    way the reference's own DummyKubernetes does (autoscaler_test.py:54-81)."""

    def __init__(self, name, resource_type, replicas=0, startup_delay=0):
        self.name = name
        self.resource_type = resource_type
        self.value = replicas
        self.patches = []
        self.fail_next_patch = False   # raise ApiException on the next patch call (it is still listed)
        # availability model (only read through `only_running`, autoscaler.py:176-177): a
        # replica is available once it has been in the spec for `startup_delay` ticks in a
        # row; the driver calls end_tick() after every scale()
        self.startup_delay = startup_delay
        self.history = [replicas] * startup_delay
        self.available = replicas
        self.ticks = 0

    def end_tick(self):
        self.available = min([int(self.value)] + [int(h) for h in self.history])
        if self.startup_delay:
            self.history[self.ticks % self.startup_delay] = self.value
        self.ticks += 1

    def _items(self):
        spec = _Bunch(replicas=str(self.value), parallelism=str(self.value),
                      completions=str(self.value))
        return _Bunch(items=[
            _Bunch(spec=_Bunch(replicas='123', parallelism='123'),
                   metadata=_Bunch(name=self.name + '-other'),
                   status=_Bunch(available_replicas=None)),
            _Bunch(spec=spec, metadata=_Bunch(name=self.name),
                   status=_Bunch(available_replicas=str(self.available if self.startup_delay
                                                        else self.value)))])

    def list_namespaced_deployment(self, namespace):
        return self._items()

    def list_namespaced_job(self, namespace):
        return self._items()

    def _patch(self, name, namespace, body, field):
        assert name == self.name
        self.patches.append(body['spec'][field])
        if self.fail_next_patch:
            raise sys.modules['kubernetes'].client.rest.ApiException('thrown on purpose')
        self.value = body['spec'][field]
        return _Bunch(items=[])

    def patch_namespaced_deployment(self, name, namespace, body):
        return self._patch(name, namespace, body, 'replicas')

    def patch_namespaced_job(self, name, namespace, body):
        return self._patch(name, namespace, body, 'parallelism')


def make_reference_scaler(redis_client, queues, queue_delim=',', kube=None, only_running=False):
    """A reference ``Autoscaler`` whose k8s clients are the given double.
    ``only_running``: ``scale()`` reads the current pods with the reference's own
    ``only_running=True`` branch (autoscaler.py:176-177), which ``scale()`` itself never
    selects (autoscaler.py:252) -- the only line of ours on that path is the flag."""
    import logging
    mod = reference_autoscaler_module()
    cls = mod.Autoscaler
    if only_running:
        class OnlyRunning(mod.Autoscaler):
            def get_current_pods(self, namespace, resource_type, name, only_running=False):
                return mod.Autoscaler.get_current_pods(self, namespace, resource_type, name,
                                                       only_running=True)
        cls = OnlyRunning
    scaler = cls(redis_client, queues=queues, queue_delim=queue_delim)
    scaler.logger.setLevel(logging.CRITICAL)
    if kube is not None:
        scaler.get_apps_v1_client = lambda: kube
        scaler.get_batch_v1_client = lambda: kube
    return scaler
