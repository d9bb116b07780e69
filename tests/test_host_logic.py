"""CPU-only checks: the C-ABI library loads and exports what include/kas.h
declares, host-side ingest helpers, table builder validation, the drop-in
package surface, scale.py configuration, the RedisClient wrapper and the
Kubernetes-side methods (mirroring the reference's own tests)."""
import ctypes
import os
import re
import sys
import types

import numpy as np
import pytest

from tests import helpers

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from autoscaler import engine
    hdr = open(os.path.join(ROOT, 'include', 'kas.h')).read()
    declared = set(re.findall(r'KAS_API\s+[\w\s\*]+?\b(kas_\w+)\s*\(', hdr))
    assert len(declared) >= 16
    L = engine.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert declared == set(engine.SYMBOLS), declared ^ set(engine.SYMBOLS)
    assert L.kas_version() == 103
    assert L.kas_strerror(-5) == b'integer division or modulo by zero'
    assert L.kas_rec_bytes_for(7) == 32 and L.kas_rec_bytes_for(4) == 16 and L.kas_rec_bytes_for(21) == 48
    # struct layouts the header promises
    assert ctypes.sizeof(engine.Policy) == 12 and ctypes.sizeof(engine.Totals) == 16
    assert engine.TOTALS_DTYPE.itemsize == 16


def test_no_cuda_means_loud_failure():
    import torch
    from autoscaler import engine
    if torch.cuda.is_available():
        pytest.skip('CUDA present')
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        engine.decide([1], [1], [0], [1], [0])
    import autoscaler
    s = autoscaler.Autoscaler(helpers.ListRedis(), 'q')
    with pytest.raises(RuntimeError):
        s.tally_queues()
    with pytest.raises(RuntimeError):
        s.get_desired_pods('q', 1, 0, 1, 0)


def test_missing_library_is_an_import_error(monkeypatch):
    from autoscaler import engine
    monkeypatch.setattr(engine, '_lib', None)
    monkeypatch.setattr(engine, 'LIB_PATH', '/nonexistent/libkas.so')
    with pytest.raises(ImportError, match='no CPU fallback'):
        engine.lib()


def test_prefix_table_builder_validation():
    from autoscaler import engine
    t = engine.build_prefix_table(['predict', 'track'], 32).view(np.uint32)
    assert t[1] == 32 and t[2] == 2 and t[4] == (1 << 10) | (1 << 8) and (t[7] >> 8) == 0
    # nested names add cut entries: "a:" (virtual) for "a:b"
    t = engine.build_prefix_table(['a:b', 'abc'], 32).view(np.uint32)
    assert t[4] == (1 << 4) | (1 << 6)   # colon positions count the "ng-" of the literal
    with pytest.raises(engine.GlobQueueNameError):
        engine.build_prefix_table(['tra?k'], 32)
    for bad in (['a', 'a'], ['x' * 21], ['a\0b']):
        with pytest.raises(engine.KasError):
            engine.build_prefix_table(bad, 32)
    with pytest.raises(ValueError):
        engine.build_prefix_table([], 32)
    # a big table falls back to two-slot placement and still builds
    t = engine.build_prefix_table(['n%d' % i for i in range(1024)], 32).view(np.uint32)
    assert (t[7] >> 8) == 1


def test_pack_and_index_helpers_match_oracle():
    from autoscaler import engine, snapshot
    from oracle import oracle
    keys = [b'processing-a:1', b'', b'x' * 100, b'processing-predict:zstack:host', b'\x00\xff\n'[:2]]
    snap = snapshot.KeySnapshot.from_keys(keys)
    for W in (16, 32, 48, 64):
        assert (snap.heads(W) == oracle.pack_heads(keys, W)).all()
    dump = b'\n'.join([b'k1', b'processing-q:abc', b'', b'last-without-newline'])
    s2 = snapshot.KeySnapshot.from_dump(dump)
    assert len(s2) == 4
    assert (s2.heads(32) == oracle.pack_heads(dump.split(b'\n'), 32)).all()
    s3 = snapshot.KeySnapshot.from_dump(dump + b'\n')
    assert len(s3) == 4 and (s3.heads(16) == s2.heads(16)).all()
    assert len(snapshot.KeySnapshot.from_dump(b'')) == 0
    with pytest.raises(ValueError):
        snapshot.KeySnapshot.from_heads(np.zeros((3, 16), np.uint8)).heads(32)


def test_synthetic_keys_are_deterministic_and_sliceable():
    from autoscaler import engine
    names = helpers.c2_queue_names(64)
    h1, b1, o1 = engine.synth_keys_host(7, 0, 5000, names, long_tail=True, rec_bytes=32, want_blob=True)
    h2, _, _ = engine.synth_keys_host(7, 1000, 100, names, long_tail=True, rec_bytes=32)
    assert (h1[1000:1100] == h2).all()
    h3, _, _ = engine.synth_keys_host(7, 0, 5000, names, long_tail=True, rec_bytes=32, threads=1)
    assert (h1 == h3).all()
    lens = np.diff(o1.astype(np.int64))
    assert lens.max() > 300 and (engine.pack_heads_host(b1, o1, 32) == h1).all()
    starts = np.array([bytes(b1[o1[i]:o1[i] + 11]) == b'processing-' for i in range(2000)])
    assert 0.55 < starts.mean() < 0.8


def test_keyspace_generator_twin_in_the_oracle():
    """The oracle's own generator (written from the spec in DESIGN.md) makes the keys the
    library's generator makes: heads, blob and offsets, plain and long-tailed, any start."""
    from autoscaler import engine
    from oracle import coracle
    names = helpers.c2_queue_names(64)
    assert coracle.rec_bytes_for(names) == engine.rec_bytes_for(names) == 32
    for long_tail in (False, True):
        h, b, o = engine.synth_keys_host(11, 123456789, 6000, names, long_tail=long_tail, rec_bytes=32,
                                         want_blob=True)
        assert (coracle.synth_heads(11, 123456789, 6000, names, 32, long_tail=long_tail) == h).all()
        b2, o2 = coracle.synth_blob(11, 123456789, 6000, names, long_tail=long_tail)
        assert (o2 == o).all() and (b2 == b).all()
    few = ['a', 'predict:zstack']
    h, _, _ = engine.synth_keys_host(3, 0, 999, few, rec_bytes=48)
    assert (coracle.synth_heads(3, 0, 999, few, 48, threads=1) == h).all()


def test_package_surface_and_init_parsing():
    import autoscaler
    assert autoscaler.Autoscaler and autoscaler.redis.RedisClient
    for case in helpers.golden_json('known_answers')['init']:
        s = autoscaler.Autoscaler(None, case['queues'], case['delim'])
        assert list(s.redis_keys) == case['keys'] and set(s.redis_keys.values()) == {0}
    s = autoscaler.Autoscaler('client')
    assert s.redis_keys == {'predict': 0} and s.redis_client == 'client'
    assert s.managed_resource_types == {'deployment', 'job'} and s.completed_statuses == {'done', 'failed'}
    assert s.logger.name == 'Autoscaler'
    assert 'llen' in autoscaler.redis.REDIS_READONLY_COMMANDS
    assert 'scan_iter' not in autoscaler.redis.REDIS_READONLY_COMMANDS
    assert len(autoscaler.redis.REDIS_READONLY_COMMANDS) == 83


def _patch_kube(scaler, kube):
    scaler.get_apps_v1_client = lambda: kube
    scaler.get_batch_v1_client = lambda: kube


def test_get_current_pods_like_reference_test():
    # autoscaler_test.py:155-182
    import autoscaler
    scaler = autoscaler.Autoscaler(helpers.ListRedis(), 'queue')
    _patch_kube(scaler, helpers.DummyKubernetes())
    with pytest.raises(ValueError):
        scaler.get_current_pods('namespace', 'bad_type', 'pod')
    assert scaler.get_current_pods('ns', 'deployment', 'pod1') == 4
    assert scaler.get_current_pods('ns', 'deployment', 'pod2') == 8
    assert scaler.get_current_pods('ns', 'deployment', 'pod2', True) == 8
    assert scaler.get_current_pods('ns', 'deployment', 'pod1', True) == 0
    assert scaler.get_current_pods('ns', 'job', 'pod1') == 1
    assert scaler.get_current_pods('ns', 'job', 'pod2') == 2
    assert scaler.get_current_pods('ns', 'job', 'absent') == 0


def test_scale_resource_like_reference_test():
    # autoscaler_test.py:212-236
    import autoscaler
    scaler = autoscaler.Autoscaler(helpers.ListRedis(), 'queue')
    _patch_kube(scaler, helpers.DummyKubernetes())
    assert not scaler.scale_resource(1, 1, 'deployment', 'ns', 'name')
    assert scaler.scale_resource(2, 1, 'job', 'ns', 'name')
    assert scaler.scale_resource(2, 1, 'deployment', 'ns', 'name')
    with pytest.raises(ValueError):
        scaler.scale_resource(2, 1, 'badvalue', 'ns', 'name')


def test_kube_wrappers_log_and_reraise_api_exception():
    # autoscaler_test.py:104-153
    import autoscaler

    class Failing(helpers.DummyKubernetes):
        def _boom(self, *a, **k):
            raise helpers.ApiException('thrown on purpose')
        list_namespaced_deployment = list_namespaced_job = _boom
        patch_namespaced_deployment = patch_namespaced_job = _boom

    scaler = autoscaler.Autoscaler(helpers.ListRedis(), 'queue')
    _patch_kube(scaler, helpers.DummyKubernetes())
    scaler.patch_namespaced_deployment('job', 'ns', {'spec': {'replicas': 1}})
    scaler.patch_namespaced_job('job', 'ns', {'spec': {'parallelism': 1}})
    assert len(scaler.list_namespaced_deployment('ns')) == 2
    _patch_kube(scaler, Failing())
    for call in (lambda: scaler.list_namespaced_deployment('ns'), lambda: scaler.list_namespaced_job('ns'),
                 lambda: scaler.patch_namespaced_deployment('p', 'ns', {}),
                 lambda: scaler.patch_namespaced_job('p', 'ns', {})):
        with pytest.raises(helpers.ApiException):
            call()


# ---- RedisClient wrapper, driven like the reference's redis_test.py ---------
def _fake_redis_module(fail_first=None):
    mod = types.ModuleType('redis')
    mod.exceptions = types.ModuleType('redis.exceptions')

    class ConnectionError(Exception):  # noqa: A001
        pass

    class ResponseError(Exception):
        pass

    mod.exceptions.ConnectionError, mod.exceptions.ResponseError = ConnectionError, ResponseError
    state = {'fail': fail_first, 'made': []}

    class StrictRedis(object):
        def __init__(self, host='h', port=1, **kw):
            self.host, self.port, self.kw = host, port, kw
            state['made'].append(self)

        def sentinel_masters(self):
            return {'mymaster': {'ip': 'master', 'port': 6379}}

        def sentinel_slaves(self, name):
            return [{'ip': 'slave%d' % i, 'port': 6379} for i in range(3)]

        def llen(self, q):
            if state['fail'] is not None:
                err, state['fail'] = state['fail'], None
                raise err
            return 7 if self.host.startswith('slave') else -1

        def lpush(self, q, v):
            return 'master' if self.host == 'master' else 'WRONG'

        def scan_iter(self, match=None, count=None):
            return iter([self.host])

        def boom(self):
            raise ValueError('other error')

    mod.StrictRedis = StrictRedis
    return mod, state


def test_redis_client_routing_and_retries(monkeypatch):
    import autoscaler
    mod, state = _fake_redis_module()
    monkeypatch.setitem(sys.modules, 'redis', mod)
    monkeypatch.setitem(sys.modules, 'redis.exceptions', mod.exceptions)
    c = autoscaler.redis.RedisClient(host='sentinel', port=26379, backoff=0)
    assert c._redis_master.host == 'master' and len(c._redis_slaves) == 3
    assert state['made'][0].kw == {'decode_responses': True, 'charset': 'utf-8'}
    assert c.llen('q') == 7                       # read-only -> a replica
    assert c.lpush('q', 'x') == 'master'          # write -> master
    assert list(c.scan_iter(match='processing-*', count=1000)) == ['master']   # not in the read-only set
    state['fail'] = mod.exceptions.ConnectionError('down')
    assert c.llen('q') == 7                       # retried
    state['fail'] = mod.exceptions.ResponseError('BUSY Redis is busy. SCRIPT KILL')
    assert c.llen('q') == 7                       # retried
    state['fail'] = mod.exceptions.ResponseError('WRONGTYPE')
    with pytest.raises(mod.exceptions.ResponseError):
        c.llen('q')
    with pytest.raises(ValueError):
        c.boom()

    # a plain server answers SENTINEL with a ResponseError: keep using it for everything
    def no_sentinel(self):
        raise mod.exceptions.ResponseError('unknown command')
    monkeypatch.setattr(mod.StrictRedis, 'sentinel_masters', no_sentinel)
    c2 = autoscaler.redis.RedisClient(host='plain', port=6379)
    assert c2._redis_master is c2._sentinel and c2._redis_slaves == [c2._sentinel]


def test_tally_queues_through_redis_client(monkeypatch):
    """The live loop of scale.py hands Autoscaler a RedisClient, whose __getattr__ answers ANY
    attribute name with a command wrapper: the snapshot seam must not be mistaken for present
    (it would forward `key_snapshot` to the server and die on the first tick).  The tally must
    take ONE scan of the keyspace through scan_iter instead."""
    import autoscaler
    mod, state = _fake_redis_module()
    monkeypatch.setitem(sys.modules, 'redis', mod)
    monkeypatch.setitem(sys.modules, 'redis.exceptions', mod.exceptions)
    client = autoscaler.redis.RedisClient(host='sentinel', port=26379, backoff=0)
    assert callable(getattr(client, 'key_snapshot', None))        # the trap
    seen = {}

    def fake_live(redis_client, names, match=None, count=None, **kw):
        seen['keys'] = list(redis_client.scan_iter(match=match, count=count))
        seen['match'], seen['count'] = match, count
        return [3] * len(names)

    def fake_tally(snapshot, names, device='cuda'):
        seen['keys'] = [bytes(snapshot.blob[snapshot.off[i]:snapshot.off[i + 1]])
                        for i in range(len(snapshot))]
        return [3] * len(names)
    monkeypatch.setattr(autoscaler.snapshot, 'tally_live', fake_live)
    monkeypatch.setattr(autoscaler.snapshot, 'tally_snapshot', fake_tally)
    scaler = autoscaler.Autoscaler(client, 'predict,track')
    scaler.tally_queues()
    assert seen['keys'] == ['master']                             # ONE scan_iter, sent to the master
    assert seen['match'] == 'processing-*' and seen['count'] == 1000
    assert scaler.redis_keys == {'predict': 7 + 3, 'track': 7 + 3}   # llen from a replica + scan term
    # the offline stand-in still short-cuts the scan
    snap = autoscaler.snapshot.KeySnapshot.from_keys([b'processing-predict:a'])
    scaler2 = autoscaler.Autoscaler(autoscaler.snapshot.SnapshotRedis(snap, {'predict': 1}), 'predict')
    scaler2.tally_queues()
    assert seen['keys'] == [b'processing-predict:a'] and scaler2.redis_keys == {'predict': 4}
    # a snapshot object can still be taken from a live client, without a list of str in between
    snap2 = autoscaler.snapshot.KeySnapshot.from_redis(client)
    assert len(snap2) == 1 and bytes(snap2.blob[:6]) == b'master'


def test_scale_honours_an_overridden_get_desired_pods(monkeypatch):
    """Reference scale() calls self.get_desired_pods once per key (autoscaler.py:254-257), so a
    subclass or a patched instance method takes part in the decision; the batched launch is
    only used when the method is the stock one."""
    import autoscaler
    from autoscaler import engine
    calls = []

    def fake_decide(a, kpp, mn, mx, cur, mode=0):
        calls.append(mode)
        return [0] * len(a)
    monkeypatch.setattr(engine, 'decide', fake_decide)
    monkeypatch.setattr(autoscaler.snapshot, 'tally_snapshot', lambda snap, names, device='cuda': [0] * len(names))

    class Mine(autoscaler.Autoscaler):
        def get_desired_pods(self, key, keys_per_pod, min_pods, max_pods, current_pods):
            return {'a': 2, 'b': 5}[key]

        def get_current_pods(self, *a, **k):
            return 1
    seen = []
    s = Mine(autoscaler.snapshot.SnapshotRedis(), 'a,b')
    s.clip_pod_count = lambda d, mn, mx, cur: seen.append(d) or d
    s.scale_resource = lambda *a: None
    s.scale('ns', 'deployment', 'x')
    assert seen == [7] and calls == []                  # per-key calls, no batched launch
    t = autoscaler.Autoscaler(autoscaler.snapshot.SnapshotRedis(), 'a,b')
    t.get_current_pods = lambda *a, **k: 1
    t.get_desired_pods = lambda key, *a: 4              # patched on the instance, as the reference's tests do
    t.clip_pod_count = lambda d, mn, mx, cur: seen.append(d) or d
    t.scale_resource = lambda *a: None
    t.scale('ns', 'deployment', 'x')
    assert seen == [7, 8] and calls == []
    u = autoscaler.Autoscaler(autoscaler.snapshot.SnapshotRedis(), 'a,b')
    u.get_current_pods = lambda *a, **k: 1
    u.clip_pod_count = lambda d, mn, mx, cur: d
    u.scale_resource = lambda *a: None
    u.scale('ns', 'deployment', 'x')
    assert calls == [0]                                 # the stock method: one batched launch


def test_api_exception_subclasses_are_swallowed_like_the_reference():
    """`except ApiException` (reference autoscaler.py:267) also catches subclasses."""
    import autoscaler
    from tests import helpers

    class Conflict(helpers.ApiException):
        pass
    s = autoscaler.Autoscaler(autoscaler.snapshot.SnapshotRedis(), 'a')
    s.tally_queues = lambda: None
    s.get_current_pods = lambda *a, **k: 0
    s._get_desired_pods_is_stock = lambda: False
    s.get_desired_pods = lambda *a: 1
    s.clip_pod_count = lambda d, mn, mx, cur: d

    def boom(*a):
        raise Conflict('409')
    s.scale_resource = boom
    s.scale('ns', 'deployment', 'x')                    # swallowed

    def boom2(*a):
        raise KeyError('other')
    s.scale_resource = boom2
    with pytest.raises(KeyError):
        s.scale('ns', 'deployment', 'x')


def test_scale_py_config(monkeypatch):
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        'kas_scale', os.path.join(ROOT, 'kiosk-autoscaler_b200', 'scale.py'))
    scale = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(scale)
    for k in ('REDIS_HOST', 'REDIS_PORT', 'REDIS_INTERVAL', 'QUEUES', 'QUEUE_DELIMITER', 'INTERVAL',
              'RESOURCE_NAMESPACE', 'RESOURCE_TYPE', 'RESOURCE_NAME', 'MIN_PODS', 'MAX_PODS', 'KEYS_PER_POD'):
        monkeypatch.delenv(k, raising=False)
    with pytest.raises(Exception):
        scale.load_config()                       # RESOURCE_NAME has no default (scale.py:88)
    monkeypatch.setenv('RESOURCE_NAME', 'consumer')
    cfg = scale.load_config()
    assert cfg == dict(redis_host='redis-master', redis_port=6379, redis_backoff=1,
                       queues='predict,track', queue_delim=',', interval=5, namespace='default',
                       resource_type='deployment', name='consumer', min_pods=0, max_pods=1,
                       keys_per_pod=1)
    monkeypatch.setenv('MAX_PODS', '12')
    monkeypatch.setenv('QUEUES', 'a|b')
    monkeypatch.setenv('QUEUE_DELIMITER', '|')
    cfg = scale.load_config()
    assert cfg['max_pods'] == 12 and cfg['queues'] == 'a|b' and cfg['queue_delim'] == '|'
    monkeypatch.setenv('KEYS_PER_POD', 'x')
    with pytest.raises(ValueError):
        scale.load_config()


def test_shard_ranges_cover_everything():
    from autoscaler import replay
    for n in (0, 1, 7, 100, 1_000_003):
        for w in (1, 2, 3, 8):
            for align in (1, 4):
                parts = [replay.shard_range(n, r, w, align) for r in range(w)]
                assert parts[0][0] == 0 and parts[-1][1] == n
                assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
                assert all(lo % align == 0 for lo, _ in parts)
    go = np.array([0, 2, 3, 7, 8, 12], dtype=np.uint64)
    seen = []
    for r in range(2):
        r_lo, r_hi, q_lo, q_hi, local = replay.shard_groups(go, r, 2)
        assert local[0] == 0 and local[-1] == q_hi - q_lo
        seen.append((r_lo, r_hi, q_lo, q_hi))
    assert seen[0][1] == seen[1][0] and seen[0][3] == seen[1][2] and seen[1][3] == 12


def test_product_never_touches_the_oracle():
    """oracle/ is test infrastructure: no file of the product imports, includes, links or names
    it outside a comment (any file, no exceptions by name)."""
    pkg = os.path.join(ROOT, 'kiosk-autoscaler_b200')
    checked = 0
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if not f.endswith(('.py', '.cu', '.cuh', '.h', '.cpp')):
                continue
            checked += 1
            in_doc = False
            for ln, line in enumerate(open(os.path.join(dirpath, f), errors='replace'), 1):
                code = line.strip()
                if f.endswith('.py'):
                    if code.count('"""') % 2 == 1 or code.count("'''") % 2 == 1:
                        in_doc = not in_doc
                        continue
                    if in_doc or code.startswith('#') or '"""' in code:
                        continue
                    code = code.split('#')[0]
                else:
                    if code.startswith(('//', '*', '/*')):
                        continue
                    code = code.split('//')[0]
                assert 'oracle' not in code.lower(), '%s:%d: %s' % (os.path.join(dirpath, f), ln, line)
    assert checked >= 14
    for f in ('include/kas.h',):
        assert 'kas_oracle' not in open(os.path.join(ROOT, f)).read()


def test_reference_arm_prints_one_json_line():
    """bench.py --impl reference: one JSON line on stdout with the agreed keys."""
    import json
    import subprocess
    env = dict(os.environ, KAS_BENCH_CPU_KEYS='200000')
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference',
                          '--steps', '1', '--warmup', '0'], capture_output=True, text=True, env=env,
                         timeout=300)
    assert out.returncode == 0, out.stderr[-500:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d['impl'] == 'reference' and d['metric'] == 'keys matched/sec' and d['unit'] == 'keys/s'
    assert d['value'] > 0 and d['higher_is_better'] is True and d['vs_baseline'] is None
    assert d['cpu_baseline']['kind'] == 'port' and d['cpu_baseline']['cores'] >= 1
    assert d['e2e'] == {'value': d['value'], 'unit': 'keys/s', 'h2d_bytes_per_step': 0,
                        'd2h_bytes_per_step': 0}
    assert 'workload' in d['config']


def test_reference_arm_decisions_and_configs():
    """--impl reference for the other metric and a multi-queue config: honours --steps/--warmup,
    names its bounded sample in config, never loads the product library."""
    import json
    import subprocess
    env = dict(os.environ, KAS_BENCH_CPU_QUEUES='300', KAS_BENCH_REPLAY_T='200')
    for extra, metric in ((['--metric', 'decisions'], 'scaling decisions/sec'),
                          (['--config', 'c5', '--metric', 'decisions'], 'scaling decisions/sec'),
                          (['--config', 'c4'], 'scaling decisions/sec')):
        out = subprocess.run([sys.executable, '-X', 'importtime', os.path.join(ROOT, 'bench.py'), '--impl', 'reference',
                              '--steps', '3', '--warmup', '1'] + extra, capture_output=True, text=True, env=env,
                             timeout=300)
        assert out.returncode == 0, out.stderr[-500:]
        lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
        assert len(lines) == 1
        d = json.loads(lines[0])
        assert d['impl'] == 'reference' and d['metric'] == metric and d['steps'] == 3 and d['warmup'] == 1
        assert d['config']['cpu_sample_queues'] == 300 and 'bounded sample' in d['config']['cpu_sample']
        assert d['e2e']['value'] == d['value'] and d['gpu_launches'] == 0
        assert 'autoscaler.engine' not in out.stderr          # -X importtime lists every imported module
    sys.path.insert(0, ROOT)
    import bench
    go = bench.fleet_groups(5, 1000, 5000)
    assert go[0] == 0 and go[-1] == 5000 and (np.diff(go.astype(np.int64)) >= 1).all()
    assert (np.diff(go.astype(np.int64)) <= 8).all() and (bench.fleet_groups(5, 1000, 5000) == go).all()
    types = bench.resource_types(2, 0, 1000)
    assert set(types) == {'job', 'deployment'} and 0.4 < (types == 'job').mean() < 0.6


def test_short_quotient_magic_numbers_are_exact():
    """The policy-sweep kernel divides small tallies with ONE multiply-high:
    floor(t / k) == ((t + 1) * (ceil(2^32 / k) - 1)) >> 32 for every t < 2^17 and every
    1 <= k <= 2^14 (csrc/kas_decide.cuh, quotient_small).  Checked exhaustively; the first
    counterexamples sit right above that range (k = 2^14 fails at t = 2^18)."""
    t = np.arange(0, 1 << 17, dtype=np.uint64)
    for k in range(1, (1 << 14) + 1):
        m = np.uint64(((1 << 32) + k - 1) // k - 1)
        assert int(m) < (1 << 32)
        q = ((t + np.uint64(1)) * m) >> np.uint64(32)
        assert (q == t // np.uint64(k)).all(), k
        if k >= 2:
            # quotient_up: the round-up magic number m + 1 = ceil(2^32 / k) needs no + 1 on the
            # dividend and is exact on the same range (k = 1 has no 32-bit round-up number)
            assert int(m) + 1 < (1 << 32)
            assert (((t * (m + np.uint64(1))) >> np.uint64(32)) == t // np.uint64(k)).all(), k
    k = 1 << 14
    m = ((1 << 32) + k - 1) // k - 1
    assert (((1 << 18) + 1) * m) >> 32 != (1 << 18) // k


def test_startup_delay_limits_agree():
    """include/kas.h, the ctypes layer and the argument checks name the same limit."""
    from autoscaler import engine
    hdr = open(os.path.join(ROOT, 'include', 'kas.h')).read()
    m = re.search(r'#define\s+KAS_MAX_STARTUP_DELAY\s+(\d+)u', hdr)
    assert m and int(m.group(1)) == engine.MAX_STARTUP_DELAY
    assert 'startup_delay' in dict(engine.ReplayArgs._fields_)
    # optional fields are appended in the order they were added; older callers zero them
    assert [f[0] for f in engine.ReplayArgs._fields_[-3:]] == ['d_patch_fail', 'startup_delay', 'd_resume']
    # the struct the header declares has the same members in the same order
    body = hdr[hdr.index('typedef struct kas_replay_args {'):hdr.index('} kas_replay_args;')]
    body = re.sub(r'/\*.*?\*/', '', body, flags=re.S)
    members = re.findall(r'\b(\w+)\s*;', body)
    assert members == [f[0] for f in engine.ReplayArgs._fields_]
    with pytest.raises(ValueError):
        engine._check_delay(engine.MAX_STARTUP_DELAY + 1)
    with pytest.raises(ValueError):
        engine._check_delay(-1)
    assert engine._check_delay(0) == 0


def test_keep_factor_form_of_the_decision_is_the_reference_clip():
    """csrc/kas_decide.cuh, finish_simple3: for keys_per_pod >= 1, min <= max, max >= 0 and a
    non-negative tally,  max(min(q, max), min, cur * keep)  with
    keep = min(q + (min > 0), max > 0)  is get_desired_pods (autoscaler.py:197-219) -- the form
    the sweep kernels compute.  Exhaustive over a small box, any current pods (also negative)."""
    from oracle import oracle
    for kpp in (1, 2, 3):
        for tally in range(0, 31):
            q = tally // kpp
            for mx in range(0, 7):
                for mn in range(-2, mx + 1):
                    keep = min(q + (1 if mn > 0 else 0), 1 if mx > 0 else 0)
                    for cur in range(-2, 10):
                        want = oracle.get_desired_pods(tally, kpp, mn, mx, cur)
                        got = max(min(q, mx), mn, cur * keep)
                        if mx == 0 and cur > 0 and q > 0:
                            # the one place the three-term form needs keep = 0: max_pods == 0
                            assert keep == 0
                        assert got == want, (kpp, tally, mn, mx, cur)
