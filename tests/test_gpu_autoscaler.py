"""The drop-in Autoscaler driven the way the reference's tests drive theirs
(autoscaler/autoscaler_test.py), with the tally and decisions on the GPU."""
import random

import pytest

from tests import helpers

pytestmark = pytest.mark.gpu


def _patch_kube(scaler, kube):
    scaler.get_apps_v1_client = lambda: kube
    scaler.get_batch_v1_client = lambda: kube


def test_tally_queues_llen_only():
    # shape of autoscaler_test.py:184-210: lists only, no processing-* keys
    import autoscaler
    expected = random.randint(1, 10)
    client = helpers.ListRedis(lists={'queue': expected}, keyspace=[b'queue'])
    scaler = autoscaler.Autoscaler(client, queues='queue', queue_delim=',')
    scaler.tally_queues()
    assert scaler.redis_keys == {'queue': expected}
    queues = 'predict,track,train'
    client = helpers.ListRedis(lists={q: expected for q in queues.split(',')},
                               keyspace=[q.encode() for q in queues.split(',')])
    scaler = autoscaler.Autoscaler(client, queues=queues, queue_delim=',')
    scaler.tally_queues()
    assert scaler.redis_keys == {q: expected for q in queues.split(',')}


def test_tally_queues_live_client_single_scan():
    import autoscaler
    from oracle import oracle
    keys = helpers.latin(helpers.golden_json('tally_cases')[0]['keyspace'])
    client = helpers.ListRedis(lists={'predict': 3}, keyspace=keys)
    scaler = autoscaler.Autoscaler(client, queues='predict,track,predict:zstack')
    scaler.tally_queues()
    want = oracle.tally_queues({q: 0 for q in scaler.redis_keys}, client.llen, keys)
    assert scaler.redis_keys == want
    assert client.scans == ['processing-*']          # one scan for all queues
    # a queue name with glob metacharacters is a live glob in the reference (autoscaler.py:69)
    g = autoscaler.Autoscaler(client, queues='pre*,tra?k,track')
    g.tally_queues()
    assert g.redis_keys == oracle.tally_queues({q: 0 for q in g.redis_keys}, client.llen, keys)


def test_scale_with_stateful_cluster_matches_fixtures():
    import autoscaler
    for case in helpers.golden_json('scale_cases')[:12]:
        names = case['queues'].split(case['delim'])
        client = helpers.ListRedis(keyspace=helpers.latin(case['keyspace']))
        kube = helpers.StatefulKube('r', case['cur0'])
        scaler = autoscaler.Autoscaler(client, case['queues'], case['delim'])
        _patch_kube(scaler, kube)
        after = []
        for row in case['depths']:
            client.lists = dict(zip(names, row))
            scaler.scale('ns', case['type'], 'r', min_pods=case['min'], max_pods=case['max'],
                         keys_per_pod=case['kpp'])
            after.append(int(kube.value))
        assert after == case['after'] and [int(p) for p in kube.patches] == case['patches']


def test_scale_does_not_raise_and_swallows_api_exception():
    # autoscaler_test.py:238-264
    import autoscaler
    scaler = autoscaler.Autoscaler(helpers.ListRedis(), queues='predict,track')
    _patch_kube(scaler, helpers.DummyKubernetes())
    for resource_type in scaler.managed_resource_types:
        scaler.scale(namespace='namespace', resource_type=resource_type, name='test')

    def bad_scale_resource(*args, **kwargs):
        raise helpers.ApiException('thrown on purpose')

    scaler.scale_resource = bad_scale_resource
    for resource_type in scaler.managed_resource_types:
        scaler.scale(namespace='namespace', resource_type=resource_type, name='test')


def test_live_scan_is_streamed_in_chunks():
    """tally_live: one pass over the client's scan_iter, keys staged in pinned chunks and counted
    chunk by chunk on the GPU (prefix kernel for literal names, glob kernel for live globs),
    against the oracle's per-queue scans of the same keyspace."""
    import autoscaler
    from autoscaler import engine, snapshot
    from oracle import coracle
    names = helpers.c2_queue_names(24) + ['pre*', 'tr[a-c]ck', 'x' * 60]
    _, blob, off = engine.synth_keys_host(31, 0, 30_000, helpers.c2_queue_names(24), long_tail=True,
                                          want_heads=False, want_blob=True)
    keys = [bytes(blob[off[i]:off[i + 1]]) for i in range(30_000)]
    keys += [b'processing-' + b'x' * 60 + b':1', b'processing-trbck:9']
    client = helpers.ListRedis(keyspace=keys)
    got = snapshot.tally_live(client, names, chunk_keys=4096, chunk_bytes=200_000)
    assert snapshot.tally_live.last['chunks'] >= 4 and snapshot.tally_live.last['keys'] == len(
        [k for k in keys if k.startswith(b'processing-')])
    live = [k for k in keys if k.startswith(b'processing-')]
    want = [sum(coracle.glob_match(b'processing-' + engine.encode_name(q) + b':*', k) for k in live) for q in names]
    assert got == want and sum(got[-3:]) > 0
    assert client.scans == ['processing-*']                      # ONE scan, not one per queue
    # str keys (decode_responses=True, as the reference's client is configured)
    client2 = helpers.ListRedis(keyspace=keys)
    client2.scan_iter = lambda match=None, count=None: (k.decode('latin-1') for k in live[:500])
    want2 = [sum(coracle.glob_match(b'processing-' + engine.encode_name(q) + b':*', k.decode('latin-1').encode('utf-8'))
                 for k in live[:500]) for q in names]
    assert snapshot.tally_live(client2, names) == want2
    with pytest.raises(ValueError):
        snapshot.tally_live(helpers.ListRedis(keyspace=[b'processing-' + b'y' * 300]), ['a'], chunk_bytes=128)
