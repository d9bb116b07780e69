"""K2 (replay) through the C ABI against the reference-made fixtures and the oracle."""
import zlib

import numpy as np
import pytest

from tests import helpers

pytestmark = pytest.mark.gpu


def stable_seed(*parts):
    """A seed that is the same in every process (hash() of a str is salted per run)."""
    return zlib.crc32(repr(parts).encode())


def test_golden_scale_cases():
    import autoscaler
    from oracle import oracle
    for case in helpers.golden_json('scale_cases'):
        names = list(oracle.parse_queues(case['queues'], case['delim']))
        split = case['queues'].split(case['delim'])
        llen = np.array([[dict(zip(split, row)).get(q, 0) for q in names] for row in case['depths']],
                        dtype=np.int32)
        snap = autoscaler.snapshot.KeySnapshot.from_keys(helpers.latin(case['keyspace']))
        inprog = autoscaler.snapshot.tally_snapshot(snap, names) if len(snap) else [0] * len(names)
        scaler = autoscaler.Autoscaler(helpers.ListRedis(), case['queues'], case['delim'])
        res = scaler.replay(llen, [(case['kpp'], case['min'], case['max'])],
                            inprog=np.array(inprog, dtype=np.int32), current_pods=case['cur0'],
                            resource_type=case['type'], want_dense=True, patch_fail=case.get('fails'))
        assert res.dense[0, :, 0].tolist() == case['after']
        assert int(res.n_patches[0, 0]) == len(case['patches'])
        field = 'replicas' if case['type'] == 'deployment' else 'parallelism'
        if not case.get('fails'):
            assert [b['spec'][field] for _, b in res.patch_bodies(case['type'])] == case['patches']
    with pytest.raises(ValueError):
        scaler.replay(llen, [(1, 0, 1)], resource_type='badvalue')
    with pytest.raises(ZeroDivisionError):
        scaler.replay(llen, [(0, 0, 1)])


def test_c1_reference_run():
    """BASELINE config 1: QUEUES='predict,track', 10k keys, 10k ticks, policy sweep."""
    import autoscaler
    g = helpers.golden_c1()
    snap = autoscaler.snapshot.KeySnapshot.from_keys(g['keys'])
    scaler = autoscaler.Autoscaler(autoscaler.snapshot.SnapshotRedis(snap), 'predict,track')
    scaler.tally_queues()
    assert [scaler.redis_keys['predict'], scaler.redis_keys['track']] == g['snapshot_counts'].tolist()
    res = scaler.replay(g['llen'], g['policies'], inprog=g['inprog'], want_dense=True)
    assert res.pod_ticks[0].tolist() == g['pod_ticks'].tolist()
    assert res.n_patches[0].tolist() == g['n_patches'].tolist()
    assert res.final[0].tolist() == g['final'].tolist()
    for k, si in enumerate(g['dense_idx']):
        assert (res.dense[si, :, 0] == g['dense'][k]).all()
    assert res.sums[:, 0].tolist() == g['pod_ticks'].tolist()
    rep = autoscaler.replay.sweep_report(res, interval_seconds=5, cost_per_pod_hour=2.0)
    assert [r['pod_ticks'] for r in rep] == g['pod_ticks'].tolist()
    assert [r['n_patches'] for r in rep] == g['n_patches'].tolist()
    assert rep[0]['pod_hours'] == g['pod_ticks'][0] * 5 / 3600.0 and rep[0]['cost'] == 2 * rep[0]['pod_hours']
    front = [r for r in rep if r['pareto']]
    assert front and all(not (o['pod_ticks'] < f['pod_ticks'] and o['n_patches'] < f['n_patches'])
                         for f in front for o in rep)


POLICY_SETS = {
    'one': [(1, 0, 1)],
    'sweep': [(k, a, b) for k in (1, 2, 5, 10, 50) for a in (0, 1) for b in (1, 4, 16, 64)],
    'quirks': [(3, 9, 5), (1, -2, 3), (2 ** 31 - 1, 0, 7), (1, 0, 2 ** 31 - 1), (7, 0, 0), (2, -5, -1)],
    'negmax': [(1, -4, -2), (3, -(2 ** 31), -1)],
    'edge': [(1, 0, 2 ** 31 - 1), (2, 0, 2 ** 31 - 1), (1, 2 ** 31 - 1, 2 ** 31 - 1), (3, -7, 0)],
    'negative': [(-1, 0, 5), (-3, -4, 6), (-(2 ** 31), 0, 3), (2, 0, 4)],
}


@pytest.mark.parametrize('pol', list(POLICY_SETS))
@pytest.mark.parametrize('shape', ['vec4', 'scalar', 'grouped', 'tiny', 'small_grouped', 'small_identity'])
@pytest.mark.parametrize('ipmode', [0, 1, 2])
def test_random_fleets_vs_oracle(pol, shape, ipmode):
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng(stable_seed(pol, shape, ipmode))
    T, Q = {'vec4': (37, 8192), 'scalar': (50, 4099), 'grouped': (41, 3000), 'tiny': (300, 3),
            'small_grouped': (700, 1500), 'small_identity': (130, 2048)}[shape]
    llen, go = helpers.random_fleet(rng, T, Q, grouped=shape in ('grouped', 'small_grouped'),
                                    neg=(pol == 'negative'))
    if pol == 'quirks':
        llen[rng.random(llen.shape) < 0.01] = 2 ** 31 - 1
    inprog = None
    if ipmode == 1:
        inprog = rng.integers(0, 5, size=Q).astype(np.int32)
    elif ipmode == 2:
        inprog = rng.integers(0, 9, size=(T, Q)).astype(np.int32)
        if pol == 'quirks':
            inprog[rng.random(inprog.shape) < 0.01] = 2 ** 31 - 1
    R = Q if go is None else len(go) - 1
    cur0 = rng.integers(0, 12, size=R).astype(np.int32) if rng.random() < 0.7 else None
    want = coracle.replay(llen, inprog, go, cur0, POLICY_SETS[pol], want_dense=True)
    got = autoscaler.replay.replay_fleet(llen, POLICY_SETS[pol], inprog=inprog, group_off=go,
                                         cur0=cur0, want_dense=True)
    assert (got.dense == want['dense']).all()
    assert (got.pod_ticks == want['pod_ticks']).all()
    assert (got.n_patches == want['n_patches']).all()
    assert (got.final == want['final']).all()
    assert (got.sums[:, 0] == want['pod_ticks'].sum(0)).all()
    assert (got.sums[:, 1] == want['n_patches'].sum(0)).all()
    assert (got.sums[:, 2] == want['final'].astype(np.int64).sum(0)).all()


RAGGED = [(1, 0, 1), (1, 0, 0), (1, 1, 1), (2, 0, 3), (2, -1, 3), (2, 2, 5), (3, 0, 0), (3, 0, 7), (3, 1, 7),
          (5, 0, 2 ** 24 - 1), (5, 2, 2 ** 24 - 1), (7, 0, 9), (7, 0, 9), (7, 0, 11), (9, 4, 4)]


@pytest.mark.parametrize('width', ['narrow', 'wide_policy', 'wide_cur0', 'two_launches'])
@pytest.mark.parametrize('dtype', ['int32', 'uint16'])
@pytest.mark.parametrize('ipmode', [0, 2])
def test_sweep_pairing_and_pod_tick_widths(width, dtype, ipmode):
    """The policy-sweep kernel pairs policies by (keys_per_pod, min > 0, max > 0), closes odd
    groups with an empty slot, and sums pod-ticks in 32 bits per tile when no decision can
    reach 2^24: groups of odd size, max_pods == 0, duplicates, decisions right at the bound,
    the 64-bit form (a large max_pods or given cur0), and more slots than one launch holds."""
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng(stable_seed(width, dtype, ipmode))
    T, Q = 150, 8192
    pols = list(RAGGED)
    if width == 'wide_policy':
        pols += [(5, 0, 2 ** 24), (11, 0, 2 ** 31 - 1), (11, 3, 2 ** 31 - 1)]
    if width == 'two_launches':
        pols += [(k, a, b) for k in (4, 6, 8, 10, 12, 13, 17) for a in (0, 2) for b in (0, 1, 40)][:-1]
    hi = 2 ** 16 - 1 if dtype == 'uint16' else 2 ** 31 - 1
    llen = rng.integers(0, 60, size=(T, Q)).astype(np.int64)
    llen[rng.random(llen.shape) < 0.3] = 0
    big = rng.random(Q) < 0.05                      # queues that sit at the largest count
    llen[:, big] = hi
    inprog = None
    if ipmode == 2:
        inprog = rng.integers(0, 9, size=(T, Q)).astype(np.int64)
        inprog[:, big] = hi if dtype == 'int32' else 0
    cur0 = None
    if width == 'wide_cur0':
        cur0 = rng.integers(0, 12, size=Q).astype(np.int32)
        cur0[rng.random(Q) < 0.02] = 2 ** 31 - 1
    np_t = np.uint16 if dtype == 'uint16' else np.int32
    want = coracle.replay(llen.astype(np.int32), None if inprog is None else inprog.astype(np.int32),
                          None, cur0, pols, want_dense=True)
    got = autoscaler.replay.replay_fleet(llen.astype(np_t), pols,
                                         inprog=None if inprog is None else inprog.astype(np_t),
                                         cur0=cur0, want_dense=True)
    assert (got.dense == want['dense']).all()
    assert (got.pod_ticks == want['pod_ticks']).all()
    assert (got.n_patches == want['n_patches']).all()
    assert (got.final == want['final']).all()
    assert (got.sums[:, 0] == want['pod_ticks'].sum(0)).all()
    assert (got.sums[:, 1] == want['n_patches'].sum(0)).all()
    if width == 'narrow':
        # the largest decision is one below the bound, held for whole tiles
        assert int(got.dense.max()) == 2 ** 24 - 1 or dtype == 'uint16'


PACKABLE = [(1, 0, 1), (1, 0, 0), (1, 1, 1), (2, 0, 3), (2, 0, 1023), (2, 2, 5), (3, 0, 0), (3, 0, 7), (3, 1, 7),
            (5, 0, 1023), (5, 1023, 1023), (7, 0, 9), (7, 0, 9), (7, 0, 11), (9, 4, 4), (16384, 0, 3), (1, 0, 1023)]


@pytest.mark.parametrize('data', ['small', 'sporadic_big', 'negatives', 'saturated'])
@pytest.mark.parametrize('dtype', ['int32', 'uint16'])
@pytest.mark.parametrize('ipmode', [0, 1, 2])
def test_sweep_packed_state(data, dtype, ipmode):
    """The sweep kernel's packed form (two resources per register, 16 bits each; policies with
    0 <= min, max <= 1023, keys_per_pod <= 2^14, no cur0): all-small counts, batches that leave
    the packed form because of a large or negative count and come back, and max_pods = 1023 held
    for whole tiles (the 16-bit pod-tick sum of a tile at 64 x 1023)."""
    if dtype == 'uint16' and data == 'negatives':
        pytest.skip('no negative uint16')
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng(stable_seed(data, dtype, ipmode))
    T, Q = 200, 8192
    hi = 2 ** 16 - 1 if dtype == 'uint16' else 2 ** 31 - 1
    llen = rng.integers(0, 60, size=(T, Q)).astype(np.int64)
    llen[rng.random(llen.shape) < 0.3] = 0
    llen[:, rng.random(Q) < 0.1] += 16000            # right below / above the packed limit of 2^14
    if data == 'sporadic_big':
        llen[rng.random(llen.shape) < 0.002] = hi
        llen[rng.random(llen.shape) < 0.002] = 16384
    if data == 'negatives':
        llen[rng.random(llen.shape) < 0.002] = -rng.integers(1, 50)
    if data == 'saturated':
        llen[:, rng.random(Q) < 0.5] = 16383
    inprog = None
    if ipmode == 1:
        inprog = rng.integers(0, 5, size=Q).astype(np.int64)
        inprog[rng.random(Q) < 0.01] = 70000
    elif ipmode == 2:
        inprog = rng.integers(0, 9, size=(T, Q)).astype(np.int64)
    np_t = np.uint16 if dtype == 'uint16' else np.int32
    llen = np.minimum(llen, hi)
    want = coracle.replay(llen.astype(np.int32), None if inprog is None else inprog.astype(np.int32),
                          None, None, PACKABLE, want_dense=True)
    ip_dev = None if inprog is None else (inprog.astype(np_t) if ipmode == 2 else inprog.astype(np.int32))
    got = autoscaler.replay.replay_fleet(llen.astype(np_t), PACKABLE, inprog=ip_dev, want_dense=True)
    assert (got.dense == want['dense']).all()
    assert (got.pod_ticks == want['pod_ticks']).all()
    assert (got.n_patches == want['n_patches']).all()
    assert (got.final == want['final']).all()
    assert (got.sums[:, 0] == want['pod_ticks'].sum(0)).all()
    assert (got.sums[:, 1] == want['n_patches'].sum(0)).all()
    # and without the dense output (the other instance of the tick code)
    got2 = autoscaler.replay.replay_fleet(llen.astype(np_t), PACKABLE, inprog=ip_dev)
    assert (got2.pod_ticks == want['pod_ticks']).all() and (got2.n_patches == want['n_patches']).all()
    assert (got2.final == want['final']).all()
    if data == 'saturated':
        assert int(want['dense'].max()) == 1023


@pytest.mark.parametrize('policy', [(1, 0, 1), (1, 0, 1023), (1, 2, 7), (1, 0, 0), (5, 0, 64), (16384, 0, 3),
                                    (3, 1, 1), (2, 1023, 1023)])
@pytest.mark.parametrize('T', [1, 7, 64, 301])
def test_uint16_single_policy_packed(policy, T):
    """One policy over a uint16 trace without in-progress term or cur0: the packed kernel
    (a trace word = two resources, never taken apart).  Counts up to 65535, max_pods = 1023
    held over the 64-tick emptying of the 16-bit sums, tick counts around the pipeline depth."""
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng(stable_seed(policy, T))
    Q = 8192
    llen = rng.integers(0, 60, size=(T, Q)).astype(np.uint16)
    llen[rng.random((T, Q)) < 0.4] = 0
    llen[rng.random((T, Q)) < 0.01] = 65535
    llen[:, rng.random(Q) < 0.2] = 40000           # sits at max_pods for the whole trace
    want = coracle.replay(llen.astype(np.int32), None, None, None, [policy], want_dense=True)
    got = autoscaler.replay.replay_fleet(llen, [policy], want_dense=True)
    assert (got.dense == want['dense']).all()
    assert (got.pod_ticks == want['pod_ticks']).all()
    assert (got.n_patches == want['n_patches']).all()
    assert (got.final == want['final']).all()
    assert (got.sums[0] == [want['pod_ticks'].sum(), want['n_patches'].sum(), want['final'].sum()]).all()
    got2 = autoscaler.replay.replay_fleet(llen, [policy])
    assert (got2.pod_ticks == want['pod_ticks']).all() and (got2.n_patches == want['n_patches']).all()
    assert (got2.final == want['final']).all()


def test_synth_trace_twin():
    from autoscaler import engine
    from oracle import coracle
    d = engine.synth_trace(9, 130, 1_000_000_007, 1000)
    assert (d.cpu().numpy() == coracle.trace_fill(9, 130, 1_000_000_007, 1000)).all()
    assert (d == 0).float().mean().item() > 0.3


def test_large_replay_sampled_queues():
    """1M queues x 2000 ticks resident in HBM, generated on the device; whole
    trajectories of sampled queues are rebuilt by the oracle's generator twin and
    replayed on the CPU; totals are also checked against a two-shard run."""
    import torch
    from autoscaler import engine
    from oracle import coracle
    T, Q, seed = 2000, 1_000_000, 21
    trace = engine.synth_trace(seed, T, 0, Q)
    pol = [(1, 0, 1), (5, 0, 16), (2, 1, 64), (50, 0, 4)]
    out = engine.replay(trace, T, Q, pol)
    tot = engine.totals_to_host(out['totals'])
    sums = out['sums'].cpu().numpy()
    rng = np.random.default_rng(1)
    for q in rng.integers(0, Q, size=24).tolist() + [0, Q - 1]:
        col = coracle.trace_fill(seed, T, q, 1)
        want = coracle.replay(col, None, None, None, pol)
        assert tot['pod_ticks'][q].tolist() == want['pod_ticks'][0].tolist()
        assert tot['n_patches'][q].tolist() == want['n_patches'][0].tolist()
        assert tot['final_current'][q].tolist() == want['final'][0].tolist()
    assert (sums[:, 0] == tot['pod_ticks'].sum(0)).all()
    assert (sums[:, 1] == tot['n_patches'].astype(np.int64).sum(0)).all()
    # patch-event stream at this size (245 scan tiles): offsets agree with the replay's own
    # patch counts, sampled logs with the oracle
    off, ev = engine.patch_events(trace, T, Q, pol[1])
    assert (np.diff(off.astype(np.int64)) == tot['n_patches'][:, 1].astype(np.int64)).all()
    assert len(ev) == int(sums[1, 1])
    for q in rng.integers(0, Q, size=8).tolist():
        col = coracle.trace_fill(seed, T, q, 1)
        dense = coracle.replay(col, None, None, None, [pol[1]], want_dense=True)['dense'][0, :, 0]
        mine = ev[int(off[q]):int(off[q + 1])]
        assert [(int(e['tick']), int(e['from_pods']), int(e['to_pods'])) for e in mine] == \
            _oracle_events(dense, 0)
    # sharding by resource: two halves generated independently give the same box-wide sums
    half = Q // 2
    acc = torch.zeros_like(out['sums'])
    for lo in (0, half):
        part = engine.synth_trace(seed, T, lo, half)
        acc += engine.replay(part, T, half, pol, want_totals=False)['sums']
    assert (acc.cpu().numpy() == sums).all()


def _oracle_events(dense_sr, cur0):
    """(tick, from, to) of one resource from its dense trajectory."""
    out, prev = [], int(cur0)
    for t, v in enumerate(dense_sr.tolist()):
        if v != prev:
            out.append((t, prev, v))
        prev = v
    return out


@pytest.mark.parametrize('shape', ['identity_small', 'identity_big', 'grouped'])
def test_patch_event_stream(shape):
    """f1: the calls scale_resource would have made, per resource in tick order."""
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng({'identity_small': 1, 'identity_big': 2, 'grouped': 3}[shape])
    T, Q = {'identity_small': (80, 700), 'identity_big': (25, 20_011), 'grouped': (60, 9000)}[shape]
    llen, go = helpers.random_fleet(rng, T, Q, grouped=(shape == 'grouped'), neg=(shape == 'grouped'))
    R = Q if go is None else len(go) - 1
    cur0 = rng.integers(0, 9, size=R).astype(np.int32)
    inprog = rng.integers(0, 3, size=Q).astype(np.int32)
    types = [('job' if i % 3 == 0 else 'deployment') for i in range(R)]
    for pol in [(1, 0, 1), (3, 1, 9), (-2, 0, 6), (2, 7, 4)]:
        want = coracle.replay(llen, inprog, go, cur0, [pol], want_dense=True)
        log = autoscaler.replay.patch_events(llen, pol, inprog=inprog, group_off=go, cur0=cur0,
                                             resource_types=types)
        assert len(log) == int(want['n_patches'].sum())
        assert (np.diff(log.off.astype(np.int64)) == want['n_patches'][:, 0]).all()
        for r in list(range(0, R, max(1, R // 50))) + [R - 1]:
            ev = log.of(r)
            assert [(int(e['tick']), int(e['from_pods']), int(e['to_pods'])) for e in ev] == \
                _oracle_events(want['dense'][0, :, r], cur0[r])
        field = 'parallelism' if types[0] == 'job' else 'replicas'
        assert all(set(b['spec']) == {field} for _, b in log.bodies(0))
    with pytest.raises(ZeroDivisionError):
        autoscaler.replay.patch_events(llen, (0, 0, 1))
    with pytest.raises(ValueError):
        autoscaler.replay.patch_events(llen, (1, 0, 1), resource_types='badvalue')


def test_patch_events_match_reference_fixture():
    """The patches the reference issued in tests/golden/scale_cases.json."""
    import autoscaler
    from oracle import oracle
    for case in helpers.golden_json('scale_cases'):
        names = list(oracle.parse_queues(case['queues'], case['delim']))
        split = case['queues'].split(case['delim'])
        llen = np.array([[dict(zip(split, row)).get(q, 0) for q in names] for row in case['depths']],
                        dtype=np.int32)
        inprog = np.array(oracle.match_counts(names, helpers.latin(case['keyspace'])), dtype=np.int32)
        fails = case.get('fails')
        log = autoscaler.replay.patch_events(llen, (case['kpp'], case['min'], case['max']), inprog=inprog,
                                             group_off=np.array([0, len(names)], dtype=np.uint64),
                                             cur0=np.array([case['cur0']], dtype=np.int32),
                                             resource_types=case['type'],
                                             patch_fail=None if fails is None else np.array(fails).reshape(-1, 1))
        field = 'replicas' if case['type'] == 'deployment' else 'parallelism'
        assert [b['spec'][field] for _, b in log.bodies(0)] == case['patches']


@pytest.mark.parametrize('shape', ['vec8', 'scalar', 'grouped', 'small'])
@pytest.mark.parametrize('pol', ['one', 'sweep', 'negative'])
@pytest.mark.parametrize('ipmode', [0, 1, 2])
def test_uint16_trace_format(shape, pol, ipmode):
    """The uint16 storage of the trace (half the bytes) gives the same integers."""
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng(stable_seed('u16', pol, shape, ipmode))
    T, Q = {'vec8': (45, 8192), 'scalar': (33, 4100), 'grouped': (29, 5000), 'small': (150, 40)}[shape]
    llen = rng.integers(0, 60, size=(T, Q)).astype(np.uint16)
    llen[rng.random((T, Q)) < 0.5] = 0
    llen[rng.random((T, Q)) < 0.01] = 65535
    go = None
    if shape == 'grouped':
        _, go = helpers.random_fleet(rng, 1, Q, grouped=True)
    inprog = None
    if ipmode == 1:
        inprog = rng.integers(-3 if pol == 'negative' else 0, 5, size=Q).astype(np.int32)
    elif ipmode == 2:
        inprog = rng.integers(0, 9, size=(T, Q)).astype(np.uint16)
    R = Q if go is None else len(go) - 1
    cur0 = rng.integers(0, 12, size=R).astype(np.int32)
    want = coracle.replay(llen.astype(np.int32), None if inprog is None else inprog.astype(np.int32), go,
                          cur0, POLICY_SETS[pol], want_dense=True)
    got = autoscaler.replay.replay_fleet(llen, POLICY_SETS[pol], inprog=inprog, group_off=go, cur0=cur0,
                                         want_dense=True)
    assert (got.dense == want['dense']).all()
    assert (got.pod_ticks == want['pod_ticks']).all() and (got.n_patches == want['n_patches']).all()
    assert (got.sums[:, 0] == want['pod_ticks'].sum(0)).all()
    log = autoscaler.replay.patch_events(llen, POLICY_SETS[pol][0], inprog=inprog, group_off=go, cur0=cur0)
    assert (np.diff(log.off.astype(np.int64)) == want['n_patches'][:, 0]).all()


@pytest.mark.parametrize('shape', ['identity', 'grouped', 'small'])
def test_failing_patch_calls(shape):
    """ApiException on a patch call is swallowed by scale() (autoscaler.py:267-273): the call
    counts, the replicas stay.  Fleet-sized check against the oracle (the reference-made
    fixtures with failures are covered by test_golden_scale_cases)."""
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng({'identity': 5, 'grouped': 6, 'small': 7}[shape])
    T, Q = {'identity': (60, 9000), 'grouped': (50, 12000), 'small': (200, 64)}[shape]
    llen, go = helpers.random_fleet(rng, T, Q, grouped=(shape != 'identity'))
    R = Q if go is None else len(go) - 1
    fail = rng.random((T, R)) < 0.3
    cur0 = rng.integers(0, 7, size=R).astype(np.int32)
    pol = [(1, 0, 3), (2, 1, 9), (-3, 0, 4)]
    want = coracle.replay(llen, None, go, cur0, pol, want_dense=True, patch_fail=fail)
    got = autoscaler.replay.replay_fleet(llen, pol, group_off=go, cur0=cur0, want_dense=True, patch_fail=fail)
    assert (got.dense == want['dense']).all()
    assert (got.n_patches == want['n_patches']).all() and (got.pod_ticks == want['pod_ticks']).all()
    log = autoscaler.replay.patch_events(llen, pol[1], group_off=go, cur0=cur0, patch_fail=fail)
    assert (np.diff(log.off.astype(np.int64)) == want['n_patches'][:, 1]).all()
    with pytest.raises(ValueError):
        autoscaler.replay.replay_fleet(llen, pol, group_off=go, patch_fail=fail[:-1])


def test_only_running_reference_fixture():
    """The start-up-delay replay against what the reference's own scale() +
    get_current_pods(only_running=True) did (tests/golden/only_running_cases.json)."""
    import autoscaler
    for case in helpers.golden_json('only_running_cases'):
        names = case['queues'].split(',')
        scaler = autoscaler.Autoscaler(None, queues=case['queues'])
        fails = case.get('fails')
        got = scaler.replay(np.array(case['depths'], dtype=np.int32),
                            [(case['kpp'], case['min'], case['max'])], current_pods=case['cur0'],
                            resource_type=case['type'], want_dense=True, patch_fail=fails,
                            startup_delay=case['delay'])
        assert got.dense[0, :, 0].tolist() == case['after']
        assert int(got.n_patches[0, 0]) == len(case['patches'])
        log = autoscaler.replay.patch_events(np.array(case['depths'], dtype=np.int32),
                                             (case['kpp'], case['min'], case['max']),
                                             group_off=np.array([0, len(names)], dtype=np.uint64),
                                             cur0=np.array([case['cur0']]), resource_types=case['type'],
                                             patch_fail=None if fails is None else np.array(fails).reshape(-1, 1),
                                             startup_delay=case['delay'])
        ev = log.of(0)
        assert [int(e['to_pods']) for e in ev] == case['patches']
        # what the controller read on the ticks it patched
        assert [int(e['from_pods']) for e in ev] == [case['seen'][int(e['tick'])] for e in ev]


@pytest.mark.parametrize('shape', ['identity', 'grouped', 'identity_u16'])
@pytest.mark.parametrize('delay', [1, 3, 16])
def test_startup_delay_fleets_vs_oracle(shape, delay):
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng(stable_seed(shape, delay))
    T, Q = 90, 6000
    llen, go = helpers.random_fleet(rng, T, Q, grouped=(shape == 'grouped'))
    R = Q if go is None else len(go) - 1
    cur0 = rng.integers(0, 9, size=R).astype(np.int32)
    fail = rng.random((T, R)) < 0.05
    pols = [(1, 0, 1), (2, 0, 8), (1, 1, 4), (5, 0, 64), (3, 9, 5), (-2, 0, 6)]
    want = coracle.replay(llen, None, go, cur0, pols, want_dense=True, patch_fail=fail, startup_delay=delay)
    trace = llen.astype(np.uint16) if shape == 'identity_u16' else llen
    if shape == 'identity_u16':
        want = coracle.replay(trace.astype(np.int32), None, go, cur0, pols, want_dense=True,
                              patch_fail=fail, startup_delay=delay)
    got = autoscaler.replay.replay_fleet(trace, pols, group_off=go, cur0=cur0, want_dense=True,
                                         patch_fail=fail, startup_delay=delay)
    assert (got.dense == want['dense']).all()
    assert (got.pod_ticks == want['pod_ticks']).all()
    assert (got.n_patches == want['n_patches']).all()
    assert (got.final == want['final']).all()
    assert (got.sums[:, 1] == want['n_patches'].sum(0)).all()
    # a job ignores the delay (autoscaler.py:185-190)
    job = autoscaler.replay.replay_fleet(trace, pols[:2], group_off=go, cur0=cur0, resource_type='job',
                                         startup_delay=delay)
    plain = autoscaler.replay.replay_fleet(trace, pols[:2], group_off=go, cur0=cur0)
    assert (job.pod_ticks == plain.pod_ticks).all() and (job.n_patches == plain.n_patches).all()
    log = autoscaler.replay.patch_events(trace, pols[1], group_off=go, cur0=cur0, patch_fail=fail,
                                         startup_delay=delay)
    assert (np.diff(log.off.astype(np.int64)) == want['n_patches'][:, 1]).all()


LEAN_POLICIES = {
    'k1': [(1, 0, 1)],
    'k1_min': [(1, 2, 9)],
    'kpp': [(3, 1, 9)],
    'max0': [(2, 0, 0)],
    'four': [(1, 0, 1), (2, 0, 4), (5, 1, 16), (10, 0, 64)],
    'edge': [(16384, 0, 3), (1, 0, 2 ** 20 - 1), (7, 1023, 1023)],
    'five': [(1, 0, 1), (2, 0, 4), (5, 1, 16), (10, 0, 64), (3, 0, 7)],
}


def _group_offsets(rng, Q, sizes):
    out, left = [0], Q
    while left > 0:
        s = int(min(left, rng.choice(sizes)))
        out.append(out[-1] + s)
        left -= s
    return np.array(out, dtype=np.uint64)


@pytest.mark.parametrize('pol', list(LEAN_POLICIES))
@pytest.mark.parametrize('groups', ['1to8', 'with_empty', 'wide', 'huge', 'uniform4'])
@pytest.mark.parametrize('dtype', ['int32', 'uint16'])
@pytest.mark.parametrize('ipmode', [0, 1, 2])
def test_grouped_lean_kernel(pol, groups, dtype, ipmode):
    """Multi-queue resources without cur0 / dense output / failing patches under realistic
    policies: the kernel with size-sorted threads and TMA-staged tiles.  Group sizes 1-8, empty
    groups, groups wider than a warp's size classes, one group too wide for a tile (that block
    reads element by element), ticks holding counts >= 2^16 or negative ones (replayed exactly),
    per-queue in-progress counts outside the 16-bit range (whole block exact)."""
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng(stable_seed('lean', pol, groups, dtype, ipmode))
    T, Q = 75, 6000
    sizes = {'1to8': list(range(1, 9)), 'with_empty': [0, 0, 1, 2, 3], 'wide': [1, 5, 9, 31, 32, 33, 40],
             'huge': [1, 2, 3, 4], 'uniform4': [4]}[groups]
    go = _group_offsets(rng, Q, sizes)
    if groups == 'huge':        # one resource owning 3500 queues in the middle
        keep = go[(go < 1200) | (go > 4700)]
        go = keep.astype(np.uint64)
    R = len(go) - 1
    hi = 2 ** 16 - 1 if dtype == 'uint16' else 2 ** 31 - 1
    llen = rng.integers(0, 60, size=(T, Q)).astype(np.int64)
    llen[rng.random(llen.shape) < 0.4] = 0
    llen[rng.random(llen.shape) < 0.002] = hi
    llen[rng.random(llen.shape) < 0.002] = 65535
    if dtype == 'int32':
        llen[rng.random(llen.shape) < 0.002] = 65536
        llen[rng.random(llen.shape) < 0.002] = -rng.integers(1, 50)
    np_t = np.uint16 if dtype == 'uint16' else np.int32
    inprog = ip_dev = None
    if ipmode == 1:
        inprog = rng.integers(0, 5, size=Q).astype(np.int64)
        inprog[rng.integers(0, Q, size=3)] = 70000           # those blocks leave the fast form
        inprog[rng.integers(0, Q, size=3)] = -2
        ip_dev = inprog.astype(np.int32)
    elif ipmode == 2:
        inprog = rng.integers(0, 9, size=(T, Q)).astype(np.int64)
        inprog[rng.random(inprog.shape) < 0.001] = hi
        ip_dev = inprog.astype(np_t)
    pols = LEAN_POLICIES[pol]
    want = coracle.replay(llen.astype(np.int32), None if inprog is None else inprog.astype(np.int32),
                          go, None, pols)
    got = autoscaler.replay.replay_fleet(llen.astype(np_t), pols, inprog=ip_dev, group_off=go)
    assert (got.pod_ticks == want['pod_ticks']).all()
    assert (got.n_patches == want['n_patches']).all()
    assert (got.final == want['final']).all()
    assert (got.sums[:, 0] == want['pod_ticks'].sum(0)).all()
    assert (got.sums[:, 1] == want['n_patches'].sum(0)).all()
    assert (got.sums[:, 2] == want['final'].astype(np.int64).sum(0)).all()
    assert R == got.pod_ticks.shape[0]


@pytest.mark.parametrize('pol', ['k1', 'kpp', 'four'])
@pytest.mark.parametrize('dtype', ['int32', 'uint16'])
@pytest.mark.parametrize('ipmode', [0, 1, 2])
def test_grouped_lean_kernel_large_fleet(pol, dtype, ipmode):
    """A fleet large enough for the 512-resource blocks of the lean multi-queue kernel (two tile
    buffers instead of three), including a block with a too-wide group and large counts."""
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng(stable_seed('lean-big', pol, dtype, ipmode))
    T, Q = 40, 420_000
    go = _group_offsets(rng, Q, list(range(1, 9)))
    go = go[(go < 100_000) | (go > 102_500)].astype(np.uint64)      # one resource of ~2500 queues
    assert len(go) - 1 >= 512 * 148
    hi = 2 ** 16 - 1 if dtype == 'uint16' else 2 ** 31 - 1
    llen = rng.integers(0, 60, size=(T, Q)).astype(np.int64)
    llen[rng.random(llen.shape) < 0.4] = 0
    llen[rng.random(llen.shape) < 0.0005] = hi
    if dtype == 'int32':
        llen[rng.random(llen.shape) < 0.0005] = -3
    np_t = np.uint16 if dtype == 'uint16' else np.int32
    inprog = ip_dev = None
    if ipmode == 1:
        inprog = rng.integers(0, 5, size=Q).astype(np.int64)
        inprog[rng.integers(0, Q, size=5)] = 70000
        ip_dev = inprog.astype(np.int32)
    elif ipmode == 2:
        inprog = rng.integers(0, 9, size=(T, Q)).astype(np.int64)
        ip_dev = inprog.astype(np_t)
    pols = LEAN_POLICIES[pol]
    want = coracle.replay(llen.astype(np.int32), None if inprog is None else inprog.astype(np.int32), go, None, pols)
    got = autoscaler.replay.replay_fleet(llen.astype(np_t), pols, inprog=ip_dev, group_off=go)
    assert (got.pod_ticks == want['pod_ticks']).all()
    assert (got.n_patches == want['n_patches']).all()
    assert (got.final == want['final']).all()
    assert (got.sums[:, 0] == want['pod_ticks'].sum(0)).all()


def test_grouped_offsets_that_overlap_or_go_back():
    """CSR offsets straight through the C ABI that are not non-decreasing: the lean kernel's
    blocks notice and read element by element, like the plain kernel and the oracle (a range
    that goes back is empty, overlapping ranges are each summed)."""
    import torch
    from autoscaler import engine
    from oracle import coracle
    rng = np.random.default_rng(77)
    T, Q, R = 40, 4096, 700
    llen = rng.integers(0, 30, size=(T, Q)).astype(np.int32)
    go = np.sort(rng.integers(0, Q + 1, size=R + 1)).astype(np.uint64)
    go[0], go[-1] = 0, Q
    go[100], go[101] = go[101], go[100]          # one range goes back, its neighbours overlap
    go[400] = go[398]
    pols = [(1, 0, 1), (2, 1, 8)]
    want = coracle.replay(llen, None, go, None, pols)
    d_go = torch.from_numpy(go.view(np.int64)).cuda()
    out = engine.replay(torch.from_numpy(llen).cuda(), T, Q, pols, d_group_off=d_go, n_resources=R)
    tot = engine.totals_to_host(out['totals'])
    assert (tot['pod_ticks'] == want['pod_ticks']).all()
    assert (tot['n_patches'] == want['n_patches']).all()
    assert (tot['final_current'] == want['final']).all()


def test_first_pod_latency_reference_fixture():
    """f4: time to the first pod after a burst, against what the reference's own scale() loop
    did (tests/golden/first_pod_cases.json)."""
    import autoscaler
    for c in helpers.golden_json('first_pod_cases'):
        depths = np.array(c['depths'], dtype=np.int32)
        nq = depths.shape[1]
        fails = None if c['fails'] is None else np.array(c['fails']).reshape(-1, 1)
        got = autoscaler.replay.first_pod_latency(
            depths, [(c['kpp'], c['min'], c['max'])], group_off=np.array([0, nq], dtype=np.uint64),
            cur0=np.array([c['cur0']]), resource_type=c['type'], patch_fail=fails,
            startup_delay=c['delay'], want_rows=True)[0]
        row = got['rows'][0]
        assert {k: int(row[k]) for k in ('n_served', 'n_unserved', 'wait_sum', 'wait_max', 'pending')} == c['stats']
        assert got['n_served'] == c['stats']['n_served'] and got['wait_max'] == c['stats']['wait_max']


@pytest.mark.parametrize('shape', ['identity', 'grouped', 'identity_u16'])
@pytest.mark.parametrize('delay', [0, 2, 16])
@pytest.mark.parametrize('ipmode', [0, 1, 2])
def test_first_pod_latency_fleets_vs_oracle(shape, delay, ipmode):
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng(stable_seed('ttfp', shape, delay, ipmode))
    T, Q = 120, 5000
    llen, go = helpers.random_fleet(rng, T, Q, grouped=(shape == 'grouped'), neg=(shape == 'grouped'))
    llen[:, rng.random(Q) < 0.5] //= 8                     # quiet queues: bursts that need several ticks
    R = Q if go is None else len(go) - 1
    cur0 = rng.integers(0, 3, size=R).astype(np.int32) * (rng.random(R) < 0.3)
    fail = rng.random((T, R)) < 0.1
    np_t = np.uint16 if shape == 'identity_u16' else np.int32
    inprog = ip_dev = None
    if ipmode == 1:
        inprog = (rng.integers(0, 3, size=Q) * (rng.random(Q) < 0.2)).astype(np.int32)
        ip_dev = inprog
    elif ipmode == 2:
        inprog = (rng.integers(0, 4, size=(T, Q)) * (rng.random((T, Q)) < 0.1)).astype(np.int32)
        ip_dev = inprog.astype(np_t)
    pols = [(1, 0, 1), (6, 0, 8), (3, 1, 4), (-2, 0, 5), (40, 0, 2)]
    got = autoscaler.replay.first_pod_latency(llen.astype(np_t), pols, inprog=ip_dev, group_off=go,
                                              cur0=cur0.astype(np.int32), patch_fail=fail, startup_delay=delay,
                                              want_rows=True)
    seen_wait = 0
    for pol, g in zip(pols, got):
        want = coracle.first_pod(llen, inprog, go, cur0.astype(np.int32), pol, patch_fail=fail, startup_delay=delay)
        for k in ('n_served', 'n_unserved', 'wait_sum', 'wait_max', 'pending'):
            assert (g['rows'][k].astype(np.int64) == want[k]).all(), (pol, k)
        assert g['n_served'] == int(want['n_served'].sum()) and g['n_unserved'] == int(want['n_unserved'].sum())
        assert g['wait_sum'] == int(want['wait_sum'].sum()) and g['wait_max'] == int(want['wait_max'].max())
        assert g['pending_ticks'] == int(want['pending'].sum())
        seen_wait += g['wait_sum']
    assert seen_wait > 0
    # the report carries the columns
    res = autoscaler.replay.replay_fleet(llen.astype(np_t), pols, inprog=ip_dev, group_off=go,
                                         cur0=cur0.astype(np.int32), patch_fail=fail, startup_delay=delay)
    rep = autoscaler.replay.sweep_report(res, first_pod=got)
    assert rep[1]['bursts_served'] == got[1]['n_served'] and rep[1]['mean_wait_seconds'] == got[1]['mean_wait'] * 5


@pytest.mark.parametrize('R', [1, 3, 50])
@pytest.mark.parametrize('T', [1024, 5000, 10007])
@pytest.mark.parametrize('dtype', ['int32', 'uint16'])
@pytest.mark.parametrize('ipmode', [0, 1, 2])
def test_tick_parallel_replay_of_single_queue_resources(R, T, dtype, ipmode):
    """Few single-queue resources with a long trace: the ticks are replayed in parallel (each
    tick is a clamp of the current pods; clamps compose).  Against the sequential oracle: sane
    policies incl. min_pods > 0 and max_pods = 0, negative and huge counts, negative and large
    cur0, dense output, thread runs that end exactly at T."""
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng(stable_seed('scan', R, T, dtype, ipmode))
    hi = 2 ** 16 - 1 if dtype == 'uint16' else 2 ** 31 - 1
    llen = rng.integers(0, 30, size=(T, R)).astype(np.int64)
    llen[rng.random((T, R)) < 0.6] = 0
    llen[rng.random((T, R)) < 0.01] = hi
    if dtype == 'int32':
        llen[rng.random((T, R)) < 0.02] = -rng.integers(1, 2 ** 31)
    np_t = np.uint16 if dtype == 'uint16' else np.int32
    inprog = ip_dev = None
    if ipmode == 1:
        inprog = rng.integers(-2, 4, size=R).astype(np.int64)
        ip_dev = inprog.astype(np.int32)
    elif ipmode == 2:
        inprog = (rng.integers(0, 5, size=(T, R)) * (rng.random((T, R)) < 0.2)).astype(np.int64)
        ip_dev = inprog.astype(np_t)
    cur0 = rng.integers(-3, 40, size=R).astype(np.int32)
    cur0[0] = 2 ** 31 - 1
    pols = [(1, 0, 1), (3, 0, 8), (2, 1, 5), (7, 0, 0), (1, -4, 6), (16384, 2, 2), (5, 0, 2 ** 31 - 1)]
    want = coracle.replay(llen.astype(np.int32), None if inprog is None else inprog.astype(np.int32), None,
                          cur0, pols, want_dense=True)
    got = autoscaler.replay.replay_fleet(llen.astype(np_t), pols, inprog=ip_dev, cur0=cur0, want_dense=True)
    assert (got.dense == want['dense']).all()
    assert (got.pod_ticks == want['pod_ticks']).all()
    assert (got.n_patches == want['n_patches']).all()
    assert (got.final == want['final']).all()
    assert (got.sums[:, 0] == want['pod_ticks'].sum(0)).all() and (got.sums[:, 1] == want['n_patches'].sum(0)).all()
    assert (got.sums[:, 2] == want['final'].astype(np.int64).sum(0)).all()
    got2 = autoscaler.replay.replay_fleet(llen.astype(np_t), pols[:2], inprog=ip_dev)      # no cur0, no dense
    want2 = coracle.replay(llen.astype(np.int32), None if inprog is None else inprog.astype(np.int32), None,
                           None, pols[:2])
    assert (got2.pod_ticks == want2['pod_ticks']).all() and (got2.n_patches == want2['n_patches']).all()
    # one resource through the Autoscaler front (a single-queue scaler)
    scaler = autoscaler.Autoscaler(None, 'predict')
    one = scaler.replay(llen[:, :1].astype(np_t), pols[:3], current_pods=int(cur0[0]))
    w1 = coracle.replay(llen[:, :1].astype(np.int32), None, None, cur0[:1], pols[:3])
    assert (one.pod_ticks == w1['pod_ticks']).all() and (one.final == w1['final']).all()


@pytest.mark.parametrize('shape', ['two_queues', 'mixed_groups', 'single_queue_failing'])
@pytest.mark.parametrize('dtype', ['int32', 'uint16'])
@pytest.mark.parametrize('ipmode', [0, 1, 2])
def test_small_fleet_long_trace_by_state_enumeration(shape, dtype, ipmode):
    """A few resources with a long trace and small max_pods: every possible pod count is pushed
    through runs of ticks in parallel and the runs are chained (kas_replay_scan.cuh).  Multi-queue
    resources, failing patch calls, a starting count outside [0, max_pods], and blocks that do
    not fit the form (min_pods < 0, max_pods > 254, nine queues) against the sequential oracle."""
    import autoscaler
    from oracle import coracle
    rng = np.random.default_rng(stable_seed('fsm', shape, dtype, ipmode))
    T = 3001
    sizes = {'two_queues': [2], 'mixed_groups': [1, 3, 8, 9, 2], 'single_queue_failing': [1, 1, 1]}[shape]
    go = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
    Q, R = int(go[-1]), len(sizes)
    hi = 2 ** 16 - 1 if dtype == 'uint16' else 2 ** 31 - 1
    llen = rng.integers(0, 12, size=(T, Q)).astype(np.int64)
    llen[rng.random((T, Q)) < 0.7] = 0
    llen[rng.random((T, Q)) < 0.003] = hi
    if dtype == 'int32':
        llen[rng.random((T, Q)) < 0.01] = -rng.integers(1, 1000)
    np_t = np.uint16 if dtype == 'uint16' else np.int32
    inprog = ip_dev = None
    if ipmode == 1:
        inprog = rng.integers(0, 3, size=Q).astype(np.int64)
        ip_dev = inprog.astype(np.int32)
    elif ipmode == 2:
        inprog = (rng.integers(0, 5, size=(T, Q)) * (rng.random((T, Q)) < 0.1)).astype(np.int64)
        ip_dev = inprog.astype(np_t)
    cur0 = rng.integers(0, 5, size=R).astype(np.int32)
    cur0[0] = 2 ** 31 - 1
    if R > 1:
        cur0[1] = -7
    fail = rng.random((T, R)) < 0.2
    pols = [(1, 0, 1), (2, 0, 4), (1, 1, 16), (5, 0, 64), (1, 0, 254), (1, 0, 255), (3, -2, 6), (-2, 0, 5),
            (7, 0, 0), (1, 3, 3), (2, 0, 1000)]
    want = coracle.replay(llen.astype(np.int32), None if inprog is None else inprog.astype(np.int32), go, cur0, pols,
                          want_dense=True, patch_fail=fail)
    got = autoscaler.replay.replay_fleet(llen.astype(np_t), pols, inprog=ip_dev, group_off=go, cur0=cur0,
                                         want_dense=True, patch_fail=fail)
    assert (got.dense == want['dense']).all()
    assert (got.pod_ticks == want['pod_ticks']).all()
    assert (got.n_patches == want['n_patches']).all()
    assert (got.final == want['final']).all()
    assert (got.sums[:, 0] == want['pod_ticks'].sum(0)).all() and (got.sums[:, 1] == want['n_patches'].sum(0)).all()
    got2 = autoscaler.replay.replay_fleet(llen.astype(np_t), pols[:4], inprog=ip_dev, group_off=go)
    want2 = coracle.replay(llen.astype(np.int32), None if inprog is None else inprog.astype(np.int32), go, None, pols[:4])
    assert (got2.pod_ticks == want2['pod_ticks']).all() and (got2.final == want2['final']).all()


@pytest.mark.parametrize('shape', ['identity', 'grouped', 'small'])
@pytest.mark.parametrize('dtype', ['int32', 'uint16'])
@pytest.mark.parametrize('ipmode', [0, 1, 2])
def test_chunked_replay_resumes_bit_for_bit(shape, dtype, ipmode):
    """A trace replayed in chunks of ticks (each chunk resumes from the totals of the one before,
    kas_replay_args.d_resume + kas_totals_chain) gives the integers of one replay of the whole
    trace and of the oracle: streamed from host memory, several policies, cur0 on the first chunk,
    chunk sizes that do not divide T; and the patch log of a resumed chunk continues the log."""
    import torch
    import autoscaler
    from autoscaler import engine
    from oracle import coracle
    rng = np.random.default_rng(stable_seed('resume', shape, dtype, ipmode))
    T, Q = {'identity': (333, 8192), 'grouped': (257, 6000), 'small': (1500, 24)}[shape]
    llen, go = helpers.random_fleet(rng, T, Q, grouped=(shape != 'identity'))
    R = Q if go is None else len(go) - 1
    np_t = np.uint16 if dtype == 'uint16' else np.int32
    inprog = ip_dev = None
    if ipmode == 1:
        inprog = rng.integers(0, 4, size=Q).astype(np.int32)
        ip_dev = inprog
    elif ipmode == 2:
        inprog = (rng.integers(0, 6, size=(T, Q)) * (rng.random((T, Q)) < 0.2)).astype(np.int32)
        ip_dev = inprog.astype(np_t)
    cur0 = rng.integers(0, 9, size=R).astype(np.int32)
    pols = [(1, 0, 1), (3, 1, 9), (5, 0, 64), (2, 0, 4), (-2, 0, 6), (1, 7, 3)]
    want = coracle.replay(llen, inprog, go, cur0, pols)
    for chunk in (64, 100, T):
        sr = autoscaler.replay.StreamingReplay(pols, Q, group_off=go, chunk_ticks=chunk, trace_dtype=dtype,
                                               per_tick_inprog=(ipmode == 2))
        host = torch.from_numpy(llen.astype(np_t).view(np.int16) if dtype == 'uint16' else llen).pin_memory()
        got = sr.run(host, inprog=ip_dev, cur0=cur0)
        assert sr.chunks == -(-T // chunk)
        assert (got.pod_ticks == want['pod_ticks']).all()
        assert (got.n_patches == want['n_patches']).all()
        assert (got.final == want['final']).all()
        assert (got.sums[:, 0] == want['pod_ticks'].sum(0)).all() and (got.sums[:, 1] == want['n_patches'].sum(0)).all()
        assert (got.sums[:, 2] == want['final'].astype(np.int64).sum(0)).all()
    # the patch log of the second half, resumed from the totals of the first
    half = T // 2
    d = torch.from_numpy(llen.astype(np_t).view(np.int16) if dtype == 'uint16' else llen).cuda()
    d_go = None if go is None else torch.from_numpy(go.view(np.int64)).cuda()
    d_ip = None if ip_dev is None else torch.from_numpy(ip_dev.view(np.int16) if ip_dev.dtype == np.uint16 else ip_dev).cuda()
    first = engine.replay(d[:half], half, Q, pols, d_inprog=None if d_ip is None else (d_ip[:half] if ipmode == 2 else d_ip),
                          inprog_mode=ipmode, d_group_off=d_go, n_resources=R, d_cur0=torch.from_numpy(cur0).cuda())
    R_ = R
    parr = (engine.Policy * len(pols))(*[engine.Policy(*p) for p in pols])
    args = engine.ReplayArgs(engine.ptr(d[half:]), engine.ptr(None if d_ip is None else (d_ip[half:] if ipmode == 2 else d_ip)),
                             ipmode, T - half, Q, R_, engine.ptr(d_go), None, parr, len(pols), None, None, None,
                             engine.trace_dtype_of(d), None, 0, engine.ptr(first['totals']))
    off = torch.empty(R_ + 1, dtype=torch.int64, device='cuda')
    scratch = torch.empty(max(8, engine.lib().kas_patch_events_scratch_bytes(R_)), dtype=torch.uint8, device='cuda')
    engine.check(engine.lib().kas_patch_event_offsets(engine.C.byref(args), 1, engine.ptr(off), engine.ptr(scratch),
                                                      scratch.numel(), engine.stream_ptr()), 'kas_patch_event_offsets')
    whole = coracle.replay(llen, inprog, go, cur0, [pols[1]], want_dense=True)['dense'][0]
    before = whole[half - 1]
    changes = ((np.diff(np.concatenate([before[None], whole[half:]]), axis=0) != 0).sum(0))
    assert (np.diff(off.cpu().numpy()) == changes).all()
    # arguments that cannot be combined
    bad = engine.ReplayArgs(engine.ptr(d), None, 0, T, Q, R_, engine.ptr(d_go), engine.ptr(torch.from_numpy(cur0).cuda()),
                            parr, len(pols), engine.ptr(first['totals']), None, None, engine.trace_dtype_of(d), None, 0,
                            engine.ptr(first['totals']))
    assert engine.lib().kas_replay(engine.C.byref(bad), engine.stream_ptr()) == engine.KAS_ERR_INVALID
