"""Pin the oracle (Python and C restatements) to fixtures produced by running
the reference's own source (oracle/make_golden.py) and to the reference test's
known answers (autoscaler_test.py:87-102)."""
import numpy as np
import pytest

from oracle import coracle, oracle
from tests import helpers


def test_known_answers_from_reference_test():
    # autoscaler_test.py:90-102 verbatim: redis_keys['queue'] = 10
    for args, want in [((2, 0, 2, 1), 2), ((5, 9, 10, 0), 9), ((3, 0, 5, 1), 3), ((10, 0, 5, 3), 3)]:
        assert oracle.get_desired_pods(10, *args) == want
        assert coracle.lib().kas_oracle_get_desired_pods(10, *args) == want
    g = helpers.golden_json('known_answers')
    assert [r['result'] for r in g['get_desired_pods']] == [2, 9, 3, 3]
    assert g['kpp_zero'] == 'ZeroDivisionError' and g['unknown_key'] == 'KeyError'
    with pytest.raises(ZeroDivisionError):
        oracle.get_desired_pods(10, 0, 0, 10, 0)
    for case in g['init']:
        assert list(oracle.parse_queues(case['queues'], case['delim'])) == case['keys']


def test_decision_grid():
    g = helpers.golden_json('decisions')
    L = coracle.lib()
    for tally, kpp, mn, mx, cur, want in g['get_desired_pods']:
        assert oracle.get_desired_pods(tally, kpp, mn, mx, cur) == want
        assert L.kas_oracle_get_desired_pods(tally, kpp, mn, mx, cur) == want
    for d, mn, mx, cur, want in g['clip_pod_count']:
        assert oracle.clip_pod_count(d, mn, mx, cur) == want
        assert L.kas_oracle_clip_pod_count(d, mn, mx, cur) == want
    # min > max keeps the if/elif order
    assert oracle.clip_pod_count(7, 9, 5, 0) == 5 and oracle.clip_pod_count(3, 9, 5, 0) == 9


def _groups_of(case):
    names = list(oracle.parse_queues(case['queues'], case['delim']))
    return names


def test_scale_cases():
    for case in helpers.golden_json('scale_cases'):
        names = _groups_of(case)
        keys = helpers.latin(case['keyspace'])
        inprog = oracle.match_counts(names, keys)
        # the fixture fed LLEN as dict(zip(split names, row)): duplicates collapse
        split = case['queues'].split(case['delim'])
        llen = [[dict(zip(split, row)).get(q, 0) for q in names] for row in case['depths']]
        tall = [[a + b for a, b in zip(row, inprog)] for row in llen]
        assert tall == case['tallies']
        fails = case.get('fails')
        after, patches = oracle.replay_resource(llen, [inprog] * len(llen), case['kpp'],
                                                case['min'], case['max'], case['cur0'], fails)
        assert after == case['after'] and patches == case['patches']
        res = coracle.replay(np.array(llen, dtype=np.int32), np.array(inprog, dtype=np.int32),
                             np.array([0, len(names)], dtype=np.uint64), [case['cur0']],
                             [(case['kpp'], case['min'], case['max'])], want_dense=True,
                             patch_fail=None if fails is None else np.array(fails).reshape(-1, 1))
        assert res['dense'][0, :, 0].tolist() == case['after']
        assert res['n_patches'][0, 0] == len(case['patches'])
        assert res['pod_ticks'][0, 0] == sum(case['after'])
        if fails is None:
            n = oracle.np_replay(llen, inprog, [0, len(names)], [case['cur0']],
                                 [(case['kpp'], case['min'], case['max'])], want_dense=True)
            assert n['dense'][0, :, 0].tolist() == case['after']


def test_only_running_cases():
    """The start-up-delay model of `only_running` (autoscaler.py:176-177): the reference's
    own scale() + get_current_pods(only_running=True) against a Kubernetes double whose
    available_replicas lag the spec; a job ignores the flag (autoscaler.py:185-190)."""
    cases = helpers.golden_json('only_running_cases')
    assert len(cases) >= 60 and any(c['delay'] >= 3 and c['fails'] for c in cases)
    for case in cases:
        delay = case['delay'] if case['type'] == 'deployment' else 0
        fails = case.get('fails')
        after, patches = oracle.replay_resource(case['depths'], None, case['kpp'], case['min'],
                                                case['max'], case['cur0'], fails, startup_delay=delay)
        assert after == case['after'] and patches == case['patches']
        nq = len(case['depths'][0])
        res = coracle.replay(np.array(case['depths'], dtype=np.int32), None,
                             np.array([0, nq], dtype=np.uint64), [case['cur0']],
                             [(case['kpp'], case['min'], case['max'])], want_dense=True,
                             patch_fail=None if fails is None else np.array(fails).reshape(-1, 1),
                             startup_delay=delay)
        assert res['dense'][0, :, 0].tolist() == case['after']
        assert res['n_patches'][0, 0] == len(case['patches'])
        assert res['pod_ticks'][0, 0] == sum(case['after'])
        assert res['final'][0, 0] == case['after'][-1]


def test_tally_cases():
    for case in helpers.golden_json('tally_cases'):
        names = list(oracle.parse_queues(case['queues'], case['delim']))
        keys = helpers.latin(case['keyspace'])
        got = oracle.tally_queues({q: 0 for q in names}, lambda q: case['lists'].get(q, 0), keys)
        assert got == case['redis_keys']
        blob, off = coracle.keys_to_blob(keys)
        c = coracle.match_counts(names, blob, off)
        assert [int(v) + case['lists'].get(q, 0) for q, v in zip(names, c)] == \
            [case['redis_keys'][q] for q in names]


def test_c1_against_reference_run():
    g = helpers.golden_c1()
    names = ['predict', 'track']
    blob, off = coracle.keys_to_blob(g['keys'])
    assert coracle.match_counts(names, blob, off).tolist() == g['snapshot_counts'].tolist()
    res = coracle.replay(g['llen'], g['inprog'], np.array([0, 2], dtype=np.uint64), [0],
                         g['policies'], want_dense=True)
    assert res['pod_ticks'][0].tolist() == g['pod_ticks'].tolist()
    assert res['n_patches'][0].tolist() == g['n_patches'].tolist()
    assert res['final'][0].tolist() == g['final'].tolist()
    for k, si in enumerate(g['dense_idx']):
        assert (res['dense'][si, :, 0] == g['dense'][k]).all()
    assert (g['llen'].astype(np.int64) + g['inprog'] == g['tallies_default']).all()
    # the Python port on a slice (it is slow): first 500 ticks, three policies
    for k, si in enumerate(g['dense_idx']):
        kpp, mn, mx = (int(v) for v in g['policies'][si])
        after, _ = oracle.replay_resource(g['llen'][:500].tolist(), g['inprog'][:500].tolist(),
                                          kpp, mn, mx, 0)
        assert after == g['dense'][k][:500].tolist()


def test_glob_c_equals_python():
    rng = np.random.default_rng(5)
    alpha = b'ab*?[]^-\\:c'
    for _ in range(4000):
        p = bytes(rng.choice(list(alpha), size=rng.integers(0, 7)).tolist())
        s = bytes(rng.choice(list(b'abc:-]'), size=rng.integers(0, 6)).tolist())
        assert coracle.glob_match(p, s) == oracle.redis_glob_match(p, s), (p, s)
    assert oracle.redis_glob_match(b'processing-a:*', b'processing-a:')
    assert not oracle.redis_glob_match(b'processing-a:*', b'processing-a')
    assert oracle.redis_glob_match(b'h[^e]llo', b'hallo') and not oracle.redis_glob_match(b'h[^e]llo', b'hello')
    assert oracle.redis_glob_match(b'h[a-b]llo', b'hbllo') and oracle.redis_glob_match(b'h\\*llo', b'h*llo')


def test_heads_equivalence_and_c_vs_numpy_replay():
    names = helpers.c2_queue_names(24)
    rng = np.random.default_rng(3)
    keys = []
    for _ in range(3000):
        q = names[rng.integers(len(names))]
        form = rng.integers(5)
        suffix = bytes(rng.integers(33, 126, size=rng.integers(0, 40)).tolist())
        keys.append([b'processing-' + q.encode() + b':' + suffix, b'processing-' + q.encode(),
                     b'processing-' + q.encode()[:-1] + b':' + suffix, q.encode() + b':' + suffix,
                     b'processing-' + q.encode() + b'x:' + suffix][form])
    blob, off = coracle.keys_to_blob(keys)
    W = oracle.rec_bytes_for(names)
    heads = oracle.pack_heads(keys, W)
    assert coracle.match_counts(names, blob, off).tolist() == \
        coracle.match_counts_heads(names, heads).tolist() == oracle.match_counts(names, keys)
    llen, go = helpers.random_fleet(rng, 60, 37, grouped=True, neg=True)
    pol = [(1, 0, 1), (3, 1, 9), (-2, 0, 5), (7, 5, 2)]
    a = coracle.replay(llen, llen[::-1].copy(), go, rng.integers(0, 6, len(go) - 1), pol, want_dense=True)
    b = oracle.np_replay(llen, llen[::-1].copy(), go, a['final'][:, 0] * 0 + 0, pol)  # cur0 differs: recompute
    c = oracle.np_replay(llen, llen[::-1].copy(), go, None, pol, want_dense=True)
    d = coracle.replay(llen, llen[::-1].copy(), go, None, pol, want_dense=True)
    assert (c['dense'] == d['dense']).all() and (c['pod_ticks'] == d['pod_ticks']).all()
    assert (c['n_patches'] == d['n_patches']).all()
    assert b['pod_ticks'].shape == a['pod_ticks'].shape


def test_python_and_c_oracle_agree_on_random_replays():
    """The two restatements (oracle.py line by line, kas_oracle.c for big sizes) against each
    other on random single resources: any keys_per_pod sign, min > max, negative counts,
    failing patches and the start-up delay of the `only_running` reading."""
    import random
    rnd = random.Random(7)
    for case in range(300):
        nq = rnd.choice([1, 1, 2, 5])
        T = rnd.randint(1, 60)
        kpp = rnd.choice([1, 1, 2, 3, 7, 50, -1, -3])
        mn = rnd.choice([0, 0, 1, 3, -2])
        mx = rnd.choice([0, 1, 4, 16, 2, -1])
        cur0 = rnd.choice([0, 0, 1, 5, 40])
        delay = rnd.choice([0, 0, 1, 2, 5, 9])
        lo = -5 if rnd.random() < 0.2 else 0
        depths = [[rnd.choice([0, 0, 0, rnd.randint(lo, 40)]) for _ in range(nq)] for _ in range(T)]
        fails = [rnd.random() < 0.2 for _ in range(T)] if rnd.random() < 0.4 else None
        after, patches = oracle.replay_resource(depths, None, kpp, mn, mx, cur0, fails, startup_delay=delay)
        res = coracle.replay(np.array(depths, dtype=np.int32), None, np.array([0, nq], dtype=np.uint64),
                             [cur0], [(kpp, mn, mx)], want_dense=True,
                             patch_fail=None if fails is None else np.array(fails).reshape(-1, 1),
                             startup_delay=delay)
        assert res['dense'][0, :, 0].tolist() == after, case
        assert res['n_patches'][0, 0] == len(patches), case
        assert res['pod_ticks'][0, 0] == sum(after), case


def test_first_pod_latency_against_the_reference_made_fixture():
    """tests/golden/first_pod_cases.json: the reference's own scale() loop (read through its
    get_current_pods, only_running under a start-up delay) with the reads before and after every
    tick recorded, and the written definition applied to them.  Both oracle forms must agree."""
    import numpy as np
    from oracle import coracle, oracle
    from tests import helpers
    cases = helpers.golden_json('first_pod_cases')
    assert len(cases) >= 80 and sum(c['stats']['n_unserved'] for c in cases) > 10
    for c in cases:
        want = c['stats']
        got = oracle.first_pod_latency(c['depths'], None, c['kpp'], c['min'], c['max'], c['cur0'],
                                       c['fails'], c['delay'])
        assert got == want
        depths = np.array(c['depths'], dtype=np.int32)
        nq = depths.shape[1]
        fails = None if c['fails'] is None else np.array(c['fails']).reshape(-1, 1)
        cg = coracle.first_pod(depths, None, np.array([0, nq], dtype=np.uint64), np.array([c['cur0']], dtype=np.int32),
                               (c['kpp'], c['min'], c['max']), patch_fail=fails, startup_delay=c['delay'])
        assert {k: int(v[0]) for k, v in cg.items()} == want
