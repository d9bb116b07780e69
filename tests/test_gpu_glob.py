"""General glob kernel (queue names with * ? [ \\) against the oracle's restatement
of Redis' matcher and against the reference-made fixtures."""
import numpy as np
import pytest

from tests import helpers

pytestmark = pytest.mark.gpu


def _gpu_glob_counts(keys, patterns):
    import torch
    from autoscaler import engine
    from oracle import coracle
    blob, off = coracle.keys_to_blob(keys)
    d_blob = torch.from_numpy(blob).cuda()
    d_off = torch.from_numpy(off.view(np.int64)).cuda()
    return engine.counts_to_host(engine.glob_count(d_blob, d_off, len(keys), patterns)).tolist()


def test_all_golden_tally_cases_including_glob_names():
    """Every fixture of tests/golden/tally_cases.json (made by running the reference's
    tally_queues), also the ones whose queue names are globs."""
    import autoscaler
    from oracle import oracle
    saw_glob = 0
    for case in helpers.golden_json('tally_cases'):
        names = list(oracle.parse_queues(case['queues'], case['delim']))
        snap = autoscaler.snapshot.KeySnapshot.from_keys(helpers.latin(case['keyspace']))
        scaler = autoscaler.Autoscaler(autoscaler.snapshot.SnapshotRedis(snap, case['lists']),
                                       case['queues'], case['delim'])
        scaler.tally_queues()
        assert scaler.redis_keys == case['redis_keys'], case['queues']
        saw_glob += any(oracle.has_glob_meta(q.encode()) for q in names)
    assert saw_glob >= 1


def test_random_patterns_vs_oracle():
    from oracle import coracle, oracle
    rng = np.random.default_rng(11)
    palpha = list(b'ab*?[]^-\\:c')
    salpha = list(b'abc:-]\\')
    pats = []
    for _ in range(400):
        pats.append(bytes(rng.choice(palpha, size=rng.integers(0, 9)).tolist()))
    pats += [b'processing-pre*:*', b'processing-tra?k:*', b'processing-[pt]r*:*', b'*', b'', b'**a**',
             b'[', b'[^', b'a[', b'\\', b'a\\', b'[a-', b'[z-a]', b'[\\]]', b'*:*:*', b'?*?']
    keys = [bytes(rng.choice(salpha, size=rng.integers(0, 10)).tolist()) for _ in range(3000)]
    keys += [b'', b'a', b'processing-predict:x', b'processing-track:', b'processing-trick:y',
             b'processing-pr:z', b'\\', b'a\\', b']']
    for lo in range(0, len(pats), 100):
        chunk = pats[lo:lo + 100]
        got = _gpu_glob_counts(keys, chunk)
        want = [sum(coracle.glob_match(p, k) for k in keys) for p in chunk]
        assert got == want, [(p, g, w) for p, g, w in zip(chunk, got, want) if g != w][:5]
    # the C matcher itself agrees with the readable Python one on these
    for p in pats[:60]:
        for k in keys[:200]:
            assert coracle.glob_match(p, k) == oracle.redis_glob_match(p, k)


def test_glob_equals_prefix_kernel_on_literal_names():
    """For literal names the glob 'processing-<q>:*' and the prefix kernel must agree."""
    import torch
    from autoscaler import engine, snapshot
    names = helpers.c2_queue_names(32)
    _, blob, off = engine.synth_keys_host(4, 0, 100_000, names, long_tail=True, want_heads=False,
                                          want_blob=True)
    snap = snapshot.KeySnapshot.from_blob(blob, off)
    want = snapshot.tally_snapshot(snap, names)
    d_blob, d_off = snap.device_blob()
    pats = [b'processing-' + q.encode() + b':*' for q in names]
    got = engine.counts_to_host(engine.glob_count(d_blob, d_off, len(snap), pats)).tolist()
    assert got == want
    # dump-indexed snapshots (one separator byte per key) go through the same kernel
    dump = b'\n'.join(bytes(blob[off[i]:off[i + 1]]) for i in range(2000))
    s2 = snapshot.KeySnapshot.from_dump(dump)
    assert snapshot.tally_snapshot(s2, ['pre*', 'predict']) == \
        snapshot.tally_snapshot(snapshot.KeySnapshot.from_blob(blob[:off[2000]], off[:2001]),
                                ['pre*', 'predict'])
    with pytest.raises(ValueError):
        snapshot.tally_snapshot(snapshot.KeySnapshot.from_heads(snap.heads(32)), ['pre*'])


def test_names_beyond_the_prefix_table_limits():
    """Names longer than a 64-byte record, a NUL inside a name, and more than 1024 names
    (also at a width where 1024 do not fit shared memory) still tally like the reference."""
    from autoscaler import snapshot
    from oracle import oracle
    long_name = 'q' * 70
    names = ['predict', long_name, 'a\x00b', 'track'] + ['n%04d' % i for i in range(1500)] + ['w' * 50]
    keys = [b'processing-predict:1', b'processing-' + long_name.encode() + b':x',
            b'processing-' + long_name.encode()[:-1] + b':x', b'processing-a\x00b:z', b'processing-a:z',
            b'processing-n0007:1', b'processing-n1499:', b'processing-n1499', b'processing-' + b'w' * 50 + b':k']
    keys = keys * 3
    snap = snapshot.KeySnapshot.from_keys(keys)
    got = snapshot.tally_snapshot(snap, names)
    assert got == oracle.match_counts(names, keys)
    assert sum(got) == 3 * 6
