"""Scalar decision API (get_desired_pods / clip_pod_count) on the GPU against
the reference's known answers and the golden grid."""
import pytest

from tests import helpers

pytestmark = pytest.mark.gpu


def test_get_desired_pods_known_answers():
    # mirrors the reference's own test (autoscaler_test.py:87-102)
    import autoscaler
    scaler = autoscaler.Autoscaler(helpers.ListRedis(), 'queue')
    scaler.redis_keys['queue'] = 10
    assert scaler.get_desired_pods('queue', 2, 0, 2, 1) == 2      # desired > max
    assert scaler.get_desired_pods('queue', 5, 9, 10, 0) == 9     # desired < min
    assert scaler.get_desired_pods('queue', 3, 0, 5, 1) == 3      # in range
    assert scaler.get_desired_pods('queue', 10, 0, 5, 3) == 3     # in range, current pods exist
    with pytest.raises(ZeroDivisionError):
        scaler.get_desired_pods('queue', 0, 0, 10, 0)
    with pytest.raises(KeyError):
        scaler.get_desired_pods('nope', 1, 0, 10, 0)
    assert scaler.clip_pod_count(7, 9, 5, 0) == 5 and scaler.clip_pod_count(3, 9, 5, 0) == 9


def test_golden_decision_grid():
    from autoscaler import engine
    g = helpers.golden_json('decisions')
    rows = g['get_desired_pods']
    got = engine.decide([r[0] for r in rows], [r[1] for r in rows], [r[2] for r in rows],
                        [r[3] for r in rows], [r[4] for r in rows], mode=0)
    assert got == [r[5] for r in rows]
    rows = g['clip_pod_count']
    got = engine.decide([r[0] for r in rows], None, [r[1] for r in rows], [r[2] for r in rows],
                        [r[3] for r in rows], mode=1)
    assert got == [r[4] for r in rows]


def test_int64_corners():
    from autoscaler import engine
    from oracle import oracle
    big = 2 ** 63 - 1
    cases = [(-(2 ** 63), -1, 0, big, 0), (-(2 ** 63), -1, 0, 5, 0), (big, 1, -big, big, 3),
             (-(2 ** 63), 1, -(2 ** 63), big, 0), (-7, 2, -big, big, 0), (7, -2, -big, big, 0)]
    got = engine.decide(*[list(c) for c in zip(*cases)], mode=0)
    assert got == [oracle.get_desired_pods(*c) if c[:2] != (-(2 ** 63), -1) or c[3] != big else big
                   for c in cases]
    with pytest.raises(OverflowError):
        engine.decide([2 ** 64], [1], [0], [1], [0])


def test_randomised_decisions_full_int64_domain():
    """20 000 seeded tuples over edge values and the whole int64 range, against the
    Python port (whose arithmetic is Python's own unbounded ints, like the reference)."""
    import random
    from autoscaler import engine
    from oracle import oracle
    rng = random.Random(31337)
    edges = [0, 1, -1, 2, -2, 3, 7, 10, 2 ** 31 - 1, -(2 ** 31), 2 ** 31, 2 ** 32, 2 ** 62, -(2 ** 62),
             2 ** 63 - 1, -(2 ** 63) + 1]

    def pick():
        u = rng.random()
        if u < 0.35:
            return rng.choice(edges)
        if u < 0.7:
            return rng.randint(-60, 60)
        return rng.randint(-(2 ** 63) + 1, 2 ** 63 - 1)

    rows = []
    while len(rows) < 20000:
        r = (pick(), pick(), pick(), pick(), pick())
        if r[1] != 0:
            rows.append(r)
    got = engine.decide(*[list(c) for c in zip(*rows)], mode=0)
    want = [oracle.get_desired_pods(*r) for r in rows]
    bad = [(r, g, w) for r, g, w in zip(rows, got, want) if g != w]
    assert not bad, bad[:5]
    got = engine.decide([r[0] for r in rows], None, [r[2] for r in rows], [r[3] for r in rows],
                        [r[4] for r in rows], mode=1)
    assert got == [oracle.clip_pod_count(r[0], r[2], r[3], r[4]) for r in rows]
    # a zero divisor anywhere rejects the call, as `x // 0` raises in the reference
    with pytest.raises(ZeroDivisionError):
        engine.decide([5, 6], [1, 0], [0, 0], [9, 9], [0, 0], mode=0)
