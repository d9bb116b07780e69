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
