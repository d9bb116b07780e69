"""Time the multi-queue (CSR groups) replay: KAS_LIB=<.so> python profiles/tools/time_grouped.py
Q queues in groups of 1..8 (BASELINE config 5), T ticks, default policy and a 4-policy set."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'kiosk-autoscaler_b200')]
import numpy as np  # noqa: E402
import torch  # noqa: E402
from autoscaler import engine  # noqa: E402

T = int(os.environ.get('KAS_BENCH_REPLAY_T', 2000))
Q = int(os.environ.get('KAS_BENCH_REPLAY_Q', 1_000_000))
rng = np.random.default_rng(5)
sizes = rng.integers(1, 9, size=Q)
off = np.concatenate([[0], np.cumsum(sizes)])
off = off[off <= Q]
if off[-1] != Q:
    off = np.append(off, Q)
R = len(off) - 1
d_go = torch.from_numpy(off.astype(np.int64)).cuda()
trace = engine.synth_trace(3, T, 0, Q)
if os.environ.get('KAS_TRACE_DTYPE', 'int32') == 'uint16':
    trace = trace.to(torch.int16)


def timed(pol, reps):
    for _ in range(2):
        out = engine.replay(trace, T, Q, pol, d_group_off=d_go, n_resources=R)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        out = engine.replay(trace, T, Q, pol, d_group_off=d_go, n_resources=R)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out


m1, o1 = timed([(1, 0, 1)], 5)
m4, o4 = timed([(1, 0, 1), (2, 0, 4), (5, 1, 16), (10, 0, 64)], 3)
import bench  # noqa: E402
m41, o41 = timed(bench.SWEEP, 2)
esz = trace.element_size()
print('grouped R=%d Q=%d T=%d %s: S=1 %.3f ms (%.0f GB/s)  S=4 %.3f ms (%.3e dec/s)  S=41 %.2f ms (%.3e dec/s)  chk=%d/%d' % (
    R, Q, T, os.environ.get('KAS_TRACE_DTYPE', 'int32'), m1, T * Q * esz / m1 / 1e6, m4, T * Q * 4 / m4 * 1e3,
    m41, T * Q * 41 / m41 * 1e3, int(o1['sums'][0, 0]), int(o41['sums'][:, 0].sum())), flush=True)
