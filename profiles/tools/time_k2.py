"""Time K2 (replay) for one build of libkas: KAS_LIB=<.so> python profiles/tools/time_k2.py
S=1 default policy and the 41-point sweep on a T x Q synthetic trace resident in HBM."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'kiosk-autoscaler_b200')]
import torch  # noqa: E402
import bench  # noqa: E402
from autoscaler import engine  # noqa: E402

T = int(os.environ.get('KAS_BENCH_REPLAY_T', 2000))
Q = int(os.environ.get('KAS_BENCH_REPLAY_Q', 1_000_000))
trace = engine.synth_trace(3, T, 0, Q)
if os.environ.get('KAS_TRACE_DTYPE', 'int32') == 'uint16':
    trace = trace.to(torch.int16)      # values < 800: same bits as uint16


def timed(pol, reps):
    for _ in range(2):
        out = engine.replay(trace, T, Q, pol)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        out = engine.replay(trace, T, Q, pol)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out


m1, o1 = timed([(1, 0, 1)], 10)
ms, os_ = timed(bench.SWEEP, 3)
print('%-28s ' % os.environ.get('KAS_TRACE_DTYPE', 'int32') + ' S=1 %.3f ms (%.0f GB/s of int32-equivalent)  sweep41 %.2f ms (%.3e dec/s)  chk=%d/%d' % (
    m1, T * Q * 4 / m1 / 1e6, ms, T * Q * 41 / ms * 1e3,
    int(o1['sums'][0, 0]), int(os_['sums'][:, 0].sum())), flush=True)
