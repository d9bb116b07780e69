"""BASELINE config 1 on the GPU: QUEUES='predict,track', 10k keys, 10k ticks, 41 policies,
from the committed fixture (tests/golden/c1.npz).  Prints the time of the tally and of the replay."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'kiosk-autoscaler_b200')]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import autoscaler  # noqa: E402
from autoscaler import engine  # noqa: E402
from tests import helpers  # noqa: E402

g = helpers.golden_c1()
snap = autoscaler.snapshot.KeySnapshot.from_keys(g['keys'])
scaler = autoscaler.Autoscaler(autoscaler.snapshot.SnapshotRedis(snap), 'predict,track')
scaler.tally_queues()
torch.cuda.synchronize()
t0 = time.perf_counter()
scaler.tally_queues()
torch.cuda.synchronize()
t_tally = time.perf_counter() - t0
llen = torch.from_numpy(g['llen']).cuda()
inprog = torch.from_numpy(g['inprog']).cuda()
go = torch.from_numpy(np.array([0, 2], dtype=np.int64)).cuda()
pol = [tuple(int(v) for v in p) for p in g['policies']]
for _ in range(2):
    out = engine.replay(llen, 10000, 2, pol, d_inprog=inprog, inprog_mode=2, d_group_off=go, n_resources=1)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    out = engine.replay(llen, 10000, 2, pol, d_inprog=inprog, inprog_mode=2, d_group_off=go, n_resources=1)
b.record()
torch.cuda.synchronize()
ok = out['sums'][:, 0].cpu().numpy().tolist() == g['pod_ticks'].tolist()
print('C1: tally_queues (host API, 10k keys) %.3f ms; replay 2 queues x 10k ticks x 41 policies %.3f ms '
      '(%.3e decisions/s); equals the reference run: %s' % (
          t_tally * 1e3, a.elapsed_time(b) / 5, 2 * 10000 * 41 / (a.elapsed_time(b) / 5) * 1e3, ok))
