"""Time K1 (match_count) for one build of libkas: KAS_LIB=<.so> python profiles/tools/time_k1.py
Prints one line: library, ms per launch (CUDA events, 30 launches after 5 warm-ups), GB/s."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'kiosk-autoscaler_b200')]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import bench  # noqa: E402
from autoscaler import engine  # noqa: E402

n = int(os.environ.get('KAS_BENCH_KEYS', 100_000_000))
names = {'c2': bench.queue_names(), 'two': ['predict', 'track'], 'one': ['predict'],
         'w16': ['q%d' % i for i in range(8)],
         'n1024': ['q%04d%s' % (i, 'abcdefghij'[:i % 11]) for i in range(1024)]}[os.environ.get('KAS_NAMES', 'c2')]
W = engine.rec_bytes_for(names)
host = torch.empty(n * W, dtype=torch.uint8).pin_memory()
engine.synth_keys_host(2, 0, n, names, rec_bytes=W, heads_out=host)
d = host.cuda()
tbl = torch.from_numpy(engine.build_prefix_table(names, W)).cuda()
c = torch.empty(len(names), dtype=torch.int64, device='cuda')
for _ in range(5):
    engine.match_count(d, n, W, tbl, len(names), d_counts=c)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(30):
    engine.match_count(d, n, W, tbl, len(names), d_counts=c)
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / 30
print('%-40s %.4f ms  %.0f GB/s  sum=%d' % (os.path.basename(engine.LIB_PATH), ms, n * W / ms / 1e6,
                                           int(engine.counts_to_host(c).sum())), flush=True)
