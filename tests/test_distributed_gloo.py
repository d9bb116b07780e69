"""world_size-2 check of the sharding + all-reduce plumbing on CPU (gloo).

The kernels cannot run here, so each rank computes its shard with the oracle
(test infrastructure) and the host logic under test is what the GPU path uses
unchanged: shard_range / shard_groups and allreduce_sum_."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out_dir):
    for p in (ROOT, os.path.join(ROOT, 'kiosk-autoscaler_b200')):
        sys.path.insert(0, p)
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from autoscaler import engine, replay
    from oracle import coracle
    from tests import helpers
    names = helpers.c2_queue_names(16)
    W = engine.rec_bytes_for(names)
    n_keys, T, Q = 20_000, 50, 1003
    pol = [(1, 0, 1), (3, 1, 9)]
    # keys sharded by contiguous range
    lo, hi = replay.shard_range(n_keys, rank, world)
    heads, _, _ = engine.synth_keys_host(5, lo, hi - lo, names, rec_bytes=W)
    counts = torch.from_numpy(coracle.match_counts_heads(names, heads).astype(np.int64))
    replay.allreduce_sum_(counts)
    # resources sharded with their queues kept together
    rng = np.random.default_rng(3)
    llen, go = helpers.random_fleet(rng, T, Q, grouped=True)
    r_lo, r_hi, q_lo, q_hi, local_go = replay.shard_groups(go, rank, world)
    part = coracle.replay(llen[:, q_lo:q_hi].copy(), None, local_go, None, pol)
    sums = torch.from_numpy(np.stack([part['pod_ticks'].sum(0), part['n_patches'].sum(0),
                                      part['final'].astype(np.int64).sum(0)], axis=1))
    replay.allreduce_sum_(sums)
    if rank == 0:
        full_heads, _, _ = engine.synth_keys_host(5, 0, n_keys, names, rec_bytes=W)
        want_counts = coracle.match_counts_heads(names, full_heads).astype(np.int64)
        full = coracle.replay(llen, None, go, None, pol)
        want_sums = np.stack([full['pod_ticks'].sum(0), full['n_patches'].sum(0),
                              full['final'].astype(np.int64).sum(0)], axis=1)
        ok = (counts.numpy() == want_counts).all() and (sums.numpy() == want_sums).all()
        open(os.path.join(out_dir, 'ok'), 'w').write('1' if ok else '0')
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_and_allreduce(tmp_path):
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert open(tmp_path / 'ok').read() == '1'


def test_allreduce_is_a_noop_without_a_group():
    sys.path.insert(0, os.path.join(ROOT, 'kiosk-autoscaler_b200'))
    from autoscaler import replay
    t = torch.tensor([1, 2, 3])
    assert replay.allreduce_sum_(t) is t and t.tolist() == [1, 2, 3]
