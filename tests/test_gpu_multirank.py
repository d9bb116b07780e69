"""The CUDA path on several GPUs of one box: one process per GPU, keys and resources split by
contiguous range (shard_range / shard_groups), ONE all-reduce of the integer totals -- through
libkas's own communicator (kas_comm_* / kas_allreduce_i64, NCCL over NVLink) and, for comparison,
through torch.distributed.  The totals must be the integers a single rank computes and the
integers the CPU oracle computes (reference: autoscaler.py:60-77 for the counts, :244-260 for the
replay), at every world size.  Skipped below 2 visible GPUs."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

N_KEYS, T, Q = 3_000_017, 300, 40_000
POLICIES = [(1, 0, 1), (3, 1, 9), (5, 0, 16), (2, 0, 4), (10, 1, 64)]


def _worker(rank, world, port, out_dir):
    for p in (ROOT, os.path.join(ROOT, 'kiosk-autoscaler_b200')):
        sys.path.insert(0, p)
    import torch
    import torch.distributed as dist
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    from autoscaler import engine, replay, snapshot
    from tests import helpers
    comm = engine.Comm.from_torch_group()
    names = helpers.c2_queue_names(64)
    W = engine.rec_bytes_for(names)

    # ---- tally: keys split by contiguous range, counts all-reduced -------------------------
    lo, hi = replay.shard_range(N_KEYS, rank, world)
    heads, _, _ = engine.synth_keys_host(5, lo, hi - lo, names, rec_bytes=W)
    table = torch.from_numpy(engine.build_prefix_table(names, W)).to(dev)
    d_heads = torch.from_numpy(heads).to(dev)
    counts = engine.match_count(d_heads, hi - lo, W, table, len(names))
    counts_torch = counts.clone()
    replay.allreduce_sum_(counts, comm=comm)          # libkas -> ncclAllReduce
    replay.allreduce_sum_(counts_torch)               # torch.distributed
    # the streamed (host memory) form with the all-reduce inside the pipeline
    st = snapshot.StreamingTally(names, chunk_keys=400_000, device=dev)
    streamed = st.run(torch.from_numpy(heads).pin_memory(), hi - lo, reduce=comm.allreduce_sum_)

    # ---- replay: resources split with their queues kept together, sums all-reduced -----------
    rng = np.random.default_rng(3)
    llen, go = helpers.random_fleet(rng, T, Q, grouped=True)
    r_lo, r_hi, q_lo, q_hi, local_go = replay.shard_groups(go, rank, world)
    # (columns padded to a multiple of 4 so the shard keeps 16-byte rows)
    width = (q_hi - q_lo + 3) // 4 * 4
    part = np.zeros((T, width), dtype=np.int32)
    part[:, :q_hi - q_lo] = llen[:, q_lo:q_hi]
    d_go = torch.from_numpy(local_go.view(np.int64)).to(dev)
    out = engine.replay(torch.from_numpy(part).to(dev), T, width, POLICIES, d_group_off=d_go,
                        n_resources=r_hi - r_lo)
    sums = out['sums']
    replay.allreduce_sum_(sums, comm=comm)
    # single-queue resources through the vectorised kernels, split by queue
    a, b = replay.shard_range(Q, rank, world, align=8)
    ident = engine.replay(torch.from_numpy(np.ascontiguousarray(llen[:, a:b])).to(dev), T, b - a, POLICIES)
    sums_id = ident['sums']
    replay.allreduce_sum_(sums_id, comm=comm)
    u16 = engine.replay(torch.from_numpy(np.ascontiguousarray(llen[:, a:b]).astype(np.uint16).view(np.int16)).to(dev),
                        T, b - a, POLICIES[:1], want_totals=False)['sums']
    replay.allreduce_sum_(u16, comm=comm)
    torch.cuda.synchronize()
    np.savez(os.path.join(out_dir, 'rank%d.npz' % rank), counts=counts.cpu().numpy(),
             counts_torch=counts_torch.cpu().numpy(), streamed=np.array(streamed, dtype=np.uint64).astype(np.int64),
             sums=sums.cpu().numpy(), sums_id=sums_id.cpu().numpy(), u16=u16.cpu().numpy())
    dist.barrier()
    comm.destroy()
    dist.destroy_process_group()


def _expected():
    """The oracle's integers for the whole keyspace / fleet (computed once per test session)."""
    from autoscaler import engine
    from oracle import coracle
    from tests import helpers
    names = helpers.c2_queue_names(64)
    W = engine.rec_bytes_for(names)
    heads = coracle.synth_heads(5, 0, N_KEYS, names, W)
    counts = coracle.match_counts_heads(names, heads).astype(np.int64)
    rng = np.random.default_rng(3)
    llen, go = helpers.random_fleet(rng, T, Q, grouped=True)

    def sums_of(res):
        return np.stack([res['pod_ticks'].sum(0), res['n_patches'].sum(0), res['final'].astype(np.int64).sum(0)], axis=1)
    grouped = sums_of(coracle.replay(llen, None, go, None, POLICIES))
    ident = sums_of(coracle.replay(llen, None, None, None, POLICIES))
    return counts, grouped, ident, llen, go


@pytest.fixture(scope='module')
def expected():
    return _expected()


@pytest.mark.parametrize('world', [2, 4, 8])
def test_sharded_totals_equal_one_rank_and_the_oracle(world, tmp_path, expected):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < world:
        pytest.skip('needs %d GPUs, %d visible' % (world, torch.cuda.device_count()))
    from autoscaler import engine
    from tests import helpers
    counts, grouped, ident, llen, go = expected
    port = 29600 + (os.getpid() % 1500) + world
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    for rank in range(world):
        got = np.load(tmp_path / ('rank%d.npz' % rank))
        assert (got['counts'] == counts).all()
        assert (got['counts_torch'] == counts).all()
        assert (got['streamed'] == counts).all()
        assert (got['sums'] == grouped).all()
        assert (got['sums_id'] == ident).all()
        assert (got['u16'][0] == ident[0]).all()
    # ... and what ONE rank computes over the whole input
    names = helpers.c2_queue_names(64)
    W = engine.rec_bytes_for(names)
    heads, _, _ = engine.synth_keys_host(5, 0, N_KEYS, names, rec_bytes=W)
    table = torch.from_numpy(engine.build_prefix_table(names, W)).cuda()
    one = engine.counts_to_host(engine.match_count(torch.from_numpy(heads).cuda(), N_KEYS, W, table, len(names)))
    assert (one.astype(np.int64) == counts).all()
    d_go = torch.from_numpy(go.view(np.int64)).cuda()
    one_g = engine.replay(torch.from_numpy(llen).cuda(), T, Q, POLICIES, d_group_off=d_go, n_resources=len(go) - 1,
                          want_totals=False)['sums'].cpu().numpy()
    assert (one_g == grouped).all()


def test_comm_of_one_rank_and_bad_arguments():
    """A one-rank communicator is legal (the all-reduce is the identity); argument checks."""
    import ctypes as C
    import torch
    from autoscaler import engine
    comm = engine.Comm(1, 0, engine.Comm.unique_id())
    t = torch.arange(7, dtype=torch.int64, device='cuda')
    engine.check(engine.lib().kas_allreduce_i64(comm._comm, engine.ptr(t), 7, engine.stream_ptr()), 'kas_allreduce_i64')
    torch.cuda.synchronize()
    assert t.tolist() == list(range(7))
    with pytest.raises(TypeError):
        comm.allreduce_sum_(torch.zeros(3, dtype=torch.int32, device='cuda'))
    assert engine.lib().kas_allreduce_i64(None, engine.ptr(t), 7, None) == engine.KAS_ERR_INVALID
    assert engine.lib().kas_comm_init(None, 1, 0, None) == engine.KAS_ERR_INVALID
    c = C.c_void_p(0)
    buf = C.create_string_buffer(engine.COMM_ID_BYTES)
    assert engine.lib().kas_comm_init(C.byref(c), 2, 2, buf) == engine.KAS_ERR_INVALID
    comm.destroy()
    with pytest.raises(ValueError):
        engine.Comm(1, 0, b'short')
