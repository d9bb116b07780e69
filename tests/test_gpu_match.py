"""K1 (prefix match and count) through the C ABI against the oracle."""
import numpy as np
import pytest

from tests import helpers

pytestmark = pytest.mark.gpu


def _gpu_counts(names, heads_np, rec_bytes, offset_bytes=0, accumulate_chunks=0):
    import torch
    from autoscaler import engine
    tbl = torch.from_numpy(engine.build_prefix_table(names, rec_bytes)).cuda()
    n = heads_np.shape[0]
    raw = torch.empty(n * rec_bytes + 64, dtype=torch.uint8, device='cuda')
    view = raw[offset_bytes:offset_bytes + n * rec_bytes]
    view.copy_(torch.from_numpy(heads_np.reshape(-1)))
    if not accumulate_chunks:
        c = engine.match_count(view, n, rec_bytes, tbl, len(names))
    else:
        c = torch.zeros(len(names), dtype=torch.int64, device='cuda')
        edges = np.linspace(0, n, accumulate_chunks + 1).astype(np.int64)
        for a, b in zip(edges[:-1], edges[1:]):
            engine.match_count(view[a * rec_bytes:], int(b - a), rec_bytes, tbl, len(names),
                               d_counts=c, accumulate=True)
    return engine.counts_to_host(c).tolist()


def test_golden_tally_cases():
    """Fixtures made by running the reference's tally_queues (literal names only)."""
    from autoscaler import engine
    from oracle import oracle
    ran = 0
    for case in helpers.golden_json('tally_cases'):
        names = list(oracle.parse_queues(case['queues'], case['delim']))
        if any(oracle.has_glob_meta(q.encode('utf-8')) for q in names):
            continue
        keys = helpers.latin(case['keyspace'])
        W = engine.rec_bytes_for(names)
        heads = oracle.pack_heads(keys, W)
        # names reach the server as UTF-8 (redis.py:159-161 of the reference); the fixture's
        # keys are raw bytes, so a latin-1 'tr\xe9' inside a key is not the queue name 'tré'
        got = _gpu_counts(names, heads, W) if len(keys) else [0] * len(names)
        want = [case['redis_keys'][q] - case['lists'].get(q, 0) for q in names]
        assert got == want, case['queues']
        ran += 1
    assert ran >= 20


@pytest.mark.parametrize('nq,maxlen', [(1, 4), (2, 20), (64, 20), (64, 36), (300, 52), (1024, 20)])
def test_synthetic_keyspace_vs_oracle(nq, maxlen):
    from autoscaler import engine
    from oracle import coracle
    rng = np.random.default_rng(nq * 1000 + maxlen)
    names = helpers.c2_queue_names(min(nq, 64))
    while len(names) < nq:
        names.append('n%d-%s' % (len(names), 'y' * int(rng.integers(0, maxlen - 8))))
    names = [q[:maxlen] for q in names]
    names = list(dict.fromkeys(names))
    names[0] = names[0].ljust(maxlen, 'z')[:maxlen]
    W = engine.rec_bytes_for(names)
    n = 200_003
    heads, blob, off = engine.synth_keys_host(11, 5, n, names, long_tail=True, rec_bytes=W,
                                              want_blob=True)
    want = coracle.match_counts(names, blob, off).tolist()
    assert sum(want) > n // 3
    assert coracle.match_counts_heads(names, heads).tolist() == want
    assert _gpu_counts(names, heads, W) == want
    assert _gpu_counts(names, heads, W, offset_bytes=16) == want      # 16-byte aligned only
    assert _gpu_counts(names, heads, W, accumulate_chunks=3) == want  # chunked accumulate


def test_edges_empty_ragged_and_adversarial():
    from autoscaler import engine
    from oracle import oracle
    names = ['predict', 'predict:zstack', 'track', 'tracker', '', 'a:b', 'a:b:c', 'a']
    keys = [b'processing-predict:zstack:h', b'processing-predict:zstack', b'processing-:', b'processing-::x',
            b'processing-a:b:c:d', b'processing-a:b:', b'processing-a:', b'processing-a', b'processing-tracker:1',
            b'processing-track:', b'processing-track:\x00\x00', b'processing-track\x00:', b'processing-trac:k',
            b'PROCESSING-track:1', b'processing-predict:\xff\xfe\xfd', b'', b'p', b'processing-',
            b'processing-predict:' + b':' * 40, b'processing-a:b:c' + b'x' * 700]
    W = engine.rec_bytes_for(names)
    assert W == 32
    for n in (0, 1, 2, 31, 32, 33, len(keys)):
        sub = (keys * 3)[:n]
        want = oracle.match_counts(names, sub)
        got = _gpu_counts(names, oracle.pack_heads(sub, W), W) if n else [0] * len(names)
        assert got == want, n
    # nested names count for every queue they match (reference: one SCAN per queue)
    want = oracle.match_counts(names, keys)
    assert want[names.index('a')] >= 4 and want[names.index('a:b')] >= 2


def test_rejections():
    import torch
    from autoscaler import engine
    with pytest.raises(engine.GlobQueueNameError):
        engine.build_prefix_table(['pre*'], 32)
    with pytest.raises(engine.KasError):
        engine.build_prefix_table(['a', 'a'], 32)
    with pytest.raises(engine.KasError):
        engine.build_prefix_table(['x' * 21], 32)
    # a table image that does not match the launch is reported, not trusted
    tbl = torch.from_numpy(engine.build_prefix_table(['a', 'b'], 32)).cuda()
    heads = torch.zeros(64 * 32, dtype=torch.uint8, device='cuda')
    c = engine.match_count(heads, 64, 32, tbl, 3)
    with pytest.raises(engine.KasError):
        engine.counts_to_host(c)
    with pytest.raises(engine.KasError):
        engine.match_count(heads[8:], 1, 32, tbl, 2)      # misaligned heads


def test_full_size_properties():
    """BASELINE config 2 size (1e8 keys x 64 prefixes): linearity over chunks and
    agreement with the oracle on a strided sample of whole chunks."""
    import torch
    from autoscaler import engine
    from oracle import coracle
    names = helpers.c2_queue_names(64)
    W = engine.rec_bytes_for(names)
    n = 100_000_000
    chunk = 12_500_000
    tbl = torch.from_numpy(engine.build_prefix_table(names, W)).cuda()
    d_heads = torch.empty(n * W, dtype=torch.uint8, device='cuda')
    host = torch.empty(chunk * W, dtype=torch.uint8).pin_memory()
    per_chunk = []
    for i in range(n // chunk):
        engine.synth_keys_host(2, i * chunk, chunk, names, rec_bytes=W, heads_out=host)
        d_heads[i * chunk * W:(i + 1) * chunk * W].copy_(host)
        c = engine.match_count(d_heads[i * chunk * W:], chunk, W, tbl, 64)
        per_chunk.append(engine.counts_to_host(c).astype(np.int64))
        if i in (0, 5):
            sample = host.numpy().reshape(chunk, W)[:2_000_000]
            cs = engine.match_count(d_heads[i * chunk * W:], 2_000_000, W, tbl, 64)
            assert engine.counts_to_host(cs).tolist() == coracle.match_counts_heads(names, sample).tolist()
    total = engine.counts_to_host(engine.match_count(d_heads, n, W, tbl, 64)).astype(np.int64)
    assert (total == np.sum(per_chunk, axis=0)).all()
    assert 0.45 * n < total.sum() < 0.75 * n


def test_device_pack_and_streaming_tally():
    """Ingest: the device packer equals the host packer / oracle, and the streaming
    (host-resident heads, chunked H2D) tally equals the resident one."""
    import torch
    from autoscaler import engine, snapshot
    from oracle import coracle, oracle
    names = helpers.c2_queue_names(40)
    n = 300_007
    heads, blob, off = engine.synth_keys_host(9, 0, n, names, long_tail=True, rec_bytes=32, want_blob=True)
    d_blob = torch.from_numpy(blob).cuda()
    d_off = torch.from_numpy(off.view(np.int64)).cuda()
    for W in (16, 32, 48, 64):
        dev = engine.pack_heads_device(d_blob, d_off, n, W).cpu().numpy()
        host = engine.pack_heads_host(blob, off, W)
        assert (dev == host).all()
        if W == 32:
            assert (host == heads).all()
            keys = [bytes(blob[off[i]:off[i + 1]]) for i in range(2000)]
            assert (host[:2000] == oracle.pack_heads(keys, W)).all()
    # newline dump, separator excluded, on the device packer too
    dump = b'\n'.join(bytes(blob[off[i]:off[i + 1]]) for i in range(5000)) + b'\n'
    s2 = snapshot.KeySnapshot.from_dump(dump)
    dd = engine.pack_heads_device(torch.from_numpy(np.frombuffer(dump, np.uint8).copy()).cuda(),
                                  torch.from_numpy(s2.off.view(np.int64)).cuda(), len(s2), 32, sep_bytes=1)
    assert (dd.cpu().numpy() == heads[:5000]).all()
    want = coracle.match_counts(names, blob, off).tolist()
    st = snapshot.StreamingTally(names, chunk_keys=65_536)
    pinned = torch.from_numpy(heads).pin_memory()
    assert st.run(pinned, n) == want
    assert st.run(pinned, n) == want            # reusable
    assert st.run(pinned, 1) == coracle.match_counts(names, blob[:off[1]], off[:2]).tolist()
