"""Key-snapshot ingest from an RDB dump (kas_rdb_keys_host).  No Redis server
exists in this image, so the dumps are written by tests/rdb_writer.py from the published format:
parity of the READER with real dumps is unpinned, parity of what is done WITH the keys is the
usual one (oracle tally of the same keys)."""
import numpy as np
import pytest

from tests import helpers
from tests.rdb_writer import Rdb, listpack, lzf_compress, ziplist


def _dump():
    long_key = b'processing-predict:' + b'abcabcabc' * 12 + b':host-77'
    r = Rdb(11).aux(b'redis-ver', b'7.2.4').aux(b'redis-bits', b'64').select(0, (9, 2))
    r.key(0, b'plain', Rdb.v_string(b'hello'))
    r.key(0, b'12345', Rdb.v_string(b'7', 'int'), how='int')
    r.key(0, b'-7', Rdb.v_string(b'x'), how='int')
    r.key(0, b'-30000', Rdb.v_string(b'x'), how='int')
    r.key(0, b'2000000000', Rdb.v_string(b'x'), how='int')
    r.key(0, long_key, Rdb.v_string(b'z' * 100, 'lzf'), how='lzf', expire_ms=1700000000000)
    r.key(1, b'predict', Rdb.v_list([b'a', b'b', b'c']), expire_s=1700000000)
    r.key(10, b'track', Rdb.v_blob(ziplist(5)))
    r.key(14, b'train', Rdb.v_quicklist([4, 7, 1]), idle=5)
    r.key(14, b'trainz', Rdb.v_quicklist([40, 70], 'lzf'), freq=3)
    r.key(18, b'segment', Rdb.v_quicklist2([('packed', 9), ('plain', b'big element'), ('packed', 2)]))
    r.key(18, b'full', Rdb.v_quicklist2([('packed', 70000)]))
    r.key(4, b'processing-track:h1', Rdb.v_hash([(b'f', b'v'), (b'g', b'w')]))
    r.key(4, b'processing-track:h2', Rdb.v_hash([(b'f', b'v')]))
    r.key(2, b'aset', Rdb.v_set([b'1', b'2']))
    r.key(3, b'z1', Rdb.v_zset_text([(b'm', 1.5), (b'n', 'inf'), (b'o', 'nan')]))
    r.key(5, b'z2', Rdb.v_zset_bin([(b'm', 1.5)]))
    r.select(3)
    for t in (9, 11, 12, 13, 16, 17, 20):
        r.key(t, b'blob%d' % t, Rdb.v_blob())
    for t in (15, 19, 21):
        r.key(t, b'stream%d' % t, Rdb.v_stream(t))
    r.key(7, b'mod', Rdb.v_module2())
    r.key(0, b'k' * 70, Rdb.v_string(b'x'), how=14)
    r.key(0, b'k' * 300, Rdb.v_string(b'x'), how=32)
    r.key(0, b'k' * 20, Rdb.v_string(b'x'), how=64)
    r.key(4, b'processing-predict:in-db3', Rdb.v_hash([(b'f', b'v')]))
    return r.finish(), long_key


def test_rdb_reader_walks_every_encoding():
    from autoscaler import engine
    data, long_key = _dump()
    d = engine.rdb_keys_host(data)
    keys = [bytes(d['blob'][d['off'][i]:d['off'][i + 1]]) for i in range(len(d['type']))]
    assert keys[:6] == [b'plain', b'12345', b'-7', b'-30000', b'2000000000', long_key]
    assert len(keys) == 32 and keys[-1] == b'processing-predict:in-db3'
    assert d['db'].tolist() == [0] * 17 + [3] * 15
    by = {k: (int(t), int(n)) for k, t, n in zip(keys, d['type'], d['list_len'])}
    assert by[b'predict'] == (1, 3) and by[b'track'] == (10, 5) and by[b'train'] == (14, 12)
    assert by[b'trainz'] == (14, 110) and by[b'segment'] == (18, 12)
    assert by[b'full'] == (18, engine.RDB_UNKNOWN_LEN)           # saturated 16-bit counter: not guessed
    assert by[b'aset'] == (2, 0) and by[b'stream21'] == (21, 0) and by[b'k' * 300] == (0, 0)
    # LZF round trip of the writer's own compressor through the reader
    for s in (b'a' * 500, b'abcabcabc' * 40 + b'xyz', bytes(range(256)) * 3, b'q'):
        k = engine.rdb_keys_host(Rdb().key(0, s, Rdb.v_string(b'v'), how='lzf').finish())
        assert bytes(k['blob']) == s and len(lzf_compress(b'a' * 500)) < 20


def test_rdb_reader_refuses_what_it_cannot_read():
    from autoscaler import engine
    data, _ = _dump()
    for bad in (data[:-20], data[:200], b'REDIS0009', b'RODIS0009\xff', b'REDIS00x9\xff', b''):
        with pytest.raises((engine.KasError, ValueError)) as e:
            engine.rdb_keys_host(bad)
        assert 'invalid' in str(e.value) or isinstance(e.value, ValueError)
    for t in (6, 8, 22, 30):                     # old module format and types this reader does not know
        with pytest.raises(engine.KasError) as e:
            engine.rdb_keys_host(Rdb().key(t, b'm', b'').finish())
        assert e.value.status == engine.KAS_ERR_UNSUPPORTED
    with pytest.raises(engine.KasError):         # unknown opcode
        engine.rdb_keys_host(b'REDIS0011\xf1\x00\xff')
    empty = engine.rdb_keys_host(b'REDIS0009\xff')
    assert len(empty['type']) == 0 and empty['off'].tolist() == [0]
    # a corrupt LZF key is an error, not garbage
    good = Rdb().key(0, b'abcabcabc' * 12, Rdb.v_string(b'v'), how='lzf').finish()
    broken = bytearray(good)
    broken[14] ^= 0xFF
    with pytest.raises(engine.KasError):
        engine.rdb_keys_host(bytes(broken))


def test_snapshot_from_rdb_filters_databases_and_carries_llen(tmp_path):
    from autoscaler import snapshot
    data, long_key = _dump()
    snap = snapshot.KeySnapshot.from_rdb(data)
    assert len(snap) == 32
    p = tmp_path / 'dump.rdb'
    p.write_bytes(data)
    snap0 = snapshot.KeySnapshot.from_rdb(str(p), db=0)
    assert len(snap0) == 17 and snap0.list_lengths[b'train'] == 12 and snap0.list_lengths[b'full'] is None
    snap3 = snapshot.KeySnapshot.from_rdb(data, db=3)
    assert len(snap3) == 15 and snap3.list_lengths == {}
    assert bytes(snap3.blob[snap3.off[-2]:snap3.off[-1]]) == b'processing-predict:in-db3'
    with pytest.raises(ValueError):
        snapshot.SnapshotRedis.from_rdb(data, db=0)        # 'full' has no LLEN in its header
    small = Rdb().select(0).key(1, b'predict', Rdb.v_list([b'a'] * 4)).key(
        18, b'track', Rdb.v_quicklist2([('packed', 6)])).key(0, b'notalist', Rdb.v_string()).finish()
    r = snapshot.SnapshotRedis.from_rdb(small)
    assert r.llen('predict') == 4 and r.llen('track') == 6 and r.llen('notalist') == 0 and r.llen('absent') == 0


@pytest.mark.gpu
def test_tally_queues_from_an_rdb_dump():
    """The whole of tally_queues (LLEN + SCAN term, autoscaler.py:60-77) from a dump file."""
    import autoscaler
    from autoscaler import engine
    from oracle import coracle
    names = helpers.c2_queue_names(12)
    _, blob, off = engine.synth_keys_host(41, 0, 4000, names, long_tail=True, want_heads=False, want_blob=True)
    keys = sorted(set(bytes(blob[off[i]:off[i + 1]]) for i in range(4000)) - set(n.encode() for n in names))
    r = Rdb(10).select(0)
    for i, k in enumerate(keys):
        how = 'lzf' if (len(k) > 60 and i % 2) else 'raw'
        r.key(4, k, Rdb.v_hash([(b'status', b'started')]), how=how, expire_ms=1700000000000 + i if i % 3 == 0 else None)
    lists = {'predict': 7, 'track': 0, 'segment': 31}
    r.key(1, b'predict', Rdb.v_list([b'j'] * 7))
    r.key(18, b'segment', Rdb.v_quicklist2([('packed', 30), ('plain', b'x' * 100)]))
    scaler = autoscaler.Autoscaler(autoscaler.snapshot.SnapshotRedis.from_rdb(r.finish()), ','.join(names))
    scaler.tally_queues()
    kb, ko = coracle.keys_to_blob(keys + [b'predict', b'segment'])
    want = coracle.match_counts(names, kb, ko).tolist()
    assert [scaler.redis_keys[q] for q in names] == [w + lists.get(q, 0) for q, w in zip(names, want)]
    assert sum(want) > 1000
