"""Key-snapshot ingest and the snapshot-backed Redis seam.

The reference walks the whole keyspace over TCP once per queue
(autoscaler.py:69-71 of the reference).  Here the keyspace is taken ONCE,
reduced to fixed-width heads (the first ``rec_bytes`` bytes of every key, zero
padded) and handed to the match kernel, which counts every queue in one pass.
"""
from __future__ import annotations

import numpy as np

from autoscaler import engine


class KeySnapshot(object):
    """A keyspace snapshot as a blob of key bytes + offsets (host side).

    Heads are packed on demand for a given record width; a snapshot built
    straight from packed heads (``from_heads``) has no blob.
    """

    def __init__(self, blob=None, off=None, sep_bytes=0, heads=None):
        self.blob = blob
        self.off = off
        self.sep_bytes = sep_bytes
        self._heads = {} if heads is None else {heads.shape[1]: heads}
        self._device = {}

    # -- constructors -------------------------------------------------------
    @classmethod
    def from_keys(cls, keys):
        """From an iterable of keys (bytes or str; str is encoded as UTF-8 the
        way the reference's client decodes them, redis.py:159-161)."""
        enc = [k if isinstance(k, (bytes, bytearray)) else str(k).encode('utf-8') for k in keys]
        off = np.zeros(len(enc) + 1, dtype=np.uint64)
        if enc:
            off[1:] = np.cumsum([len(k) for k in enc], dtype=np.uint64)
        blob = np.frombuffer(b''.join(enc) + b'\0', dtype=np.uint8)
        return cls(blob=blob, off=off)

    @classmethod
    def from_blob(cls, blob, off):
        return cls(blob=np.ascontiguousarray(blob, dtype=np.uint8),
                   off=np.ascontiguousarray(off, dtype=np.uint64))

    @classmethod
    def from_dump(cls, data):
        """From a newline-separated key dump (``redis-cli --scan`` output)."""
        blob = np.frombuffer(bytes(data), dtype=np.uint8)
        return cls(blob=blob, off=engine.index_lines_host(blob), sep_bytes=1)

    @classmethod
    def from_rdb(cls, data, db=None):
        """From a Redis RDB dump (bytes, or a path): every key of the dump -- of database
        ``db`` only when given -- without a server (``kas_rdb_keys_host``: values are skipped by
        type, LZF-compressed and integer-encoded keys restored).  The snapshot also carries
        ``list_lengths``: {key bytes: elements} for the keys that hold lists, i.e. what ``LLEN``
        would answer (autoscaler.py:67 of the reference)."""
        if isinstance(data, str):
            with open(data, 'rb') as f:
                data = f.read()
        d = engine.rdb_keys_host(data)
        off = d['off']
        keep = np.ones(len(d['type']), dtype=bool) if db is None else d['db'] == db
        lens = np.diff(off.astype(np.int64))
        if not keep.all():
            idx = np.flatnonzero(keep)
            parts = [d['blob'][off[i]:off[i + 1]] for i in idx]
            blob = np.concatenate(parts + [np.zeros(1, np.uint8)]) if len(parts) else np.zeros(1, np.uint8)
            new_off = np.zeros(len(idx) + 1, dtype=np.uint64)
            new_off[1:] = np.cumsum(lens[idx])
        else:
            idx = np.arange(len(lens))
            blob, new_off = d['blob'], off
        snap = cls(blob=blob, off=new_off)
        snap.list_lengths = {}
        for j, i in enumerate(idx):
            if int(d['type'][i]) in engine.RDB_LIST_TYPES:
                n = int(d['list_len'][i])
                snap.list_lengths[bytes(blob[new_off[j]:new_off[j + 1]])] = None if n == engine.RDB_UNKNOWN_LEN else n
        return snap

    @classmethod
    def from_heads(cls, heads):
        """From already packed heads: uint8 [n][rec_bytes] numpy array or tensor."""
        return cls(heads=heads)

    @classmethod
    def from_redis(cls, redis_client, match='processing-*', count=1000):
        """One SCAN of a live server through any client exposing ``scan_iter``
        (the RedisClient seam, redis.py:163-202 of the reference), kept as a snapshot
        object.  Only keys that can match some ``processing-<q>:*`` are fetched.  (The live
        tally does not build this object: ``tally_live`` streams the scan to the GPU.)"""
        return cls.from_keys(redis_client.scan_iter(match=match, count=count))

    # -- access -------------------------------------------------------------
    def __len__(self):
        if self.off is not None:
            return len(self.off) - 1
        return next(iter(self._heads.values())).shape[0]

    def heads(self, rec_bytes):
        """uint8 [n][rec_bytes] host heads for this width."""
        if rec_bytes not in self._heads:
            if self.off is None:
                have = next(iter(self._heads))
                raise ValueError('snapshot was built from %d-byte heads; %d-byte heads needed '
                                 '(a queue name is too long for it)' % (have, rec_bytes))
            self._heads[rec_bytes] = engine.pack_heads_host(self.blob, self.off, rec_bytes,
                                                            sep_bytes=self.sep_bytes)
        return self._heads[rec_bytes]

    def device_blob(self, device='cuda'):
        """(blob, offsets) resident in HBM, for the glob kernel."""
        torch = engine.require_cuda()
        key = ('blob', str(device))
        if key not in self._device:
            blob = np.ascontiguousarray(self.blob)
            if blob.size == 0:
                blob = np.zeros(1, dtype=np.uint8)
            self._device[key] = (torch.from_numpy(blob.copy()).to(device),
                                 torch.from_numpy(self.off.view(np.int64).copy()).to(device))
        return self._device[key]

    def device_heads(self, rec_bytes, device='cuda'):
        """Heads resident in HBM (cached per width)."""
        torch = engine.require_cuda()
        key = (rec_bytes, str(device))
        if key not in self._device:
            h = self.heads(rec_bytes)
            if isinstance(h, np.ndarray):
                h = torch.from_numpy(h)
            self._device[key] = h.to(device, non_blocking=True) if h.device.type == 'cpu' else h
        return self._device[key]


def _is_glob(q):
    return any(ch in engine.GLOB_META for ch in engine.encode_name(q))


MAX_PREFIX_NAME = engine.MAX_REC_BYTES - 12      # longest name a 64-byte record holds


def _needs_glob_kernel(q):
    """Names the prefix table cannot hold: live globs (metacharacters), names longer
    than a record, names with a NUL byte.  The glob kernel matches them on whole keys."""
    b = engine.encode_name(q)
    return _is_glob(q) or len(b) > MAX_PREFIX_NAME or 0 in b


def _glob_escape(name):
    """A literal name as a Redis glob (only used for over-long / NUL names, which
    by definition hold no metacharacter)."""
    return bytes(name)


def tally_snapshot(snapshot, queue_names, device='cuda'):
    """SCAN-count term for every queue over one snapshot: list of int, in the
    order of ``queue_names`` (autoscaler.py:69-71 of the reference).  Literal
    names go through the prefix kernel in one pass (several passes beyond 1024
    names); names holding glob metacharacters -- live globs in the reference --
    and names too long for a record go through the glob kernel."""
    names = list(queue_names)
    out = [0] * len(names)
    plain = [i for i, q in enumerate(names) if not _needs_glob_kernel(q)]
    globs = [i for i, q in enumerate(names) if _needs_glob_kernel(q)]
    for lo in range(0, len(plain), engine.MAX_PREFIXES):
        part = plain[lo:lo + engine.MAX_PREFIXES]
        for i, c in zip(part, _tally_prefixes(snapshot, [names[i] for i in part], device)):
            out[i] = c
    if globs:
        if snapshot.off is None:
            raise ValueError('queue names with glob metacharacters (or longer than %d bytes) need '
                             'the whole keys; this snapshot only holds fixed-width heads'
                             % MAX_PREFIX_NAME)
        d_blob, d_off = snapshot.device_blob(device)
        for lo in range(0, len(globs), 256):
            part = globs[lo:lo + 256]
            pats = [b'processing-' + _glob_escape(engine.encode_name(names[i])) + b':*' for i in part]
            counts = engine.glob_count(d_blob, d_off, len(snapshot), pats, sep_bytes=snapshot.sep_bytes)
            for i, c in zip(part, engine.counts_to_host(counts)):
                out[i] = int(c)
    return out


def _tally_prefixes(snapshot, names, device):
    torch = engine.require_cuda()
    rec_bytes = engine.rec_bytes_for(names)
    try:
        image = engine.build_prefix_table(names, rec_bytes)
    except engine.KasError as err:
        # the table of this many names does not fit shared memory at this record width
        if err.status != engine.KAS_ERR_UNSUPPORTED or len(names) < 2:
            raise
        half = len(names) // 2
        return (_tally_prefixes(snapshot, names[:half], device) +
                _tally_prefixes(snapshot, names[half:], device))
    table = torch.from_numpy(image).to(device)
    heads = snapshot.device_heads(rec_bytes, device)
    counts = engine.match_count(heads, len(snapshot), rec_bytes, table, len(names))
    return [int(c) for c in engine.counts_to_host(counts)]


class StreamingTally(object):
    """Tally a snapshot whose packed heads sit in (pinned) host memory: the heads
    are copied to the GPU in chunks on a copy stream into two staging buffers
    while the match kernel consumes the previous chunk, and only the uint64
    counts come back.  This is the end-to-end form of ``tally_queues``'s SCAN
    term for snapshots that are not resident in HBM (or larger than it)."""

    def __init__(self, queue_names, chunk_keys=8 << 20, device='cuda'):
        torch = engine.require_cuda()
        self.names = list(queue_names)
        for q in self.names:
            if _is_glob(q):
                engine.check(engine.KAS_ERR_GLOB_META, 'streaming tally of queue %r' % (q,))
        self.rec_bytes = engine.rec_bytes_for(self.names)
        self.device = torch.device(device)
        self.chunk_keys = int(chunk_keys)
        self.table = torch.from_numpy(engine.build_prefix_table(self.names, self.rec_bytes)).to(self.device)
        self.stage = [torch.empty(self.chunk_keys * self.rec_bytes, dtype=torch.uint8, device=self.device)
                      for _ in range(2)]
        self.free = [torch.cuda.Event() for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.counts = torch.empty(len(self.names), dtype=torch.int64, device=self.device)
        self.host_counts = torch.empty(len(self.names), dtype=torch.int64).pin_memory()
        self.launches = 0

    def enqueue(self, host_heads, n_keys, reduce=None):
        """Queue the whole pipeline on the current stream (no host sync).
        ``host_heads``: uint8 tensor of n_keys*rec_bytes bytes, ideally pinned.
        ``reduce``: optional callable applied to the device counts before the copy
        back (the box-wide all-reduce).  Returns the pinned host tensor the counts
        land in once the stream has drained."""
        torch = engine.require_cuda()
        W = self.rec_bytes
        flat = host_heads.reshape(-1)
        main = torch.cuda.current_stream(self.device)
        self.counts.zero_()
        for b in range(2):
            self.free[b].record(main)
        i = 0
        for lo in range(0, n_keys, self.chunk_keys):
            hi = min(n_keys, lo + self.chunk_keys)
            b = i & 1
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(self.free[b])
                self.stage[b][:(hi - lo) * W].copy_(flat[lo * W:hi * W], non_blocking=True)
                ready = torch.cuda.Event()
                ready.record(self.copy_stream)
            main.wait_event(ready)
            engine.match_count(self.stage[b], hi - lo, W, self.table, len(self.names),
                               d_counts=self.counts, accumulate=True)
            self.free[b].record(main)
            i += 1
        self.launches += i
        if reduce is not None:
            reduce(self.counts)
        self.host_counts.copy_(self.counts, non_blocking=True)
        return self.host_counts

    def plan_blob_chunks(self, off, blob_bytes_per_chunk):
        """Key ranges [lo, hi) whose bytes fit one blob staging buffer and whose count fits
        ``chunk_keys`` (host side, from the offsets of a blob + offsets snapshot)."""
        off = np.ascontiguousarray(off, dtype=np.uint64)
        n = len(off) - 1
        out, lo = [], 0
        while lo < n:
            hi = int(np.searchsorted(off, off[lo] + np.uint64(blob_bytes_per_chunk), side='right')) - 1
            hi = max(lo + 1, min(hi, n, lo + self.chunk_keys))
            if int(off[hi] - off[lo]) > blob_bytes_per_chunk:
                raise ValueError('a key of %d bytes does not fit the %d-byte staging buffer'
                                 % (int(off[hi] - off[lo]), blob_bytes_per_chunk))
            out.append((lo, hi))
            lo = hi
        return out

    def enqueue_blob(self, host_blob, host_off, chunks, blob_bytes_per_chunk, sep_bytes=0, reduce=None):
        """The same pipeline from the RAW snapshot: variable-length keys as a blob + uint64
        offsets in (pinned) host memory.  Per chunk of keys the bytes and the offsets are
        copied to the GPU, ``kas_pack_heads`` cuts the fixed-width heads there, and the match
        kernel counts them; nothing is packed on the host.  ``chunks`` comes from
        ``plan_blob_chunks``.  host_blob: uint8 tensor, host_off: int64-typed tensor [n+1]."""
        torch = engine.require_cuda()
        W = self.rec_bytes
        if not hasattr(self, 'blob_stage') or self.blob_stage[0].numel() < blob_bytes_per_chunk + 16:
            self.blob_stage = [torch.empty(blob_bytes_per_chunk + 16, dtype=torch.uint8, device=self.device)
                               for _ in range(2)]
            self.off_stage = [torch.empty(self.chunk_keys + 1, dtype=torch.int64, device=self.device)
                              for _ in range(2)]
        main = torch.cuda.current_stream(self.device)
        self.counts.zero_()
        for b in range(2):
            self.free[b].record(main)
        for i, (lo, hi) in enumerate(chunks):
            b = i & 1
            b0, b1 = int(host_off[lo]), int(host_off[hi])
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(self.free[b])
                self.blob_stage[b][:b1 - b0].copy_(host_blob[b0:b1], non_blocking=True)
                self.off_stage[b][:hi - lo + 1].copy_(host_off[lo:hi + 1], non_blocking=True)
                ready = torch.cuda.Event()
                ready.record(self.copy_stream)
            main.wait_event(ready)
            # offsets stay absolute: the blob pointer is moved back by the chunk's first offset
            # (only addresses inside the staging buffer are ever read)
            engine.check(engine.lib().kas_pack_heads(
                engine.C.c_void_p(self.blob_stage[b].data_ptr() - b0), engine.ptr(self.off_stage[b]),
                hi - lo, sep_bytes, W, engine.ptr(self.stage[b]), engine.stream_ptr()), 'kas_pack_heads')
            engine.match_count(self.stage[b], hi - lo, W, self.table, len(self.names),
                               d_counts=self.counts, accumulate=True)
            self.free[b].record(main)
        self.launches += 2 * len(chunks)
        if reduce is not None:
            reduce(self.counts)
        self.host_counts.copy_(self.counts, non_blocking=True)
        return self.host_counts

    def run(self, host_heads, n_keys, reduce=None):
        """enqueue() + wait; returns the counts as a list of int."""
        torch = engine.require_cuda()
        out = self.enqueue(host_heads, n_keys, reduce)
        torch.cuda.current_stream(self.device).synchronize()
        return [int(v) for v in out.numpy().view(np.uint64)]


def tally_live(redis_client, queue_names, match='processing-*', count=1000, chunk_keys=1 << 18,
               chunk_bytes=16 << 20, device='cuda'):
    """SCAN-count term of every queue from ONE scan of a live server, streamed.

    The reference walks the keyspace once per queue and materialises every match in a Python
    list (``len([x for x in scan])``, autoscaler.py:70-71).  Here the client's iterator
    (``scan_iter``, the RedisClient seam, redis.py:163-202 of the reference) is drained once:
    key bytes go straight into one of two pinned staging buffers (blob + offsets), and as soon
    as a buffer is full it is copied to the GPU, cut to heads by ``kas_pack_heads`` and counted
    by the match kernel while the scan fills the other buffer.  Nothing but the counts comes
    back.  Names that are live globs (or too long for a record) are counted by the glob kernel
    on the same chunks.  Returns a list of int in the order of ``queue_names``."""
    torch = engine.require_cuda()
    names = list(queue_names)
    plain = [i for i, q in enumerate(names) if not _needs_glob_kernel(q)]
    globs = [i for i, q in enumerate(names) if _needs_glob_kernel(q)]
    dev = torch.device(device)
    tables = []          # (indices, rec_bytes, device table, device counts)
    todo = [plain[lo:lo + engine.MAX_PREFIXES] for lo in range(0, len(plain), engine.MAX_PREFIXES)]
    while todo:
        part = todo.pop()
        part_names = [names[i] for i in part]
        W = engine.rec_bytes_for(part_names)
        try:
            image = engine.build_prefix_table(part_names, W)
        except engine.KasError as err:
            if err.status != engine.KAS_ERR_UNSUPPORTED or len(part) < 2:
                raise
            todo += [part[:len(part) // 2], part[len(part) // 2:]]
            continue
        tables.append((part, W, torch.from_numpy(image).to(dev),
                       torch.zeros(len(part), dtype=torch.int64, device=dev)))
    patterns = [b'processing-' + _glob_escape(engine.encode_name(names[i])) + b':*' for i in globs]
    glob_totals = [0] * len(globs)
    max_w = max([t[1] for t in tables], default=16)

    host_blob = [torch.empty(chunk_bytes + 16, dtype=torch.uint8).pin_memory() for _ in range(2)]
    host_off = [torch.empty(chunk_keys + 1, dtype=torch.int64).pin_memory() for _ in range(2)]
    views = [memoryview(t.numpy()) for t in host_blob]
    d_blob = [torch.empty(chunk_bytes + 16, dtype=torch.uint8, device=dev) for _ in range(2)]
    d_off = [torch.empty(chunk_keys + 1, dtype=torch.int64, device=dev) for _ in range(2)]
    d_heads = torch.empty(chunk_keys * max_w, dtype=torch.uint8, device=dev)
    done = [None, None]
    stream = torch.cuda.current_stream(dev)
    state = {'buf': 0, 'fill': 0, 'lens': [], 'chunks': 0, 'keys': 0}

    def flush():
        b, n = state['buf'], len(state['lens'])
        if n == 0:
            return
        off = host_off[b].numpy()
        off[0] = 0
        np.cumsum(state['lens'], out=off[1:n + 1])
        fill = state['fill']
        d_blob[b][:fill].copy_(host_blob[b][:fill], non_blocking=True)
        d_off[b][:n + 1].copy_(host_off[b][:n + 1], non_blocking=True)
        for part, W, table, counts in tables:
            engine.check(engine.lib().kas_pack_heads(engine.ptr(d_blob[b]), engine.ptr(d_off[b]), n, 0, W,
                                                     engine.ptr(d_heads), engine.stream_ptr()), 'kas_pack_heads')
            engine.match_count(d_heads, n, W, table, len(part), d_counts=counts, accumulate=True)
        if patterns:
            for lo in range(0, len(patterns), 256):
                c = engine.glob_count(d_blob[b], d_off[b], n, patterns[lo:lo + 256])
                for k, v in enumerate(engine.counts_to_host(c)):
                    glob_totals[lo + k] += int(v)
        ev = torch.cuda.Event()
        ev.record(stream)
        done[b] = ev
        state['chunks'] += 1
        state['keys'] += n
        # fill the other buffer next; the scan may not overwrite it before its last use is over
        state['buf'] = b ^ 1
        state['fill'] = 0
        state['lens'] = []
        if done[b ^ 1] is not None:
            done[b ^ 1].synchronize()

    for key in redis_client.scan_iter(match=match, count=count):
        kb = key if isinstance(key, (bytes, bytearray)) else str(key).encode('utf-8')
        n = len(kb)
        if n > chunk_bytes:
            raise ValueError('a key of %d bytes does not fit the %d-byte staging buffer' % (n, chunk_bytes))
        if state['fill'] + n > chunk_bytes or len(state['lens']) == chunk_keys:
            flush()
        a = state['fill']
        views[state['buf']][a:a + n] = kb
        state['fill'] = a + n
        state['lens'].append(n)
    flush()
    out = [0] * len(names)
    for part, W, table, counts in tables:
        for i, c in zip(part, counts.cpu().numpy().view(np.uint64)):
            out[i] = int(c)
    for i, c in zip(globs, glob_totals):
        out[i] = c
    tally_live.last = {'chunks': state['chunks'], 'keys': state['keys']}
    return out


class SnapshotRedis(object):
    """A Redis stand-in over a key snapshot: what the offline replay feeds the
    unchanged ``Autoscaler``.  ``llen`` answers from ``lists``;
    ``key_snapshot()`` hands the snapshot to ``Autoscaler.tally_queues``."""

    def __init__(self, snapshot=None, lists=None):
        self.snapshot = snapshot if snapshot is not None else KeySnapshot.from_keys([])
        self.lists = dict(lists or {})

    @classmethod
    def from_rdb(cls, data, db=0):
        """The whole seam from one RDB dump: the keyspace for the SCAN term and the lengths of
        the list keys for ``llen`` (a key that is absent or not a list answers 0 here; the live
        server raises WRONGTYPE for a non-list, which the reference re-raises, redis.py:194-195)."""
        snap = KeySnapshot.from_rdb(data, db=db)
        lists = {}
        for k, n in snap.list_lengths.items():
            if n is None:
                raise ValueError('list %r has a node with a saturated element counter; its LLEN is not '
                                 'in the dump header' % (k,))
            lists[k.decode('utf-8', 'surrogateescape')] = n
        return cls(snap, lists)

    def llen(self, name):
        return int(self.lists.get(name, 0))

    def key_snapshot(self):
        return self.snapshot

    def scan_iter(self, match=None, count=None):
        raise NotImplementedError(
            'SnapshotRedis does not iterate keys on the host; the tally runs on the GPU '
            '(Autoscaler.tally_queues uses key_snapshot())')
