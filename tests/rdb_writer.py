"""A small RDB writer for the tests of the dump reader (kas_rdb_keys_host): builds dump bytes
with the encodings the published format defines -- 6/14/32/64-bit lengths, raw / integer / LZF
strings, expiry and LRU/LFU prefixes, several databases, AUX fields, and the value types a
kiosk Redis holds (strings, lists in linked / ziplist / quicklist / listpack form, hashes, sets,
sorted sets, streams, module values).  Test infrastructure only: the values' payloads are
plausible fillers, only their framing matters to a reader that skips them."""
import struct


def enc_len(n):
    if n < 64:
        return bytes([n])
    if n < 16384:
        return bytes([0x40 | (n >> 8), n & 0xFF])
    if n < 2 ** 32:
        return b'\x80' + struct.pack('>I', n)
    return b'\x81' + struct.pack('>Q', n)


def enc_len_forced(n, form):
    """The same length in a wider form than needed (readers must accept it)."""
    if form == 14:
        return bytes([0x40 | (n >> 8), n & 0xFF])
    if form == 32:
        return b'\x80' + struct.pack('>I', n)
    return b'\x81' + struct.pack('>Q', n)


def lzf_compress(data):
    """Greedy LZF: literal runs and back references (3..264 bytes, distance <= 8191)."""
    out = bytearray()
    lit = bytearray()
    i, n = 0, len(data)
    table = {}

    def flush():
        for k in range(0, len(lit), 32):
            run = lit[k:k + 32]
            out.append(len(run) - 1)
            out.extend(run)
        del lit[:]
    while i < n:
        best = 0
        if i + 3 <= n:
            j = table.get(bytes(data[i:i + 3]))
            if j is not None and 0 < i - j <= 8191:
                k = 0
                while i + k < n and k < 264 and data[j + k] == data[i + k]:
                    k += 1
                best, ref = k, j
            table[bytes(data[i:i + 3])] = i
        if best >= 3:
            flush()
            dist = i - ref - 1
            ln = best - 2
            if ln < 7:
                out.append((ln << 5) | (dist >> 8))
            else:
                out.append((7 << 5) | (dist >> 8))
                out.append(ln - 7)
            out.append(dist & 0xFF)
            i += best
        else:
            lit.append(data[i])
            i += 1
    flush()
    return bytes(out)


def enc_string(s, how='raw'):
    s = bytes(s)
    if how == 'raw':
        return enc_len(len(s)) + s
    if how == 'int':                      # the decimal text of a small integer
        v = int(s)
        if -128 <= v < 128:
            return b'\xC0' + struct.pack('<b', v)
        if -32768 <= v < 32768:
            return b'\xC1' + struct.pack('<h', v)
        return b'\xC2' + struct.pack('<i', v)
    if how == 'lzf':
        c = lzf_compress(s)
        return b'\xC3' + enc_len(len(c)) + enc_len(len(s)) + c
    if how in (14, 32, 64):
        return enc_len_forced(len(s), how) + s
    raise ValueError(how)


def ziplist(n_entries):
    body = b'\x00\x01a' * n_entries        # filler entries
    total = 10 + len(body) + 1
    return struct.pack('<IIH', total, 10, min(n_entries, 65535)) + body + b'\xFF'


def listpack(n_entries):
    body = b'\x81a\x02' * n_entries         # filler entries
    total = 6 + len(body) + 1
    return struct.pack('<IH', total, min(n_entries, 65535)) + body + b'\xFF'


class Rdb(object):
    def __init__(self, version=9):
        self.buf = bytearray(b'REDIS%04d' % version)

    def aux(self, k, v):
        self.buf += b'\xFA' + enc_string(k) + enc_string(v)
        return self

    def select(self, db, sizes=None):
        self.buf += b'\xFE' + enc_len(db)
        if sizes:
            self.buf += b'\xFB' + enc_len(sizes[0]) + enc_len(sizes[1])
        return self

    def key(self, vtype, key, value, how='raw', expire_ms=None, expire_s=None, idle=None, freq=None):
        if expire_ms is not None:
            self.buf += b'\xFC' + struct.pack('<Q', expire_ms)
        if expire_s is not None:
            self.buf += b'\xFD' + struct.pack('<I', expire_s)
        if idle is not None:
            self.buf += b'\xF8' + enc_len(idle)
        if freq is not None:
            self.buf += b'\xF9' + bytes([freq])
        self.buf += bytes([vtype]) + enc_string(key, how) + value
        return self

    # values -------------------------------------------------------------------------------
    @staticmethod
    def v_string(s=b'v', how='raw'):
        return enc_string(s, how)

    @staticmethod
    def v_list(items):
        return enc_len(len(items)) + b''.join(enc_string(i) for i in items)

    v_set = v_list

    @staticmethod
    def v_hash(pairs):
        return enc_len(len(pairs)) + b''.join(enc_string(k) + enc_string(v) for k, v in pairs)

    @staticmethod
    def v_zset_text(pairs):
        out = enc_len(len(pairs))
        for m, score in pairs:
            out += enc_string(m)
            if score == 'inf':
                out += b'\xFE'
            elif score == 'nan':
                out += b'\xFD'
            else:
                t = repr(float(score)).encode()
                out += bytes([len(t)]) + t
        return out

    @staticmethod
    def v_zset_bin(pairs):
        return enc_len(len(pairs)) + b''.join(enc_string(m) + struct.pack('<d', float(sc)) for m, sc in pairs)

    @staticmethod
    def v_blob(payload=b'\x00' * 12, how='raw'):
        return enc_string(payload, how)

    @staticmethod
    def v_quicklist(node_sizes, how='raw'):
        return enc_len(len(node_sizes)) + b''.join(enc_string(ziplist(n), how) for n in node_sizes)

    @staticmethod
    def v_quicklist2(nodes, how='raw'):
        """nodes: ('plain', bytes) or ('packed', n_entries)."""
        out = enc_len(len(nodes))
        for kind, x in nodes:
            if kind == 'plain':
                out += enc_len(1) + enc_string(x)
            else:
                out += enc_len(2) + enc_string(listpack(x), how)
        return out

    @staticmethod
    def v_stream(vtype, n_packs=2, groups=1, pel=2, consumers=2):
        out = enc_len(n_packs)
        for i in range(n_packs):
            out += enc_string(b'\x00' * 16) + enc_string(listpack(3))
        out += enc_len(6) + enc_len(1700000000000) + enc_len(5)
        if vtype >= 19:
            out += enc_len(1) + enc_len(0) + enc_len(0) + enc_len(0) + enc_len(6)
        out += enc_len(groups)
        for g in range(groups):
            out += enc_string(b'group%d' % g) + enc_len(1700000000000) + enc_len(2)
            if vtype >= 19:
                out += enc_len(4)
            out += enc_len(pel)
            for i in range(pel):
                out += b'\x11' * 16 + struct.pack('<Q', 1700000000123) + enc_len(1)
            out += enc_len(consumers)
            for c in range(consumers):
                out += enc_string(b'consumer%d' % c) + struct.pack('<Q', 1700000000456)
                if vtype >= 21:
                    out += struct.pack('<Q', 1700000000789)
                out += enc_len(1) + b'\x22' * 16
        return out

    @staticmethod
    def v_module2():
        return (enc_len(0x1234567890) + enc_len(2) + enc_len(77) + enc_len(5) + enc_string(b'payload') +
                enc_len(4) + struct.pack('<d', 1.5) + enc_len(3) + struct.pack('<f', 2.5) +
                enc_len(1) + enc_len(9) + enc_len(0))

    def finish(self, checksum=True):
        self.buf += b'\xFF'
        if checksum:
            self.buf += b'\x00' * 8          # a zero CRC64 means "not checked"
        return bytes(self.buf)
