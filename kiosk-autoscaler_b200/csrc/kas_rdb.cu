// kas_rdb.cu -- key-snapshot ingest from a Redis RDB dump (host code only).
//
// The reference learns the keyspace by walking it over TCP, once per queue
// (autoscaler.py:67-71: LLEN q, then SCAN MATCH processing-<q>:*).  An RDB file is the same
// keyspace at rest: this reader walks the dump once and hands back every key (a blob + offsets,
// the form kas_pack_heads / kas_glob_count take), its value type, its database and -- for list
// keys -- the list's length, i.e. what LLEN would have answered.  Values are skipped by type,
// never decoded, apart from the element counts of list encodings.
//
// Format restated from the published RDB layout (versions 1-12): length encoding, string
// encoding (raw, 8/16/32-bit integer, LZF), opcodes AUX / SELECTDB / RESIZEDB / EXPIRETIME(_MS) /
// IDLE / FREQ / FUNCTION2 / SLOT_INFO / EOF, value types 0-5, 7, 9-21.  Anything else is refused
// (KAS_ERR_UNSUPPORTED): a dump is never half-read.
#include "kas_common.cuh"

#include <string.h>
#include <string>

namespace kas {

namespace {

struct Reader {
    const uint8_t *p, *end;
    bool ok;
    bool need(uint64_t n) {
        if ((uint64_t)(end - p) < n) ok = false;
        return ok;
    }
    uint8_t u8() {
        if (!need(1)) return 0;
        return *p++;
    }
    void skip(uint64_t n) {
        if (need(n)) p += n;
    }
    uint64_t be(int n) {
        uint64_t v = 0;
        if (!need((uint64_t)n)) return 0;
        for (int i = 0; i < n; ++i) v = (v << 8) | *p++;
        return v;
    }
    uint64_t le(int n) {
        uint64_t v = 0;
        if (!need((uint64_t)n)) return 0;
        for (int i = 0; i < n; ++i) v |= (uint64_t)p[i] << (8 * i);
        p += n;
        return v;
    }
};

// Length encoding.  *special is set for the "11" prefix, where the value is an encoding id
// (0/1/2: integer of 1/2/4 bytes, 3: LZF) instead of a length.
uint64_t read_len(Reader &r, bool *special) {
    if (special) *special = false;
    const uint8_t b = r.u8();
    switch (b >> 6) {
        case 0: return b & 0x3F;
        case 1: return ((uint64_t)(b & 0x3F) << 8) | r.u8();
        case 2:
            if (b == 0x80) return r.be(4);
            if (b == 0x81) return r.be(8);
            r.ok = false;
            return 0;
        default:
            if (special) *special = true; else r.ok = false;
            return b & 0x3F;
    }
}

bool lzf_decompress(const uint8_t *in, uint64_t in_len, uint8_t *out, uint64_t out_len) {
    const uint8_t *ip = in, *ie = in + in_len;
    uint8_t *op = out, *oe = out + out_len;
    while (ip < ie) {
        uint32_t ctrl = *ip++;
        if (ctrl < 32) {                       // literal run of ctrl + 1 bytes
            ctrl += 1;
            if ((uint64_t)(ie - ip) < ctrl || (uint64_t)(oe - op) < ctrl) return false;
            memcpy(op, ip, ctrl);
            op += ctrl;
            ip += ctrl;
        } else {                               // back reference
            uint32_t len = ctrl >> 5;
            if (ip >= ie) return false;
            if (len == 7) {
                len += *ip++;
                if (ip >= ie) return false;
            }
            const uint64_t dist = (((uint64_t)(ctrl & 0x1F)) << 8) + *ip++ + 1;
            len += 2;
            if ((uint64_t)(op - out) < dist || (uint64_t)(oe - op) < len) return false;
            const uint8_t *ref = op - dist;
            for (uint32_t i = 0; i < len; ++i) op[i] = ref[i];   // may overlap: byte by byte
            op += len;
        }
    }
    return op == oe;
}

// One string.  keep == nullptr: skipped.  Otherwise the decoded bytes are appended to *keep
// (integers as their decimal text, which is what the key is called on the wire).
// raw / raw_len (optional): where an uncompressed, non-integer payload lies inside the dump.
void read_string(Reader &r, std::string *keep, const uint8_t **raw = nullptr, uint64_t *raw_len = nullptr) {
    if (raw) *raw = nullptr;
    bool special = false;
    const uint64_t n = read_len(r, &special);
    if (!r.ok) return;
    if (!special) {
        if (!r.need(n)) return;
        if (keep) keep->append(reinterpret_cast<const char *>(r.p), n);
        if (raw) { *raw = r.p; *raw_len = n; }
        r.p += n;
        return;
    }
    if (n <= 2) {
        const int bytes = n == 0 ? 1 : (n == 1 ? 2 : 4);
        const uint64_t v = r.le(bytes);
        if (keep && r.ok) {
            const int64_t sv = bytes == 1 ? (int8_t)v : (bytes == 2 ? (int16_t)v : (int32_t)v);
            keep->append(std::to_string(sv));
        }
        return;
    }
    if (n == 3) {
        const uint64_t clen = read_len(r, nullptr), ulen = read_len(r, nullptr);
        if (!r.ok || !r.need(clen)) return;
        if (keep) {
            const size_t at = keep->size();
            keep->resize(at + ulen);
            if (!lzf_decompress(r.p, clen, reinterpret_cast<uint8_t *>(&(*keep)[at]), ulen)) r.ok = false;
        }
        r.p += clen;
        return;
    }
    r.ok = false;
}

constexpr uint64_t kUnknownLen = ~0ull;

// Entries of a ziplist / listpack payload, from its header (the 16-bit counter saturates at
// 65535; a node that full is reported as unknown rather than walked).
uint64_t packed_entries(const uint8_t *raw, uint64_t n, bool listpack) {
    if (!raw) return kUnknownLen;               // LZF-compressed node: header not at hand
    if (listpack) {
        if (n < 6) return kUnknownLen;
        const uint32_t c = raw[4] | (raw[5] << 8);
        return c == 65535 ? kUnknownLen : c;
    }
    if (n < 10) return kUnknownLen;
    const uint32_t c = raw[8] | (raw[9] << 8);
    return c == 65535 ? kUnknownLen : c;
}

uint64_t add_len(uint64_t a, uint64_t b) { return (a == kUnknownLen || b == kUnknownLen) ? kUnknownLen : a + b; }

// A string node whose header we want even when it is LZF-compressed: decode into scratch.
uint64_t node_entries(Reader &r, bool listpack, std::string &scratch) {
    const uint8_t *raw = nullptr;
    uint64_t raw_len = 0;
    scratch.clear();
    const uint8_t *before = r.p;
    read_string(r, nullptr, &raw, &raw_len);
    if (!r.ok) return kUnknownLen;
    if (raw) return packed_entries(raw, raw_len, listpack);
    // compressed (or integer-encoded, which no list node is): read it again, decoded
    Reader again = {before, r.end, true};
    read_string(again, &scratch);
    if (!again.ok) { r.ok = false; return kUnknownLen; }
    return packed_entries(reinterpret_cast<const uint8_t *>(scratch.data()), scratch.size(), listpack);
}

// Skip one value of the given type; for list types return the number of elements through *llen.
bool skip_value(Reader &r, int type, uint64_t *llen, std::string &scratch) {
    *llen = 0;
    switch (type) {
        case 0: read_string(r, nullptr); break;                                      // string
        case 1: {                                                                    // list (linked)
            const uint64_t n = read_len(r, nullptr);
            for (uint64_t i = 0; i < n && r.ok; ++i) read_string(r, nullptr);
            *llen = n;
            break;
        }
        case 2: {                                                                    // set
            const uint64_t n = read_len(r, nullptr);
            for (uint64_t i = 0; i < n && r.ok; ++i) read_string(r, nullptr);
            break;
        }
        case 3: {                                                                    // zset, scores as text
            const uint64_t n = read_len(r, nullptr);
            for (uint64_t i = 0; i < n && r.ok; ++i) {
                read_string(r, nullptr);
                const uint8_t l = r.u8();
                if (l < 253) r.skip(l);
            }
            break;
        }
        case 4: {                                                                    // hash
            const uint64_t n = read_len(r, nullptr);
            for (uint64_t i = 0; i < 2 * n && r.ok; ++i) read_string(r, nullptr);
            break;
        }
        case 5: {                                                                    // zset, binary scores
            const uint64_t n = read_len(r, nullptr);
            for (uint64_t i = 0; i < n && r.ok; ++i) { read_string(r, nullptr); r.skip(8); }
            break;
        }
        case 7: {                                                                    // module value, tagged
            read_len(r, nullptr);                                                    // module id
            while (r.ok) {
                const uint64_t op = read_len(r, nullptr);
                if (op == 0) break;
                if (op == 1 || op == 2) read_len(r, nullptr);
                else if (op == 3) r.skip(4);
                else if (op == 4) r.skip(8);
                else if (op == 5) read_string(r, nullptr);
                else r.ok = false;
            }
            break;
        }
        case 9: case 11: case 12: case 13: case 16: case 17: case 20:               // one packed blob
            read_string(r, nullptr);
            break;
        case 10: *llen = node_entries(r, false, scratch); break;                      // list as one ziplist
        case 14: {                                                                   // quicklist of ziplists
            const uint64_t n = read_len(r, nullptr);
            uint64_t total = 0;
            for (uint64_t i = 0; i < n && r.ok; ++i) total = add_len(total, node_entries(r, false, scratch));
            *llen = total;
            break;
        }
        case 18: {                                                                   // quicklist of listpacks
            const uint64_t n = read_len(r, nullptr);
            uint64_t total = 0;
            for (uint64_t i = 0; i < n && r.ok; ++i) {
                const uint64_t container = read_len(r, nullptr);
                if (container == 1) { read_string(r, nullptr); total = add_len(total, 1); }      // one plain element
                else if (container == 2) total = add_len(total, node_entries(r, true, scratch));
                else r.ok = false;
            }
            *llen = total;
            break;
        }
        case 15: case 19: case 21: {                                                 // stream
            const uint64_t packs = read_len(r, nullptr);
            for (uint64_t i = 0; i < packs && r.ok; ++i) { read_string(r, nullptr); read_string(r, nullptr); }
            for (int i = 0; i < 3; ++i) read_len(r, nullptr);                        // length, last id
            if (type >= 19) for (int i = 0; i < 5; ++i) read_len(r, nullptr);        // first id, max deleted id, added
            const uint64_t groups = read_len(r, nullptr);
            for (uint64_t g = 0; g < groups && r.ok; ++g) {
                read_string(r, nullptr);
                read_len(r, nullptr); read_len(r, nullptr);
                if (type >= 19) read_len(r, nullptr);
                const uint64_t pel = read_len(r, nullptr);
                for (uint64_t i = 0; i < pel && r.ok; ++i) { r.skip(16 + 8); read_len(r, nullptr); }
                const uint64_t consumers = read_len(r, nullptr);
                for (uint64_t c = 0; c < consumers && r.ok; ++c) {
                    read_string(r, nullptr);
                    r.skip(8);
                    if (type >= 21) r.skip(8);
                    const uint64_t cpel = read_len(r, nullptr);
                    r.skip(16 * cpel);
                }
            }
            break;
        }
        default:
            return false;                                                            // unknown type
    }
    return r.ok;
}

}  // namespace

}  // namespace kas

using namespace kas;

extern "C" {

int kas_rdb_keys_host(const uint8_t *rdb, uint64_t rdb_len, uint8_t *out_blob, uint64_t blob_cap,
                      uint64_t *out_off, uint64_t off_cap, uint8_t *out_type, uint32_t *out_db,
                      uint64_t *out_list_len, uint64_t *n_keys, uint64_t *blob_len) {
    if (!rdb || !n_keys || !blob_len) return KAS_ERR_INVALID;
    if (rdb_len < 9 || memcmp(rdb, "REDIS", 5) != 0) return KAS_ERR_INVALID;
    for (int i = 5; i < 9; ++i)
        if (rdb[i] < '0' || rdb[i] > '9') return KAS_ERR_INVALID;
    Reader r = {rdb + 9, rdb + rdb_len, true};
    uint64_t n = 0, bytes = 0;
    uint32_t db = 0;
    std::string key, scratch;
    bool eof = false;
    while (r.ok && !eof) {
        int type = r.u8();
        if (!r.ok) break;
        switch (type) {
            case 0xFF: eof = true; continue;
            case 0xFE: db = (uint32_t)read_len(r, nullptr); continue;
            case 0xFB: read_len(r, nullptr); read_len(r, nullptr); continue;
            case 0xFA: read_string(r, nullptr); read_string(r, nullptr); continue;
            case 0xFD: r.skip(4); continue;                   // the value type follows as the next byte
            case 0xFC: r.skip(8); continue;
            case 0xF9: r.skip(1); continue;
            case 0xF8: read_len(r, nullptr); continue;
            case 0xF5: read_string(r, nullptr); continue;     // function library source
            case 0xF4: read_len(r, nullptr); read_len(r, nullptr); read_len(r, nullptr); continue;   // slot info
            default: break;
        }
        if (type >= 0xF0) return KAS_ERR_UNSUPPORTED;          // an opcode this reader does not know
        key.clear();
        read_string(r, &key);
        if (!r.ok) break;
        uint64_t llen = 0;
        if (!skip_value(r, type, &llen, scratch)) {
            if (r.ok) return KAS_ERR_UNSUPPORTED;              // a value type this reader does not know
            break;
        }
        if (out_off) {
            if (n + 1 >= off_cap) return KAS_ERR_TOO_SMALL;
            out_off[n] = bytes;
        }
        if (out_blob) {
            if (bytes + key.size() > blob_cap) return KAS_ERR_TOO_SMALL;
            memcpy(out_blob + bytes, key.data(), key.size());
        }
        if (out_type) out_type[n] = (uint8_t)type;
        if (out_db) out_db[n] = db;
        if (out_list_len) out_list_len[n] = llen;
        bytes += key.size();
        ++n;
    }
    if (!r.ok || !eof) return KAS_ERR_INVALID;                 // truncated or malformed
    if (out_off) {
        if (n >= off_cap) return KAS_ERR_TOO_SMALL;
        out_off[n] = bytes;
    }
    *n_keys = n;
    *blob_len = bytes;
    return KAS_OK;
}

}  // extern "C"
