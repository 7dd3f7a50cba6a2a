// decode_json.cu — batched JsonDeviceRequestDecoder.decode
// (reference: service-event-sources/src/main/java/com/sitewhere/sources/decoder/json/
//  JsonDeviceRequestDecoder.java:48-62 and JsonDeviceRequestMarshaler.java:57-160; Jackson 2.9 semantics:
//  strict tokenizer, last-wins duplicate keys, FAIL_ON_UNKNOWN_PROPERTIES, scalar coercions, StdDateFormat,
//  NumberInput.parseDouble / Double.parseDouble).
//
// Same tiling as the protobuf decoder: a CTA stages the contiguous bytes of its 128 records into shared
// memory with coalesced 16-byte streaming loads; each thread then makes ONE structural pass over its
// record (iterative, bit-stack for nesting, no recursion) that validates the grammar and records, for
// the four envelope keys and the fourteen request properties, the LAST occurrence's value token
// (kind, span, escape flags).  After the pass the envelope rules and the bean rules are applied to
// those slots, numbers go through dparse.cuh (Clinger / Eisel-Lemire / exact fallback), dates through
// the StdDateFormat rules.  Before that general pass, a record is tried against the exact text the reference's own
// serialiser writes (json_template_record): same row, an order of magnitude fewer instructions.  Rows are placed by
// the optimistic dense pass + ordered redo, as in the protobuf decoder.
#include "dparse.cuh"
#include "merge_table.cuh"
#include "swgpu_common.cuh"

#define JS_THREADS 128
#define JS_WARPS (JS_THREADS / 32)
#define JS_AHEAD 4096         // template pass: warp tiles between an L2 prefetch and its use (about one wave of resident warps)
#define JS_SLACK 64            // bytes after a warp's stage: the template's word-wise reads run up to 36 bytes past a record's end
#define JS_WSTAGE (8 * 1024)  // bytes staged per warp tile: 32 records of up to 256 B on average (more CTAs per SM than 12 KiB)
#define JS_MAX_DEPTH 256

struct JsArgs {
    const uint8_t *bytes;
    const uint64_t *rec_off;
    uint32_t n;
    uint32_t max_rows;
    swgpu_decoded_cols cols;
    uint8_t *status;
    uint32_t *first_row;
    uint32_t *out_rows;
    uint32_t *ticket;
    uint64_t *tile_state;
    uint32_t rec_base;
    const uint32_t *row_base;
    uint32_t *row_next;
    uint32_t *not_dense;
    uint32_t force_ordered;  // small batches: only the ordered pass is launched and it decodes the batch
    // fused decode + merge (json_decode_merge_kernel): the state table, per-RECORD event ids, sequence base
    MgTable mt;
    const uint64_t *id_hi, *id_lo;
    uint64_t seq_base;
    uint32_t fast;  // option "json.fast_path": try the template recogniser first (default 1)
    uint32_t *pending;  // template pass -> dense pass: non-zero when the template pass left records
    uint32_t *hits; // option "json.stats": records the template recogniser took (NULL = not counted)
};

enum { JK_NONE = 0, JK_OBJECT, JK_ARRAY, JK_STRING, JK_INT, JK_FLOAT, JK_TRUE, JK_FALSE, JK_NULL };
#define JS_ST_PENDING 0xffu  // internal record status between the template pass and the dense pass
#define JI_ESCAPED 0x10u
#define JI_NONASCII 0x20u
#define JI_BADCHILD 0x40u  // metadata object holds a container value
enum {
    S_TYPE = 0, S_REQUEST, S_TOKEN, S_ORIG,
    S_ALT, S_ETYPE, S_EDATE, S_UPD, S_META, S_NAME, S_VALUE, S_LAT, S_LON, S_ELEV, S_SOURCE, S_LEVEL, S_ATYPE, S_MSG,
    S_COUNT
};
struct JSlot {
    uint32_t off, len, info;
};

template <typename B>
struct JsonScan {
    B b;
    uint32_t n, pos;
    bool esc_key, req_unknown, root_object;

    __device__ __forceinline__ void skip_ws() {
        while (pos < n) {
            uint32_t c = b.byte(pos);
            if (c > ' ') break;  // the usual case (compact JSON): one comparison
            if (c == ' ' || c == '\t' || c == '\n' || c == '\r')
                pos++;
            else
                break;
        }
    }
    static __device__ __forceinline__ bool is_hex(uint32_t c) {
        return (c - '0') <= 9u || ((c | 0x20) - 'a') <= 5u;
    }
    // string body after the opening quote; false = JsonParseException
    __device__ bool scan_string(uint32_t &off, uint32_t &len, uint32_t &info) {
        const uint32_t start = pos;
        info = 0;
        for (;;) {
            // eight bytes at a time while they are plain string content: stop at the first byte that is a
            // quote, a backslash, a control character or non-ASCII (exact per byte: the low seven bits are
            // tested with carry-free adds, the high bit separately), and let the byte-wise rules below take it
            while (pos + 8 <= n) {
                const uint64_t w = b.le64(pos);
                const uint64_t l = w & 0x7f7f7f7f7f7f7f7full;
                const uint64_t q = l ^ 0x2222222222222222ull, bs = l ^ 0x5c5c5c5c5c5c5c5cull;
                const uint64_t m = (w | ~(l + 0x6060606060606060ull) | ~(q + 0x7f7f7f7f7f7f7f7full) |
                                    ~(bs + 0x7f7f7f7f7f7f7f7full)) &
                                   0x8080808080808080ull;
                if (m == 0) {
                    pos += 8;
                    continue;
                }
                pos += (uint32_t)(__ffsll((long long)m) - 1) >> 3;
                break;
            }
            if (pos >= n) return false;
            uint32_t c = b.byte(pos);
            if (c == '"') break;
            if (c < 0x20) return false;
            if (c == '\\') {
                info |= JI_ESCAPED;
                pos++;
                if (pos >= n) return false;
                uint32_t e = b.byte(pos);
                if (e == 'u') {
                    if (pos + 4 >= n) return false;
                    if (!is_hex(b.byte(pos + 1)) || !is_hex(b.byte(pos + 2)) || !is_hex(b.byte(pos + 3)) ||
                        !is_hex(b.byte(pos + 4)))
                        return false;
                    pos += 5;
                } else if (e == '"' || e == '\\' || e == '/' || e == 'b' || e == 'f' || e == 'n' || e == 'r' || e == 't') {
                    pos++;
                } else
                    return false;
                continue;
            }
            if (c < 0x80) {
                pos++;
                continue;
            }
            info |= JI_NONASCII;
            uint32_t need;
            if ((c & 0xE0) == 0xC0)
                need = 1;
            else if ((c & 0xF0) == 0xE0)
                need = 2;
            else if ((c & 0xF8) == 0xF0)
                need = 3;
            else
                return false;
            pos++;
            for (uint32_t k = 0; k < need; k++) {
                if (pos >= n) return false;
                if ((b.byte(pos) & 0xC0) != 0x80) return false;
                pos++;
            }
        }
        off = start;
        len = pos - start;
        pos++;
        return true;
    }
    __device__ __forceinline__ bool digit_at(uint32_t p) const { return p < n && (b.byte(p) - '0') <= 9u; }
    // JSON number grammar; kind = JK_INT / JK_FLOAT
    __device__ bool scan_number(uint32_t &kind) {
        kind = JK_INT;
        if (b.byte(pos) == '-') {
            pos++;
            if (pos >= n) return false;
        }
        uint32_t c = b.byte(pos);
        if (c == '0') {
            pos++;
            if (digit_at(pos)) return false;  // leading zeroes not allowed
        } else if (c >= '1' && c <= '9') {
            while (digit_at(pos)) pos++;
        } else
            return false;
        if (pos < n && b.byte(pos) == '.') {
            kind = JK_FLOAT;
            pos++;
            if (!digit_at(pos)) return false;
            while (digit_at(pos)) pos++;
        }
        if (pos < n && ((b.byte(pos) | 0x20) == 'e')) {
            kind = JK_FLOAT;
            pos++;
            if (pos < n && (b.byte(pos) == '+' || b.byte(pos) == '-')) pos++;
            if (!digit_at(pos)) return false;
            while (digit_at(pos)) pos++;
        }
        return true;
    }
    __device__ bool scan_literal(const char *lit, uint32_t l) {
        if (pos + l > n) return false;
        for (uint32_t k = 0; k < l; k++)
            if (b.byte(pos + k) != (uint32_t)(uint8_t)lit[k]) return false;
        pos += l;
        if (pos < n) {  // _checkMatchEnd
            uint32_t c = b.byte(pos);
            bool ident = (c - '0') <= 9u || ((c | 0x20) - 'a') <= 25u || c == '_' || c == '$' || c >= 0x80 || c < 0x09 ||
                         (c >= 0x0e && c <= 0x1b) || c == 0x7f;
            if (ident) return false;
        }
        return true;
    }
    // bytes [off, off+l) == lit (l <= 16; the caller has scanned the key, so off + l <= n).  One or two
    // 8-byte reads against constants instead of a byte loop; byte-wise only for a short key in the last
    // bytes of the record.
    static __device__ __forceinline__ uint64_t lit64(const char *lit, uint32_t l) {
        uint64_t v = 0;
#pragma unroll
        for (uint32_t k = 0; k < 8; k++)
            if (k < l) v |= (uint64_t)(uint8_t)lit[k] << (8 * k);
        return v;
    }
    __device__ __forceinline__ bool key_is(uint32_t off, const char *lit, uint32_t l) const {
        if (l <= 8) {
            if (off + 8 <= n) {
                const uint64_t mask = l == 8 ? ~0ull : ((1ull << (8 * (l & 7))) - 1ull);
                return (b.le64(off) & mask) == lit64(lit, l);
            }
        } else {
            return b.le64(off) == lit64(lit, 8) && b.le64(off + l - 8) == lit64(lit + (l - 8), 8);
        }
        for (uint32_t k = 0; k < l; k++)
            if (b.byte(off + k) != (uint32_t)(uint8_t)lit[k]) return false;
        return true;
    }
    __device__ int envelope_key(uint32_t off, uint32_t len) const {
        switch (len) {
        case 4: return key_is(off, "type", 4) ? S_TYPE : -1;
        case 7: return key_is(off, "request", 7) ? S_REQUEST : -1;
        case 10: return key_is(off, "originator", 10) ? S_ORIG : -1;
        case 11: return key_is(off, "deviceToken", 11) ? S_TOKEN : -1;
        }
        return -1;
    }
    __device__ int request_key(uint32_t off, uint32_t len) const {
        switch (len) {
        case 4:
            if (key_is(off, "name", 4)) return S_NAME;
            if (key_is(off, "type", 4)) return S_ATYPE;
            return -1;
        case 5:
            if (key_is(off, "value", 5)) return S_VALUE;
            if (key_is(off, "level", 5)) return S_LEVEL;
            return -1;
        case 6: return key_is(off, "source", 6) ? S_SOURCE : -1;
        case 7: return key_is(off, "message", 7) ? S_MSG : -1;
        case 8:
            if (key_is(off, "latitude", 8)) return S_LAT;
            if (key_is(off, "metadata", 8)) return S_META;
            return -1;
        case 9:
            if (key_is(off, "eventDate", 9)) return S_EDATE;
            if (key_is(off, "eventType", 9)) return S_ETYPE;
            if (key_is(off, "longitude", 9)) return S_LON;
            if (key_is(off, "elevation", 9)) return S_ELEV;
            return -1;
        case 11:
            if (key_is(off, "updateState", 11)) return S_UPD;
            if (key_is(off, "alternateId", 11)) return S_ALT;
            return -1;
        }
        return -1;
    }

    // One structural pass over the first JSON value.  0 = ok, 1 = JsonParseException, 2 = nesting too deep.
    __device__ int run(JSlot *slots) {
        uint32_t stk[JS_MAX_DEPTH / 32];
        uint32_t depth = 0;
        bool in_request = false, in_meta = false;
        int cur_slot = -1, pend1 = -1, pend2 = -1;
        esc_key = false;
        req_unknown = false;
        root_object = false;
#pragma unroll
        for (int k = 0; k < S_COUNT; k++) slots[k].info = JK_NONE;
        skip_ws();
        if (pos >= n) return 1;  // no content
        root_object = b.byte(pos) == '{';
        enum { ST_VALUE, ST_KEY, ST_DONE } st = ST_VALUE;
        for (;;) {
            if (st == ST_VALUE) {
                skip_ws();
                if (pos >= n) return 1;
                const uint32_t c = b.byte(pos);
                const int slot = cur_slot;
                cur_slot = -1;
                if (c == '{' || c == '[') {
                    if (depth + 1 > JS_MAX_DEPTH) return 2;
                    const bool obj = c == '{';
                    if (slot >= 0) {
                        slots[slot].off = pos;
                        slots[slot].len = 0;
                        slots[slot].info = obj ? JK_OBJECT : JK_ARRAY;
                        if (depth == 1)
                            pend1 = slot;
                        else
                            pend2 = slot;
                    }
                    if (depth == 1 && slot == S_REQUEST && obj) {
                        in_request = true;
                        req_unknown = false;
#pragma unroll
                        for (int k = S_ALT; k < S_COUNT; k++) slots[k].info = JK_NONE;
                    }
                    if (depth == 2 && in_request && slot == S_META && obj) in_meta = true;
                    if (depth == 3 && in_meta) slots[S_META].info |= JI_BADCHILD;
                    if (obj)
                        stk[depth >> 5] |= 1u << (depth & 31);
                    else
                        stk[depth >> 5] &= ~(1u << (depth & 31));
                    depth++;
                    pos++;
                    skip_ws();
                    if (pos < n && b.byte(pos) == (obj ? '}' : ']')) {
                        pos++;
                        depth--;
                        // empty container closes immediately
                        if (depth == 1) {
                            if (pend1 >= 0) {
                                slots[pend1].len = pos - slots[pend1].off;
                                pend1 = -1;
                            }
                            in_request = false;
                        } else if (depth == 2) {
                            if (pend2 >= 0) {
                                slots[pend2].len = pos - slots[pend2].off;
                                pend2 = -1;
                            }
                            in_meta = false;
                        }
                        st = ST_DONE;
                    } else
                        st = obj ? ST_KEY : ST_VALUE;
                    continue;
                }
                uint32_t off = pos, len = 0, info = 0;
                if (c == '"') {
                    pos++;
                    if (!scan_string(off, len, info)) return 1;
                    info |= JK_STRING;
                } else if (c == '-' || (c - '0') <= 9u) {
                    uint32_t kind;
                    if (!scan_number(kind)) return 1;
                    len = pos - off;
                    info = kind;
                    if (depth == 0 && pos < n) {  // _verifyRootSpace
                        uint32_t d = b.byte(pos);
                        if (!(d == ' ' || d == '\t' || d == '\n' || d == '\r')) return 1;
                    }
                } else if (c == 't') {
                    if (!scan_literal("true", 4)) return 1;
                    len = 4;
                    info = JK_TRUE;
                } else if (c == 'f') {
                    if (!scan_literal("false", 5)) return 1;
                    len = 5;
                    info = JK_FALSE;
                } else if (c == 'n') {
                    if (!scan_literal("null", 4)) return 1;
                    len = 4;
                    info = JK_NULL;
                } else
                    return 1;
                if (slot >= 0) {
                    slots[slot].off = off;
                    slots[slot].len = len;
                    slots[slot].info = info;
                }
                st = ST_DONE;
                continue;
            }
            if (st == ST_KEY) {
                skip_ws();
                if (pos >= n || b.byte(pos) != '"') return 1;
                pos++;
                uint32_t koff, klen, kinfo;
                if (!scan_string(koff, klen, kinfo)) return 1;
                if ((kinfo & JI_ESCAPED) && depth <= 2) esc_key = true;
                if (depth == 1)
                    cur_slot = envelope_key(koff, klen);
                else if (depth == 2 && in_request) {
                    cur_slot = request_key(koff, klen);
                    if (cur_slot < 0) req_unknown = true;
                } else
                    cur_slot = -1;
                skip_ws();
                if (pos >= n || b.byte(pos) != ':') return 1;
                pos++;
                st = ST_VALUE;
                continue;
            }
            // ST_DONE: a value at nesting `depth` is complete
            if (depth == 0) return 0;
            skip_ws();
            if (pos >= n) return 1;
            const uint32_t c = b.byte(pos);
            const bool obj = (stk[(depth - 1) >> 5] >> ((depth - 1) & 31)) & 1u;
            if (c == ',') {
                pos++;
                if (obj)
                    st = ST_KEY;
                else {
                    skip_ws();
                    if (pos < n && b.byte(pos) == ']') return 1;  // trailing comma
                    st = ST_VALUE;
                }
                continue;
            }
            if (c == (obj ? '}' : ']')) {
                pos++;
                depth--;
                if (depth == 1) {
                    if (pend1 >= 0) {
                        slots[pend1].len = pos - slots[pend1].off;
                        pend1 = -1;
                    }
                    in_request = false;
                } else if (depth == 2) {
                    if (pend2 >= 0) {
                        slots[pend2].len = pos - slots[pend2].off;
                        pend2 = -1;
                    }
                    in_meta = false;
                }
                continue;  // stay in ST_DONE for the enclosing container
            }
            return 1;
        }
    }
};

// ---- value conversions (Jackson StdDeserializer rules) --------------------------------------------------------------
template <typename B>
__device__ __forceinline__ void js_trim(const B &b, uint32_t &off, uint32_t &len) {
    while (len && b.byte(off) <= 0x20) {
        off++;
        len--;
    }
    while (len && b.byte(off + len - 1) <= 0x20) len--;
}
template <typename B>
__device__ bool js_eq(const B &b, uint32_t off, uint32_t len, const char *lit, uint32_t l) {
    if (len != l) return false;
    for (uint32_t k = 0; k < l; k++)
        if (b.byte(off + k) != (uint32_t)(uint8_t)lit[k]) return false;
    return true;
}
// integer token: low 64 bits (two's complement) and whether every digit is zero
template <typename B>
__device__ uint64_t js_int_low64(const B &b, uint32_t off, uint32_t len, bool &all_zero) {
    bool neg = b.byte(off) == '-';
    uint64_t v = 0;
    all_zero = true;
    for (uint32_t i = neg ? 1 : 0; i < len; i++) {
        uint32_t d = b.byte(off + i) - '0';
        if (d) all_zero = false;
        v = v * 10 + d;
    }
    return neg ? (uint64_t)0 - v : v;
}
// number TOKEN -> double bits (NumberInput.parseDouble for floats; node.doubleValue() for integers)
template <typename B>
__device__ uint64_t js_token_double(const B &b, const JSlot &s) {
    const uint32_t kind = s.info & 0xf;
    if (kind == JK_FLOAT && js_eq(b, s.off, s.len, "2.2250738585072012e-308", 23)) return 1ull;  // Double.MIN_VALUE
    uint64_t bits = sw_parse_double_bits(b, s.off, s.len);
    if (kind == JK_INT && (bits << 1) == 0) bits = 0;  // the integer 0, however written, is +0.0
    return bits;
}
// rc: 0 ok, 1 decode error, 2 pass-through; isnull for "" / "null" / JSON null
template <typename B>
__device__ int js_to_double(const B &b, const JSlot &s, uint64_t &bits, bool &isnull) {
    isnull = false;
    bits = 0;
    const uint32_t kind = s.info & 0xf;
    if (kind == JK_INT || kind == JK_FLOAT) {
        bits = js_token_double(b, s);
        return 0;
    }
    if (kind == JK_NULL) {
        isnull = true;
        return 0;
    }
    if (kind != JK_STRING) return 1;
    if (s.info & JI_ESCAPED) return 2;
    uint32_t off = s.off, len = s.len;
    js_trim(b, off, len);
    if (len == 0 || js_eq(b, off, len, "null", 4)) {
        isnull = true;
        return 0;
    }
    if (js_eq(b, off, len, "Infinity", 8) || js_eq(b, off, len, "INF", 3)) {
        bits = 0x7ff0000000000000ull;
        return 0;
    }
    if (js_eq(b, off, len, "-Infinity", 9) || js_eq(b, off, len, "-INF", 4)) {
        bits = 0xfff0000000000000ull;
        return 0;
    }
    if (js_eq(b, off, len, "NaN", 3)) {
        bits = 0x7ff8000000000000ull;
        return 0;
    }
    // plain decimal grammar of Double.parseDouble
    uint32_t i = 0, nd = 0;
    uint32_t c = b.byte(off);
    if (c == '+' || c == '-') i++;
    while (i < len && (b.byte(off + i) - '0') <= 9u) {
        i++;
        nd++;
    }
    if (i < len && b.byte(off + i) == '.') {
        i++;
        while (i < len && (b.byte(off + i) - '0') <= 9u) {
            i++;
            nd++;
        }
    }
    if (nd == 0) return 2;
    if (i < len && ((b.byte(off + i) | 0x20) == 'e')) {
        i++;
        if (i < len && (b.byte(off + i) == '+' || b.byte(off + i) == '-')) i++;
        uint32_t ne = 0;
        while (i < len && (b.byte(off + i) - '0') <= 9u) {
            i++;
            ne++;
        }
        if (ne == 0) return 2;
    }
    if (i != len) return 2;
    bits = sw_parse_double_bits(b, off, len);
    return 0;
}

__device__ __forceinline__ int64_t js_days_from_civil(int64_t y, int64_t m, int64_t d) {
    y -= m <= 2;
    const int64_t era = (y >= 0 ? y : y - 399) / 400;
    const int64_t yoe = y - era * 400;
    const int64_t doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
    const int64_t doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
    return era * 146097 + doe - 719468;
}
__device__ __forceinline__ int64_t js_lenient_ms(int year, int month0, int day, int hour, int mi, int sec, int ms) {
    const int64_t y = year + (month0 >= 0 ? month0 / 12 : -((11 - month0) / 12));
    const int64_t m = ((month0 % 12) + 12) % 12;
    const int64_t days = js_days_from_civil(y, m + 1, 1) + (day - 1);
    return ((days * 24 + hour) * 60 + mi) * 60000 + (int64_t)sec * 1000 + ms;
}
// StdDateFormat on a trimmed ASCII string.  rc: 0 ok, 1 error, 2 pass-through
template <typename B>
__device__ int js_parse_date(const B &b, uint32_t off, uint32_t len, int64_t &out) {
    auto ch = [&](uint32_t i) -> uint32_t { return b.byte(off + i); };
    auto isd = [&](uint32_t i) -> bool { return (b.byte(off + i) - '0') <= 9u; };
    auto d2 = [&](uint32_t i) -> int { return (int)((ch(i) - '0') * 10 + (ch(i + 1) - '0')); };
    if (len >= 7 && isd(0) && isd(3) && ch(4) == '-' && isd(5)) {  // looksLikeISO8601
        if (len <= 10) {
            if (len == 10 && isd(0) && isd(1) && isd(2) && isd(3) && ch(4) == '-' && isd(5) && isd(6) && ch(7) == '-' &&
                isd(8) && isd(9)) {
                const int year = d2(0) * 100 + d2(2);
                if (year < 1583) return 2;
                out = js_lenient_ms(year, d2(5) - 1, d2(8), 0, 0, 0, 0);
                return 0;
            }
            return 1;
        }
        if (!(len >= 16 && isd(0) && isd(1) && isd(2) && isd(3) && ch(4) == '-' && isd(5) && isd(6) && ch(7) == '-' &&
              isd(8) && isd(9) && ch(10) == 'T' && isd(11) && isd(12) && ch(13) == ':' && isd(14) && isd(15)))
            return 1;
        uint32_t i = 16;
        int seconds = 0, msecs = 0;
        if (i < len && ch(i) == ':') {
            if (i + 3 > len || !isd(i + 1) || !isd(i + 2)) return 1;
            seconds = d2(i + 1);
            i += 3;
        }
        if (i < len && ch(i) == '.') {
            uint32_t j = i + 1;
            while (j < len && isd(j)) j++;
            const uint32_t fl = j - (i + 1);
            if (fl == 0 || fl > 9) return 1;
            msecs += 100 * (int)(ch(i + 1) - '0');
            if (fl >= 2) msecs += 10 * (int)(ch(i + 2) - '0');
            if (fl >= 3) msecs += (int)(ch(i + 3) - '0');
            i = j;
        }
        int64_t offset_ms = 0;
        if (i < len) {
            const uint32_t z = ch(i);
            if (z == 'Z') {
                i++;
            } else if (z == '+' || z == '-') {
                if (i + 3 > len || !isd(i + 1) || !isd(i + 2)) return 1;
                const int hh = d2(i + 1);
                int mm = 0;
                i += 3;
                if (i < len) {
                    if (ch(i) == ':') i++;
                    if (i + 2 > len || !isd(i) || !isd(i + 1)) return 1;
                    mm = d2(i);
                    i += 2;
                }
                offset_ms = ((int64_t)hh * 3600 + (int64_t)mm * 60) * 1000;
                if (z == '-') offset_ms = -offset_ms;
            } else
                return 1;
        }
        if (i != len) return 1;
        const int year = d2(0) * 100 + d2(2);
        if (year < 1583) return 2;
        out = js_lenient_ms(year, d2(5) - 1, d2(8), d2(11), d2(14), seconds, msecs) - offset_ms;
        return 0;
    }
    // stringified timestamp
    bool all = len > 0;
    for (uint32_t k = 0; k < len && all; k++) {
        if (isd(k)) continue;
        if (k == 0 && ch(k) == '-') continue;
        all = false;
    }
    if (!all) return 2;  // RFC-1123 parser territory
    const bool neg = ch(0) == '-';
    const uint32_t nd = len - (neg ? 1 : 0);
    if (neg) {
        if (nd == 0 || nd > 19) return 1;
    } else {
        if (nd > 19) return 2;
    }
    // up to 19 digits fit 64 bits unsigned
    uint64_t v = 0;
    for (uint32_t k = neg ? 1 : 0; k < len; k++) v = v * 10 + (ch(k) - '0');
    if (neg) {
        if (v > 9223372036854775808ull) return 1;
        out = (int64_t)((uint64_t)0 - v);
    } else {
        if (v > 9223372036854775807ull) return 2;
        out = (int64_t)v;
    }
    return 0;
}

// enum property.  rc 0 ok / 1 error / 2 pass-through
template <typename B>
__device__ int js_to_enum(const B &b, const JSlot &s, int which /*0 eventType 1 source 2 level*/, uint32_t &ord, bool &isnull) {
    isnull = false;
    ord = 0;
    const uint32_t kind = s.info & 0xf;
    const uint32_t count = which == 0 ? 6u : (which == 1 ? 2u : 4u);
    if (kind == JK_NULL) {
        isnull = true;
        return 0;
    }
    if (kind == JK_STRING) {
        if (s.info & JI_ESCAPED) return 2;
        const uint32_t o = s.off, l = s.len;
        if (which == 0) {
            if (js_eq(b, o, l, "Measurement", 11)) { ord = 0; return 0; }
            if (js_eq(b, o, l, "Location", 8)) { ord = 1; return 0; }
            if (js_eq(b, o, l, "Alert", 5)) { ord = 2; return 0; }
            if (js_eq(b, o, l, "CommandInvocation", 17)) { ord = 3; return 0; }
            if (js_eq(b, o, l, "CommandResponse", 15)) { ord = 4; return 0; }
            if (js_eq(b, o, l, "StateChange", 11)) { ord = 5; return 0; }
        } else if (which == 1) {
            if (js_eq(b, o, l, "Device", 6)) { ord = 0; return 0; }
            if (js_eq(b, o, l, "System", 6)) { ord = 1; return 0; }
        } else {
            if (js_eq(b, o, l, "Info", 4)) { ord = 0; return 0; }
            if (js_eq(b, o, l, "Warning", 7)) { ord = 1; return 0; }
            if (js_eq(b, o, l, "Error", 5)) { ord = 2; return 0; }
            if (js_eq(b, o, l, "Critical", 8)) { ord = 3; return 0; }
        }
        uint32_t off = o, len = l;
        js_trim(b, off, len);
        if (len > 0 && (b.byte(off) - '0') <= 9u) {
            uint64_t v = 0;
            bool ok = true;
            for (uint32_t k = 0; k < len; k++) {
                uint32_t d = b.byte(off + k) - '0';
                if (d > 9u) {
                    ok = false;
                    break;
                }
                v = v * 10 + d;
                if (v > 2147483647ull) {
                    ok = false;
                    break;
                }
            }
            if (ok && v < count) {
                ord = (uint32_t)v;
                return 0;
            }
        }
        return 1;
    }
    if (kind == JK_INT) {
        bool z;
        const int32_t idx = (int32_t)(uint32_t)js_int_low64(b, s.off, s.len, z);
        if (idx >= 0 && (uint32_t)idx < count) {
            ord = (uint32_t)idx;
            return 0;
        }
        return 1;
    }
    return 1;
}

template <typename B>
__device__ int js_to_bool(const B &b, const JSlot &s, bool &val) {
    val = false;
    const uint32_t kind = s.info & 0xf;
    switch (kind) {
    case JK_TRUE: val = true; return 0;
    case JK_FALSE:
    case JK_NULL: return 0;
    case JK_INT: {
        bool z;
        js_int_low64(b, s.off, s.len, z);
        val = !z;
        return 0;
    }
    case JK_STRING: {
        if (s.info & JI_ESCAPED) return 2;
        uint32_t off = s.off, len = s.len;
        js_trim(b, off, len);
        if (js_eq(b, off, len, "true", 4) || js_eq(b, off, len, "True", 4)) {
            val = true;
            return 0;
        }
        if (js_eq(b, off, len, "false", 5) || js_eq(b, off, len, "False", 5)) return 0;
        if (len == 0 || js_eq(b, off, len, "null", 4)) return 0;
        return 1;
    }
    }
    return 1;
}

struct JsRow {
    uint32_t status;
    uint64_t meta, th, nh, tspan, nspan, ospan, mspan, mdspan, aspan;
    int64_t date;
    uint64_t v0, v1, v2;
};

// ---- template recogniser -------------------------------------------------------------------------------------------------
// The request as the reference itself serialises it: DeviceRequest{deviceToken, originator, type, request} and the
// request POJO written by a default Jackson ObjectMapper, properties in declaration order, compact, no escapes
// (sitewhere-core-api/.../rest/model/device/communication/DeviceRequest.java:15-88,
//  .../event/request/Device{Event,Measurement,Location,Alert}CreateRequest.java; dates per JsonDateSerializer.java:48-51):
//   {"deviceToken":"T","originator":null,"type":"DeviceMeasurement","request":{"eventType":"Measurement",
//    "eventDate":"yyyy-MM-ddTHH:mm:ss.SSSZ","updateState":B,"name":"N","value":X}}
//   ... "type":"DeviceLocation","request":{"eventType":"Location", ... ,"latitude":X,"longitude":X,"elevation":X}}
//   ... "type":"DeviceAlert","request":{"eventType":"Alert", ... ,"source":"S","level":"L","type":"A","message":"M"}}
// A record of exactly this shape gives the same row through the general pass (json_decode_record below); this path
// gets there in a few hundred instructions: the fixed text is compared four bytes at a time against constants, the
// variable strings are closed by a four-byte scan that keeps their words for the hash, the date and the number are
// converted while they are read.  Anything else — another key order, white space, an escape, a byte >= 0x80, an
// optional property, a number outside the plain decimal form — returns false and the general pass decodes the
// record from the start, so this path never decides an error.  Shared-memory records only (reads run up to seven
// bytes past the record's end, into the tile's slack).
struct JsWin {  // sequential little-endian 4-byte reads at any alignment
    const uint32_t *wp;
    uint32_t sh, cur;
    __device__ __forceinline__ void seek(const uint8_t *p) {
        const uintptr_t a = (uintptr_t)p;
        wp = (const uint32_t *)(a & ~(uintptr_t)3);
        sh = (uint32_t)(a & 3) * 8;
        cur = *wp;
    }
    __device__ __forceinline__ uint32_t next() {
        const uint32_t nx = *++wp;
        const uint32_t v = __funnelshift_r(cur, nx, sh);
        cur = nx;
        return v;
    }
};
static constexpr __host__ __device__ uint32_t js_lit32(const char *lit, int k, int n) {
    uint32_t v = 0;
    for (int j = 0; j < 4 && k + j < n; j++) v |= (uint32_t)(uint8_t)lit[k + j] << (8 * j);
    return v;
}
// bytes [p, p+N) == lit ?  (N = sizeof(lit) - 1)
template <int NP1>
__device__ __forceinline__ bool js_expect(const uint8_t *p, const char (&lit)[NP1]) {
    constexpr int N = NP1 - 1;
    JsWin w;
    w.seek(p);
    uint32_t acc = 0;
#pragma unroll
    for (int k = 0; k + 4 <= N; k += 4) acc |= w.next() ^ js_lit32(lit, k, N);
    if (N & 3) acc |= (w.next() ^ js_lit32(lit, N & ~3, N)) & ((1u << (8 * (N & 3))) - 1u);
    return acc == 0;
}
__device__ __forceinline__ uint32_t js_stop32(uint32_t v) {  // 0x80 in every byte that ends plain string content
    const uint32_t l = v & 0x7f7f7f7fu;
    return (v | ~(l + 0x60606060u) | ~((l ^ 0x22222222u) + 0x7f7f7f7fu) | ~((l ^ 0x5c5c5c5cu) + 0x7f7f7f7fu)) & 0x80808080u;
}
// plain string body at p (just after the opening quote) of at most 4*W - 1 bytes: its length; false = longer, not
// plain, or not closed
template <int W>
__device__ __forceinline__ bool js_plain_string(const uint8_t *p, uint32_t &len) {
    JsWin w;
    w.seek(p);
#pragma unroll 1
    for (uint32_t i = 0; i < (uint32_t)W; i++) {
        const uint32_t v = w.next();
        const uint32_t m = js_stop32(v);
        if (m) {
            const uint32_t k = (uint32_t)(__ffs((int)m) - 1) >> 3;
            len = 4 * i + k;
            return ((v >> (8 * k)) & 0xffu) == '"';
        }
    }
    return false;
}
// sw_hash64 of the len bytes at p (same word order as the byte-stream form); reads up to 3 bytes past them
__device__ __forceinline__ uint64_t js_hash_at(const uint8_t *p, uint32_t len) {
    const uint32_t m = 0x5bd1e995u;
    uint32_t h = 0x9747b28cu ^ len, f = 0x811c9dc5u;
    const uint32_t nw = len >> 2, rem = len & 3;
    JsWin w;
    w.seek(p);
#pragma unroll 1
    for (uint32_t i = 0; i < nw; i++) {
        uint32_t k = w.next();
        f = sw_hash_hi_step(f, k);
        k *= m;
        k ^= k >> 24;
        k *= m;
        h *= m;
        h ^= k;
    }
    uint32_t tail = 0;
    if (rem) {
        tail = w.next() & ((1u << (8 * rem)) - 1u);
        h ^= tail;
        h *= m;
    }
    h ^= h >> 13;
    h *= m;
    h ^= h >> 15;
    return ((uint64_t)sw_hash_hi_final(f, tail, len) << 32) | h;
}
// "yyyy-MM-ddTHH:mm:ss.SSSZ" at p -> ms since the epoch (StdDateFormat's lenient field arithmetic, as js_parse_date);
// false = another form (the general pass handles those), or a year the general pass hands back
__device__ __forceinline__ bool js_date24(const uint8_t *p, int64_t &ms) {
    JsWin w;
    w.seek(p);
    uint32_t v[6];
#pragma unroll
    for (int i = 0; i < 6; i++) v[i] = w.next();
    // separators: - at 4, 7; T at 10; : at 13, 16; . at 19; Z at 23
    const bool sep = ((v[1] & 0xffu) == '-') && (((v[1] >> 24) & 0xffu) == '-') && (((v[2] >> 16) & 0xffu) == 'T') &&
                     (((v[3] >> 8) & 0xffu) == ':') && ((v[4] & 0xffu) == ':') && (((v[4] >> 24) & 0xffu) == '.') &&
                     (((v[5] >> 24) & 0xffu) == 'Z');
    // digits everywhere else: byte - '0' <= 9 for each of the 17 digit positions
    auto dg = [&](int pos) -> uint32_t { return ((v[pos >> 2] >> (8 * (pos & 3))) & 0xffu) - '0'; };
    uint32_t bad = 0;
    const int dpos[17] = {0, 1, 2, 3, 5, 6, 8, 9, 11, 12, 14, 15, 17, 18, 20, 21, 22};
    uint32_t d[17];
#pragma unroll
    for (int i = 0; i < 17; i++) {
        d[i] = dg(dpos[i]);
        bad |= d[i] > 9u;
    }
    if (!sep || bad) return false;
    const int year = (int)(d[0] * 1000 + d[1] * 100 + d[2] * 10 + d[3]);
    if (year < 1583) return false;
    ms = js_lenient_ms(year, (int)(d[4] * 10 + d[5]) - 1, (int)(d[6] * 10 + d[7]), (int)(d[8] * 10 + d[9]), (int)(d[10] * 10 + d[11]),
                       (int)(d[12] * 10 + d[13]), (int)(d[14] * 100 + d[15] * 10 + d[16]));
    return true;
}
// JSON number at b[pos..) in the plain decimal form  -?(0|[1-9][0-9]*)(\.[0-9]+)?  with at most 19 significant digits
// (all of them fit one 64-bit integer w, so the value is exactly w / 10^frac): up to 2^53 and 10^22 one exact
// integer, one exact power of ten and one correctly rounded division give the double Double.parseDouble gives;
// beyond that the same Eisel-Lemire step as the general conversion (sw_parse_double_bits) when it can decide.
// Jackson's integer rule: an integer token that is zero is +0.0.  false = any other form, or undecided: the caller
// takes the general conversion.  `end` = first byte after the token.
__constant__ double js_p10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                  1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
__device__ __noinline__ bool js_eisel(uint64_t w, int q, uint64_t *bits) {
    uint64_t v = 0;
    const bool ok = sw_eisel_lemire(w, (int64_t)q, v);
    *bits = v;
    return ok;
}
template <typename B>
__device__ __forceinline__ bool js_plain_number(const B &b, uint32_t n, uint32_t pos, uint32_t &end, uint64_t &bits) {
    const bool neg = pos < n && b.byte(pos) == '-';
    uint32_t i = pos + (neg ? 1u : 0u);
    if (i >= n) return false;
    uint64_t w = 0;
    uint32_t nd = 0, frac = 0;
    uint32_t c = b.byte(i);
    if (c == '0') {
        i++;
        if (i < n && (b.byte(i) - '0') <= 9u) return false;  // leading zero: a grammar error, not ours to report
    } else if ((c - '1') <= 8u) {
        while (i < n && (c = b.byte(i) - '0') <= 9u) {
            w = w * 10 + c;
            nd++;
            i++;
        }
    } else
        return false;
    bool is_float = false;
    if (i < n && b.byte(i) == '.') {
        is_float = true;
        i++;
        const uint32_t f0 = i;
        while (i < n && (c = b.byte(i) - '0') <= 9u) {
            w = w * 10 + c;
            nd += (w != 0);
            i++;
        }
        frac = i - f0;
        if (frac == 0) return false;
    }
    if (i < n && ((b.byte(i) | 0x20) == 'e')) return false;  // exponent form: general conversion
    if (nd > 19 || frac > 300) return false;                  // (w may have wrapped)
    end = i;
    const uint64_t sign = neg ? 0x8000000000000000ull : 0ull;
    if (w == 0) {
        bits = is_float ? sign : 0ull;  // -0.0 keeps its sign, the integer -0 is +0.0
        return true;
    }
    if (w <= (1ull << 53) && frac <= 22) {
        double v = (double)w;
        if (frac) v = __ddiv_rn(v, js_p10[frac]);
        bits = (uint64_t)__double_as_longlong(v) | sign;
        return true;
    }
    uint64_t v;
    if (!js_eisel(w, -(int)frac, &v)) return false;
    bits = v | sign;
    return true;
}

// Any other JSON number token (17 digits, exponent form ...): the general pass's grammar (JsonScan::scan_number) and
// conversion (js_token_double).  Out of line: one lane in a hundred gets here.
__device__ __noinline__ bool js_number_general(const uint8_t *s, uint32_t n, uint32_t pos, uint32_t *end, uint64_t *bits) {
    JsonScan<SmemBytes> sc;
    sc.b = SmemBytes{s};
    sc.n = n;
    sc.pos = pos;
    uint32_t kind;
    if (pos >= n || !sc.scan_number(kind)) return false;
    JSlot t;
    t.off = pos;
    t.len = sc.pos - pos;
    t.info = kind;
    *bits = js_token_double(sc.b, t);
    *end = sc.pos;
    return true;
}
__device__ __forceinline__ bool js_number(const SmemBytes &b, uint32_t n, uint32_t pos, uint32_t &end, uint64_t &bits) {
    if (js_plain_number(b, n, pos, end, bits)) return true;
    uint32_t e = 0;
    uint64_t v = 0;
    if (!js_number_general(b.s, n, pos, &e, &v)) return false;
    end = e;
    bits = v;
    return true;
}

template <typename B> struct JsTemplateOk { static constexpr bool value = false; };
template <> struct JsTemplateOk<SmemBytes> { static constexpr bool value = true; };

// advance over a literal, or leave the record to the general pass
#define JS_LIT(lit) ((pos + (uint32_t)sizeof(lit) - 1 <= n && js_expect(s + pos, lit)) ? (pos += (uint32_t)sizeof(lit) - 1, true) : false)
__device__ __forceinline__ bool json_template_record(const SmemBytes &bytes, uint32_t n, uint32_t rec, JsRow &r) {
    const uint8_t *s = bytes.s;
    uint32_t pos = 0;
    if (!JS_LIT("{\"deviceToken\":\"")) return false;
    uint32_t tl;
    if (!js_plain_string<8>(s + pos, tl) || pos + tl >= n) return false;
    const uint32_t tok_off = pos;
    pos += tl;
    if (!JS_LIT("\",\"originator\":null,\"type\":\"Device")) return false;
    // the event type decides the rest
    uint32_t kind;
    if (JS_LIT("Measurement\",\"request\":{\"eventType\":\"Measurement\",\"eventDate\":\""))
        kind = SWGPU_KIND_MEASUREMENT;
    else if (JS_LIT("Location\",\"request\":{\"eventType\":\"Location\",\"eventDate\":\""))
        kind = SWGPU_KIND_LOCATION;
    else if (JS_LIT("Alert\",\"request\":{\"eventType\":\"Alert\",\"eventDate\":\""))
        kind = SWGPU_KIND_ALERT;
    else
        return false;
    if (pos + 24 > n) return false;
    int64_t date;
    if (!js_date24(s + pos, date)) return false;
    pos += 24;
    if (!JS_LIT("\",\"updateState\":")) return false;
    uint32_t flags = SWGPU_F_HAS_DATE;
    if (JS_LIT("false,\"")) {
    } else if (JS_LIT("true,\"")) {
        flags |= SWGPU_F_UPDATE_STATE;
    } else
        return false;
    uint64_t v0 = 0, v1 = 0, v2 = 0, nh = 0, nspan = 0, mspan = 0;
    uint32_t level = 0, source = 0;
    if (kind == SWGPU_KIND_MEASUREMENT) {
        if (!JS_LIT("name\":\"")) return false;
        uint32_t nl;
        if (!js_plain_string<8>(s + pos, nl) || pos + nl >= n) return false;
        nspan = sw_mk_span(pos, nl);
        nh = js_hash_at(s + pos, nl);
        pos += nl;
        if (!JS_LIT("\",\"value\":")) return false;
        uint32_t end;
        if (!js_number(bytes, n, pos, end, v0)) return false;
        pos = end;
        flags |= SWGPU_F_HAS_NAME | SWGPU_F_HAS_V0;
    } else if (kind == SWGPU_KIND_LOCATION) {
        uint32_t end;
        if (!JS_LIT("latitude\":")) return false;
        if (!js_number(bytes, n, pos, end, v0)) return false;
        pos = end;
        if (!JS_LIT(",\"longitude\":")) return false;
        if (!js_number(bytes, n, pos, end, v1)) return false;
        pos = end;
        if (!JS_LIT(",\"elevation\":")) return false;
        if (!js_number(bytes, n, pos, end, v2)) return false;
        pos = end;
        flags |= SWGPU_F_HAS_V0 | SWGPU_F_HAS_V1 | SWGPU_F_HAS_V2;
    } else {
        if (!JS_LIT("source\":\"")) return false;
        if (JS_LIT("Device\",\"level\":\""))
            source = 0;
        else if (JS_LIT("System\",\"level\":\""))
            source = 1;
        else
            return false;
        if (JS_LIT("Info\",\""))
            level = 0;
        else if (JS_LIT("Warning\",\""))
            level = 1;
        else if (JS_LIT("Error\",\""))
            level = 2;
        else if (JS_LIT("Critical\",\""))
            level = 3;
        else
            return false;
        if (!JS_LIT("type\":\"")) return false;
        uint32_t nl;
        if (!js_plain_string<8>(s + pos, nl) || pos + nl >= n) return false;
        nspan = sw_mk_span(pos, nl);
        nh = js_hash_at(s + pos, nl);
        pos += nl;
        if (!JS_LIT("\",\"message\":\"")) return false;
        // the message may be long: scan it four bytes at a time without keeping it
        JsWin w;
        w.seek(s + pos);
        uint32_t ml = 0;
        for (;;) {
            if (pos + ml >= n) return false;
            const uint32_t v = w.next(), m = js_stop32(v);
            if (m) {
                const uint32_t k = (uint32_t)(__ffs((int)m) - 1) >> 3;
                if (((v >> (8 * k)) & 0xffu) != '"') return false;
                ml += k;
                break;
            }
            ml += 4;
        }
        if (pos + ml >= n) return false;
        mspan = sw_mk_span(pos, ml);
        pos += ml + 1;
        flags |= SWGPU_F_HAS_NAME | SWGPU_F_HAS_MESSAGE;
    }
    // the two closing braces must end the record (what follows the first value is ignored by the reference too, but
    // then this is not the serialiser's output: leave it to the general pass)
    if (pos + 2 != n || s[pos] != '}' || s[pos + 1] != '}') return false;
    r.status = SWGPU_ST_OK;
    r.meta = sw_mk_meta(kind, flags, level, source, 1, kind, rec);
    r.th = js_hash_at(s + tok_off, tl);
    r.nh = nh;
    r.tspan = sw_mk_span(tok_off, tl);
    r.nspan = nspan;
    r.ospan = 0;
    r.mspan = mspan;
    r.mdspan = 0;
    r.aspan = 0;
    r.date = date;
    r.v0 = v0;
    r.v1 = v1;
    r.v2 = v2;
    return true;
}
#undef JS_LIT
template <typename B>
__device__ __forceinline__ bool json_template_try(const B &, uint32_t, uint32_t, JsRow &) { return false; }
template <>
__device__ __forceinline__ bool json_template_try<SmemBytes>(const SmemBytes &b, uint32_t n, uint32_t rec, JsRow &r) {
    return json_template_record(b, n, rec, r);
}
template <>  // (the caller asks for the template only for a record staged in shared memory)
__device__ __forceinline__ bool json_template_try<AnyBytes>(const AnyBytes &b, uint32_t n, uint32_t rec, JsRow &r) {
    return json_template_record(SmemBytes{b.s}, n, rec, r);
}

template <typename B, bool TPL = true>
__device__ void json_decode_record(B bytes, uint32_t n, uint32_t rec, JsRow &r, bool try_template = true,
                                   uint32_t *template_hits = nullptr) {
    if (TPL && try_template && json_template_try(bytes, n, rec, r)) {
        if (template_hits) atomicAdd(template_hits, 1u);
        return;
    }
    r.status = SWGPU_ST_DECODE_EXCEPTION;
    r.meta = r.th = r.nh = r.tspan = r.nspan = r.ospan = r.mspan = r.mdspan = r.aspan = 0;
    r.date = 0;
    r.v0 = r.v1 = r.v2 = 0;
    uint32_t start = 0;
    if (n >= 2) {  // ByteSourceJsonBootstrapper: UTF-16/32 by zero bytes or BOM
        const uint32_t b0 = bytes.byte(0), b1 = bytes.byte(1);
        if (b0 == 0 || b1 == 0 || (b0 == 0xFE && b1 == 0xFF) || (b0 == 0xFF && b1 == 0xFE)) {
            r.status = SWGPU_ST_PASSTHROUGH;
            return;
        }
        if (n >= 3 && b0 == 0xEF && b1 == 0xBB && bytes.byte(2) == 0xBF) start = 3;
    }
    JSlot slots[S_COUNT];
    JsonScan<B> sc;
    sc.b = bytes;
    sc.n = n;
    sc.pos = start;
    const int rc = sc.run(slots);
    if (rc == 2) {
        r.status = SWGPU_ST_PASSTHROUGH;
        return;
    }
    if (rc != 0 || !sc.root_object) return;
    if (sc.esc_key) {
        r.status = SWGPU_ST_PASSTHROUGH;
        return;
    }
    const JSlot &ty = slots[S_TYPE];
    if ((ty.info & 0xf) != JK_STRING) return;  // missing, or asText() of a non-string is no enum constant
    if (ty.info & JI_ESCAPED) {
        r.status = SWGPU_ST_PASSTHROUGH;
        return;
    }
    int t = -1;
    if (js_eq(bytes, ty.off, ty.len, "RegisterDevice", 14)) t = 0;
    else if (js_eq(bytes, ty.off, ty.len, "DeviceLocation", 14)) t = 1;
    else if (js_eq(bytes, ty.off, ty.len, "DeviceAlert", 11)) t = 2;
    else if (js_eq(bytes, ty.off, ty.len, "DeviceMeasurement", 17)) t = 3;
    else if (js_eq(bytes, ty.off, ty.len, "DeviceStream", 12)) t = 4;
    else if (js_eq(bytes, ty.off, ty.len, "DeviceStreamData", 16)) t = 5;
    else if (js_eq(bytes, ty.off, ty.len, "Acknowledge", 11)) t = 6;
    if (t < 0) return;
    const JSlot &rq = slots[S_REQUEST], &tk = slots[S_TOKEN], &og = slots[S_ORIG];
    if ((rq.info & 0xf) == JK_NONE) return;  // "Request is missing."
    if ((tk.info & 0xf) == JK_NONE) return;  // "Device token is missing."
    if (t != 1 && t != 2 && t != 3) {
        r.status = SWGPU_ST_PASSTHROUGH;
        return;
    }
    const uint32_t kind = t == 1 ? SWGPU_KIND_LOCATION : (t == 2 ? SWGPU_KIND_ALERT : SWGPU_KIND_MEASUREMENT);
    uint32_t flags = 0;
    bool pass = false, err = false;
    if ((tk.info & 0xf) == JK_STRING) {
        if (tk.info & JI_ESCAPED) pass = true;
        r.tspan = sw_mk_span(tk.off, tk.len);
        r.th = sw_hash64(bytes, tk.off, tk.len);
    } else
        flags |= SWGPU_F_NULL_TOKEN;
    if ((og.info & 0xf) == JK_STRING) {
        flags |= SWGPU_F_HAS_ORIGINATOR;
        if (og.info & JI_ESCAPED) flags |= SWGPU_F_STR_ESCAPED;
        r.ospan = sw_mk_span(og.off, og.len);
    }
    if ((rq.info & 0xf) == JK_NULL) {
        if (pass) {
            r.status = SWGPU_ST_PASSTHROUGH;
            return;
        }
        r.meta = sw_mk_meta(kind, flags | SWGPU_F_NULL_REQUEST, 0, 0, 0, 0, rec);
        r.status = SWGPU_ST_OK;
        return;
    }
    if ((rq.info & 0xf) != JK_OBJECT) return;
    if (sc.req_unknown) return;  // UnrecognizedPropertyException
    // properties that belong to another request class are unknown for this one
    for (int k = S_NAME; k < S_COUNT; k++) {
        if ((slots[k].info & 0xf) == JK_NONE) continue;
        bool mine;
        if (kind == SWGPU_KIND_MEASUREMENT)
            mine = k == S_NAME || k == S_VALUE;
        else if (kind == SWGPU_KIND_LOCATION)
            mine = k == S_LAT || k == S_LON || k == S_ELEV;
        else
            mine = k == S_SOURCE || k == S_LEVEL || k == S_ATYPE || k == S_MSG;
        if (!mine) return;
    }
    uint32_t has_etype = 1, etype = kind, level = 0, source = 0;
    // String-valued properties
    auto str_prop = [&](const JSlot &s, bool key_forming, uint32_t present_flag, uint64_t &span_out, uint64_t *hash_out) {
        const uint32_t k = s.info & 0xf;
        if (k == JK_NONE || k == JK_NULL) return;
        if (k == JK_OBJECT || k == JK_ARRAY) {
            err = true;
            return;
        }
        const bool coerced = k != JK_STRING, escaped = (s.info & JI_ESCAPED) != 0;
        if (key_forming) {
            if (coerced || escaped) pass = true;
        } else {
            if (coerced) flags |= SWGPU_F_NAME_COERCED;
            if (escaped) flags |= SWGPU_F_STR_ESCAPED;
        }
        flags |= present_flag;
        span_out = sw_mk_span(s.off, s.len);
        if (hash_out) *hash_out = sw_hash64(bytes, s.off, s.len);
    };
    str_prop(slots[S_ALT], false, SWGPU_F_HAS_ALT_ID, r.aspan, nullptr);
    {
        const JSlot &s = slots[S_ETYPE];
        if ((s.info & 0xf) != JK_NONE) {
            uint32_t ord;
            bool isnull;
            const int rc2 = js_to_enum(bytes, s, 0, ord, isnull);
            if (rc2 == 1) err = true;
            if (rc2 == 2) pass = true;
            if (rc2 == 0) {
                has_etype = isnull ? 0u : 1u;
                etype = isnull ? 0u : ord;
            }
        }
    }
    {
        const JSlot &s = slots[S_EDATE];
        const uint32_t k = s.info & 0xf;
        if (k == JK_NONE || k == JK_NULL) {
        } else if (k == JK_INT) {
            bool z;
            r.date = (int64_t)js_int_low64(bytes, s.off, s.len, z);
            flags |= SWGPU_F_HAS_DATE;
        } else if (k == JK_STRING) {
            if (s.info & (JI_ESCAPED | JI_NONASCII))
                pass = true;
            else {
                uint32_t off = s.off, len = s.len;
                js_trim(bytes, off, len);
                if (len == 0 || js_eq(bytes, off, len, "null", 4)) {
                } else {
                    int64_t ms = 0;
                    const int rc2 = js_parse_date(bytes, off, len, ms);
                    if (rc2 == 1) err = true;
                    if (rc2 == 2) pass = true;
                    if (rc2 == 0) {
                        r.date = ms;
                        flags |= SWGPU_F_HAS_DATE;
                    }
                }
            }
        } else
            err = true;
    }
    {
        const JSlot &s = slots[S_UPD];
        if ((s.info & 0xf) != JK_NONE) {
            bool v;
            const int rc2 = js_to_bool(bytes, s, v);
            if (rc2 == 1) err = true;
            if (rc2 == 2) pass = true;
            if (rc2 == 0 && v) flags |= SWGPU_F_UPDATE_STATE;
        }
    }
    {
        const JSlot &s = slots[S_META];
        const uint32_t k = s.info & 0xf;
        if (k == JK_NONE || k == JK_NULL) {
        } else if (k == JK_OBJECT) {
            if (s.info & JI_BADCHILD) err = true;
            flags |= SWGPU_F_HAS_METADATA;
            r.mdspan = sw_mk_span(s.off, s.len);
        } else
            err = true;
    }
    auto dbl_prop = [&](const JSlot &s, uint64_t &out, uint32_t has_flag, bool primitive) {
        if ((s.info & 0xf) == JK_NONE) return;
        uint64_t bits;
        bool isnull;
        const int rc2 = js_to_double(bytes, s, bits, isnull);
        if (rc2 == 1) err = true;
        if (rc2 == 2) pass = true;
        if (rc2 == 0) {
            out = isnull ? 0ull : bits;
            if (!primitive) flags = isnull ? (flags & ~has_flag) : (flags | has_flag);
        }
    };
    if (kind == SWGPU_KIND_MEASUREMENT) {
        str_prop(slots[S_NAME], true, SWGPU_F_HAS_NAME, r.nspan, &r.nh);
        dbl_prop(slots[S_VALUE], r.v0, SWGPU_F_HAS_V0, true);
        flags |= SWGPU_F_HAS_V0;
    } else if (kind == SWGPU_KIND_LOCATION) {
        dbl_prop(slots[S_LAT], r.v0, SWGPU_F_HAS_V0, false);
        dbl_prop(slots[S_LON], r.v1, SWGPU_F_HAS_V1, false);
        dbl_prop(slots[S_ELEV], r.v2, SWGPU_F_HAS_V2, false);
    } else {
        str_prop(slots[S_ATYPE], true, SWGPU_F_HAS_NAME, r.nspan, &r.nh);
        str_prop(slots[S_MSG], false, SWGPU_F_HAS_MESSAGE, r.mspan, nullptr);
        for (int which = 1; which <= 2; which++) {
            const JSlot &s = slots[which == 1 ? S_SOURCE : S_LEVEL];
            if ((s.info & 0xf) == JK_NONE) continue;
            uint32_t ord;
            bool isnull;
            const int rc2 = js_to_enum(bytes, s, which, ord, isnull);
            if (rc2 == 1) err = true;
            if (rc2 == 2) pass = true;
            if (rc2 == 0) {
                if (which == 1) {
                    if (isnull) flags |= SWGPU_F_NULL_SOURCE; else source = ord;
                } else {
                    if (isnull) flags |= SWGPU_F_NULL_LEVEL; else level = ord;
                }
            }
        }
    }
    if (err) return;
    if (pass) {
        r.status = SWGPU_ST_PASSTHROUGH;
        return;
    }
    r.meta = sw_mk_meta(kind, flags, level, source, has_etype, etype, rec);
    r.status = SWGPU_ST_OK;
}

__device__ __forceinline__ void js_write_row(const JsArgs &a, const uint64_t row, const JsRow &r) {
    a.cols.meta[row] = r.meta;
    a.cols.token_hash[row] = r.th;
    a.cols.name_hash[row] = r.nh;
    a.cols.event_date[row] = r.date;
    a.cols.v0[row] = __longlong_as_double((long long)r.v0);
    a.cols.v1[row] = __longlong_as_double((long long)r.v1);
    a.cols.v2[row] = __longlong_as_double((long long)r.v2);
    a.cols.token_span[row] = r.tspan;
    a.cols.name_span[row] = r.nspan;
    if (a.cols.originator_span) a.cols.originator_span[row] = r.ospan;
    if (a.cols.message_span) a.cols.message_span[row] = r.mspan;
    if (a.cols.metadata_span) a.cols.metadata_span[row] = r.mdspan;
    if (a.cols.alternate_id_span) a.cols.alternate_id_span[row] = r.aspan;
}

// one tile = JS_THREADS records, one per thread.  As in the protobuf decoder: each warp stages and decodes
// its own 32 records (no block barrier for staging).
template <bool DENSE>
__device__ __forceinline__ void json_decode_tile(const JsArgs &a, const uint32_t tile, uint8_t *smem, uint64_t *s_warp) {
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t rec0_64 = (uint64_t)tile * JS_THREADS + warp * 32;
    const uint32_t rec0 = rec0_64 < a.n ? (uint32_t)rec0_64 : a.n;
    const uint32_t nrec = min(32u, a.n - rec0);
    uint8_t *ws = smem + warp * (JS_WSTAGE + JS_SLACK);
    uint64_t r0 = 0, r1 = 0;
    if (lane < nrec) {
        r0 = a.rec_off[rec0 + lane];
        r1 = a.rec_off[rec0 + lane + 1];
    }
    // after the template pass (json_template_kernel) the dense pass decodes only the records that one left
    bool mine = lane < nrec;
    if (DENSE && a.fast) {
        mine = mine && a.status[rec0 + lane] == JS_ST_PENDING;
        if (!__any_sync(SW_FULL, mine)) return;
    }
    __syncwarp();  // (the dense pass walks tiles in a loop: every lane is done with the previous tile's bytes)
    const uint64_t b0 = __shfl_sync(SW_FULL, r0, 0);
    const uint64_t b1 = __shfl_sync(SW_FULL, r1, nrec ? nrec - 1 : 0);
    const uintptr_t addr = (uintptr_t)(a.bytes + b0);
    const uint32_t lead = (uint32_t)(addr & 15);
    const uint64_t span = (uint64_t)lead + (b1 - b0);
    const bool staged = span <= JS_WSTAGE;
    if (staged && nrec) {
        const uint4 *src = (const uint4 *)(addr - lead);
        uint4 *dst = (uint4 *)ws;
        const uint32_t nvec = (uint32_t)((span + 15) >> 4);
        for (uint32_t i = lane; i < nvec; i += 32) dst[i] = sw_ld_stream_u4(src + i);
    }
    __syncwarp();
    JsRow r;
    r.status = 0;
    uint32_t count = 0;
    if (mine) {
        const uint64_t l64 = r1 - r0;
        const uint32_t len = l64 > 0x7fffffffull ? 0x7fffffffu : (uint32_t)l64;
        // one copy of the parser for staged and unstaged records (generic loads): half the code of two
        const uint8_t *rp = staged ? ws + lead + (uint32_t)(r0 - b0) : a.bytes + r0;
        json_decode_record<AnyBytes, !DENSE>(AnyBytes{rp}, len, a.rec_base + rec0 + lane, r, a.fast != 0 && staged, a.hits);
        count = r.status == SWGPU_ST_OK ? 1u : 0u;
    }
    uint64_t tile_total = 0, row;
    if (DENSE) {
        row = (uint64_t)rec0 + lane;
        if (mine && count != 1) atomicOr(a.not_dense, 1u);
    } else {
        row = sw_block_ordered_offset<JS_THREADS>(count, a.tile_state, tile, s_warp, &tile_total);
    }
    const uint64_t row_base = a.row_base ? (uint64_t)__ldg(a.row_base) : 0ull;
    const uint32_t tid_rec = lane;  // record index within the tile
    if (mine) {
        a.status[rec0 + tid_rec] = (uint8_t)r.status;
        a.first_row[rec0 + tid_rec] = (uint32_t)(row_base + row);
        if (count && row < a.max_rows) js_write_row(a, row, r);
    }
    if ((uint64_t)tile * JS_THREADS + JS_THREADS >= a.n && tid == 0) {  // the tile holding the last record
        const uint64_t total = DENSE ? (uint64_t)a.n : row + tile_total;
        a.first_row[a.n] = (uint32_t)(row_base + total);
        *a.out_rows = (uint32_t)total;
        if (a.row_next) *a.row_next = (uint32_t)(row_base + total);
    }
}

// Template pass (option "json.fast_path"): row index = record index as in the dense pass; a record in the
// serialiser's layout is decoded here, every other one is marked JS_ST_PENDING for the dense pass that follows.
// Nothing of the general parser is in this kernel: it stays small enough for the instruction cache.
__global__ void __launch_bounds__(JS_THREADS) json_template_kernel(JsArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];  // (JS_WSTAGE + JS_SLACK) per warp of the CTA
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t rec0_64 = (uint64_t)blockIdx.x * blockDim.x + warp * 32;
    const uint32_t rec0 = rec0_64 < a.n ? (uint32_t)rec0_64 : a.n;
    const uint32_t nrec = min(32u, a.n - rec0);
    uint8_t *ws = smem + warp * (JS_WSTAGE + JS_SLACK);
    uint64_t r0 = 0, r1 = 0;
    if (lane < nrec) {
        r0 = a.rec_off[rec0 + lane];
        r1 = a.rec_off[rec0 + lane + 1];
    }
    // the bytes of the warp tile JS_AHEAD tiles further on are asked into L2 now (one bulk prefetch by lane 0 at
    // the end of this tile), so that the warp that gets there finds its offsets and its records in L2
    uint64_t p0 = 0, p1 = 0;
    const uint64_t prec = rec0_64 + (uint64_t)JS_AHEAD * 32;
    if (lane == 0 && prec < a.n) {
        p0 = a.rec_off[prec];
        p1 = a.rec_off[prec + 32 <= a.n ? prec + 32 : a.n];
    }
    const uint64_t b0 = __shfl_sync(SW_FULL, r0, 0);
    const uint64_t b1 = __shfl_sync(SW_FULL, r1, nrec ? nrec - 1 : 0);
    const uintptr_t addr = (uintptr_t)(a.bytes + b0);
    const uint32_t lead = (uint32_t)(addr & 15);
    const uint64_t span = (uint64_t)lead + (b1 - b0);
    const bool staged = span <= JS_WSTAGE;
    if (staged && nrec) {
        const uint4 *src = (const uint4 *)(addr - lead);
        uint4 *dst = (uint4 *)ws;
        const uint32_t nvec = (uint32_t)((span + 15) >> 4);
        for (uint32_t i = lane; i < nvec; i += 32) dst[i] = sw_ld_stream_u4(src + i);
    }
    __syncwarp();
    if (lane == 0 && p1 > p0) {
        const uintptr_t pa = (uintptr_t)(a.bytes + p0);
        const uint64_t pspan = (pa & 15) + (p1 - p0);
        if (pspan <= JS_WSTAGE)
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(pa & ~(uintptr_t)15),
                         "r"((uint32_t)((pspan + 15) & ~15ull)) : "memory");
    }
    const uint64_t row_base = a.row_base ? (uint64_t)__ldg(a.row_base) : 0ull;
    bool hit = false;
    if (lane < nrec) {
        JsRow r;
        if (staged) hit = json_template_record(SmemBytes{ws + lead + (uint32_t)(r0 - b0)}, (uint32_t)(r1 - r0), a.rec_base + rec0 + lane, r);
        const uint64_t row = (uint64_t)rec0 + lane;
        a.status[rec0 + lane] = hit ? (uint8_t)SWGPU_ST_OK : (uint8_t)JS_ST_PENDING;
        a.first_row[rec0 + lane] = (uint32_t)(row_base + row);
        if (hit && row < a.max_rows) js_write_row(a, row, r);
    }
    const uint32_t hm = __ballot_sync(SW_FULL, hit);
    if (lane == 0) {
        if (a.hits && hm) atomicAdd(a.hits, (uint32_t)__popc(hm));
        if ((uint32_t)__popc(hm) != nrec) atomicOr(a.pending, 1u);  // the dense pass has records to decode
    }
    if ((uint64_t)blockIdx.x * blockDim.x + blockDim.x >= a.n && tid == 0) {  // the tile holding the last record
        a.first_row[a.n] = (uint32_t)(row_base + a.n);
        *a.out_rows = a.n;
        if (a.row_next) *a.row_next = (uint32_t)(row_base + a.n);
    }
}

// DENSE = true: optimistic pass, row index = record index, no inter-tile dependency; raises *not_dense
// when some record does not decode to exactly one row.  DENSE = false: ordered redo (ticketed tiles and
// look-back, a few CTAs per SM taking tiles in turn), which returns at once unless *not_dense was raised.
// (Same scheme as pb_decode_kernel.)
template <bool DENSE>
__global__ void __launch_bounds__(JS_THREADS) json_decode_kernel(JsArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];  // JS_WARPS * (JS_WSTAGE + JS_SLACK)
    __shared__ uint64_t s_warp[JS_THREADS / 32 + 2];
    __shared__ uint32_t s_tile;
    if (DENSE) {
        if (a.fast && __ldg(a.pending) == 0) return;  // the template pass took every record
        for (uint32_t tile = blockIdx.x; tile < (a.n + JS_THREADS - 1) / JS_THREADS; tile += gridDim.x)
            json_decode_tile<true>(a, tile, smem, s_warp);
        return;
    }
    if (!a.force_ordered && __ldg(a.not_dense) == 0) return;
    const uint32_t ntiles = (uint32_t)(((uint64_t)a.n + JS_THREADS - 1) / JS_THREADS);
    for (;;) {
        if (threadIdx.x == 0) s_tile = atomicAdd(a.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= ntiles) return;
        json_decode_tile<false>(a, tile, smem, s_warp);
        __syncthreads();  // s_tile, s_warp and the staging buffers are reused
    }
}

// ---- fused decode + merge: the same staging and record decode, the row goes straight into the device-state
// table (merge_table.cuh) with sequence number seq_base + (record << SWGPU_SEQ_RECORD_SHIFT); no columns.
__global__ void __launch_bounds__(JS_THREADS) json_decode_merge_kernel(JsArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];  // JS_WARPS * (JS_WSTAGE + JS_SLACK)
    __shared__ MgLocalCnt s_cnt[JS_WARPS];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (lane == 0) s_cnt[warp].keys = s_cnt[warp].bytes = 0;
    // newest records first (CTA 0 takes the end of the batch): see pb_decode_dense_kernel<true>
    const uint64_t rec0_64 = (uint64_t)(gridDim.x - 1 - blockIdx.x) * JS_THREADS + (JS_WARPS - 1 - warp) * 32;
    const uint32_t rec0 = rec0_64 < a.n ? (uint32_t)rec0_64 : a.n;
    const uint32_t nrec = min(32u, a.n - rec0);
    uint8_t *ws = smem + warp * (JS_WSTAGE + JS_SLACK);
    uint64_t r0 = 0, r1 = 0;
    if (lane < nrec) {
        r0 = a.rec_off[rec0 + lane];
        r1 = a.rec_off[rec0 + lane + 1];
    }
    const uint64_t b0 = __shfl_sync(SW_FULL, r0, 0);
    const uint64_t b1 = __shfl_sync(SW_FULL, r1, nrec ? nrec - 1 : 0);
    const uintptr_t addr = (uintptr_t)(a.bytes + b0);
    const uint32_t lead = (uint32_t)(addr & 15);
    const uint64_t span = (uint64_t)lead + (b1 - b0);
    const bool staged = span <= JS_WSTAGE;
    if (staged && nrec) {
        const uint4 *src = (const uint4 *)(addr - lead);
        uint4 *dst = (uint4 *)ws;
        const uint32_t nvec = (uint32_t)((span + 15) >> 4);
        for (uint32_t i = lane; i < nvec; i += 32) dst[i] = sw_ld_stream_u4(src + i);
    }
    __syncwarp();
    uint32_t count = 0;
    if (lane < nrec) {
        const uint32_t rec = rec0 + lane;
        const uint64_t l64 = r1 - r0;
        const uint32_t len = l64 > 0x7fffffffull ? 0x7fffffffu : (uint32_t)l64;
        JsRow r;
        const uint8_t *rp = staged ? ws + lead + (uint32_t)(r0 - b0) : a.bytes + r0;
        json_decode_record(AnyBytes{rp}, len, a.rec_base + rec, r, a.fast != 0 && staged);
        a.status[rec] = (uint8_t)r.status;
        count = r.status == SWGPU_ST_OK ? 1u : 0u;
        const uint32_t kind = SWGPU_META_KIND(r.meta), fl = SWGPU_META_FLAGS(r.meta);
        if (count && !(fl & (SWGPU_F_NULL_TOKEN | SWGPU_F_NULL_REQUEST))) {
            const uint32_t tl = SWGPU_SPAN_LEN(r.tspan), nl = kind == SWGPU_KIND_LOCATION ? 0u : SWGPU_SPAN_LEN(r.nspan);
            if (tl > 0xffffu || nl > 0xffffu) {
                atomicAdd(&a.mt.ctr->key_too_long, 1ull);
            } else {
                const uint32_t kindbyte = mg_kindbyte(kind, fl);
                const uint64_t h = mg_mix(r.th, kind == SWGPU_KIND_LOCATION ? 0ull : r.nh, kindbyte);
                const uint64_t seq1 = a.seq_base + ((uint64_t)(a.rec_base + rec) << SWGPU_SEQ_RECORD_SHIFT) + 1;
                const uint64_t ihi = __ldg(a.id_hi + rec), ilo = __ldg(a.id_lo + rec);
                if (staged) {
                    MgKeyBytes<SmemBytes> key{SmemBytes{ws + lead + (uint32_t)(r0 - b0)}, SWGPU_SPAN_OFF(r.tspan), tl,
                                              SWGPU_SPAN_OFF(r.nspan), nl};
                    mg_upsert(a.mt, &s_cnt[warp], h, r.th, tl, nl, kindbyte, key, seq1, ihi, ilo, (uint64_t)r.date, r.v0);
                } else {
                    MgKeyBytes<GmemBytes> key{GmemBytes{a.bytes + r0}, SWGPU_SPAN_OFF(r.tspan), tl, SWGPU_SPAN_OFF(r.nspan), nl};
                    mg_upsert(a.mt, &s_cnt[warp], h, r.th, tl, nl, kindbyte, key, seq1, ihi, ilo, (uint64_t)r.date, r.v0);
                }
            }
        }
    }
    __syncwarp();
    count = __reduce_add_sync(SW_FULL, count);
    if (lane == 0) {
        if (count) atomicAdd(a.out_rows, count);
        mg_flush_counts(a.mt, &s_cnt[warp]);
    }
}

int sw_decode_merge_json_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, const MgTable &t,
                                uint64_t seq_base, const uint64_t *id_hi, const uint64_t *id_lo, uint8_t *status,
                                uint32_t *out_rows_dev) {
    JsArgs a;
    memset(&a, 0, sizeof a);
    a.bytes = bytes;
    a.rec_off = rec_off;
    a.n = n;
    a.status = status;
    a.out_rows = out_rows_dev;
    a.mt = t;
    a.id_hi = id_hi;
    a.id_lo = id_lo;
    a.seq_base = seq_base;
    a.fast = ctx->opt_json_fast;
    const int smem = JS_WARPS * (JS_WSTAGE + JS_SLACK);
    {
        static std::mutex mu;
        static bool done[64] = {false};
        std::lock_guard<std::mutex> lk(mu);
        const int d = ctx->device & 63;
        if (!done[d]) {
            SW_CUDA(ctx, cudaFuncSetAttribute(json_decode_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            done[d] = true;
        }
    }
    json_decode_merge_kernel<<<(n + JS_THREADS - 1) / JS_THREADS, JS_THREADS, smem, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}

int sw_stage_decode(swgpu_ctx *ctx, bool json, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                    uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status, uint32_t *first_row,
                    uint32_t *out_rows);

int sw_decode_json_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint32_t max_rows,
                          const swgpu_decoded_cols *out, uint8_t *status, uint32_t *first_row, uint32_t *out_rows_dev,
                          uint32_t rec_base, const uint32_t *row_base, uint32_t *row_next, bool small) {
    if (!ctx) return SWGPU_E_INVALID;
    if (!out || !out->meta || !out->token_hash || !out->name_hash || !out->event_date || !out->v0 || !out->v1 ||
        !out->v2 || !out->token_span || !out->name_span) {
        SW_SET_ERR(ctx, "decoded columns: the nine mandatory column pointers must be non-NULL");
        return SWGPU_E_INVALID;
    }
    if (!rec_off || !first_row || !out_rows_dev || (n && !status)) {
        SW_SET_ERR(ctx, "swgpu_decode_json_dev: NULL argument");
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    if (n == 0) {
        SW_CUDA(ctx, cudaMemsetAsync(first_row, 0, sizeof(uint32_t), ctx->stream));
        SW_CUDA(ctx, cudaMemsetAsync(out_rows_dev, 0, sizeof(uint32_t), ctx->stream));
        return SWGPU_OK;
    }
    const uint32_t ntiles = (n + JS_THREADS - 1) / JS_THREADS;
    const size_t sbytes = SW_SCRATCH_HDR + (size_t)ntiles * 8;  // one scan state per tile
    if (!small) {  // (small: the caller has sized and cleared the scratch; see sw_decode_pb_launch)
        if (int rc = sw_ensure_scratch(ctx, sbytes)) return rc;
        SW_CUDA(ctx, cudaMemsetAsync(ctx->scratch, 0, sbytes, ctx->stream));
    }
    JsArgs a;
    a.force_ordered = small ? 1u : 0u;
    a.bytes = bytes;
    a.rec_off = rec_off;
    a.n = n;
    a.max_rows = max_rows;
    a.cols = *out;
    a.status = status;
    a.first_row = first_row;
    a.out_rows = out_rows_dev;
    a.ticket = (uint32_t *)ctx->scratch;
    a.tile_state = (uint64_t *)((uint8_t *)ctx->scratch + SW_SCRATCH_HDR);
    a.rec_base = rec_base;
    a.row_base = row_base;
    a.row_next = row_next;
    a.not_dense = (uint32_t *)ctx->scratch + 1;
    a.fast = ctx->opt_json_fast;
    a.pending = (uint32_t *)ctx->scratch + 3;
    a.hits = ctx->opt_json_stats ? (uint32_t *)ctx->scratch + 2 : nullptr;
    const int smem = JS_WARPS * (JS_WSTAGE + JS_SLACK);
    {
        // per device, once: the attribute is a property of the loaded function
        static std::mutex mu;
        static bool done[64] = {false};
        std::lock_guard<std::mutex> lk(mu);
        const int d = ctx->device & 63;
        if (!done[d]) {
            SW_CUDA(ctx, cudaFuncSetAttribute(json_decode_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            SW_CUDA(ctx, cudaFuncSetAttribute(json_decode_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            SW_CUDA(ctx, cudaFuncSetAttribute(json_template_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            done[d] = true;
        }
    }
    const uint32_t nredo = ntiles < (uint32_t)ctx->sm_count * 4u ? ntiles : (uint32_t)ctx->sm_count * 4u;
    if (small) {
        json_decode_kernel<false><<<nredo, JS_THREADS, smem, ctx->stream>>>(a);  // the ordered pass decodes the batch
        SW_CUDA(ctx, cudaGetLastError());
        return SWGPU_OK;
    }
    if (a.fast) {
        // one warp per CTA: a CTA's shared memory goes back the moment its 32 records are done (measured: 27 resident
        // warps per SM against 17 with four-warp CTAs, 0.114 -> 0.105 ms per 1 M records)
        json_template_kernel<<<(n + 31) / 32, 32, JS_WSTAGE + JS_SLACK, ctx->stream>>>(a);  // the serialiser's layout
        SW_CUDA(ctx, cudaGetLastError());
    }
    // after the template pass the dense pass usually has little or nothing to do: a few CTAs per SM walk the tiles
    json_decode_kernel<true><<<a.fast ? nredo : ntiles, JS_THREADS, smem, ctx->stream>>>(a);  // optimistic: rows = records
    SW_CUDA(ctx, cudaGetLastError());
    json_decode_kernel<false><<<nredo, JS_THREADS, smem, ctx->stream>>>(a);  // exact redo iff not dense
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}

// Records of the last JSON decode the template recogniser took (option "json.stats" = 1; synchronises the stream).
extern "C" int32_t swgpu_json_stats(swgpu_ctx *ctx, uint64_t out[2]) {
    if (!ctx || !out) return SWGPU_E_INVALID;
    out[0] = out[1] = 0;
    if (!ctx->scratch) return SWGPU_OK;
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    uint32_t w[3];
    SW_CUDA(ctx, cudaMemcpyAsync(w, ctx->scratch, sizeof(w), cudaMemcpyDeviceToHost, ctx->stream));
    SW_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    out[0] = w[2];
    out[1] = w[1];
    return SWGPU_OK;
}

extern "C" int32_t swgpu_decode_json_dev(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                                         uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                                         uint32_t *first_row, uint32_t *out_rows_dev) {
    return sw_decode_json_launch(ctx, bytes, rec_off, n, max_rows, out, status, first_row, out_rows_dev, 0, nullptr,
                                 nullptr, false);
}

extern "C" int32_t swgpu_decode_json(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                                     uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                                     uint32_t *first_row, uint32_t *out_rows) {
    return sw_stage_decode(ctx, true, bytes, rec_off, n, max_rows, out, status, first_row, out_rows);
}

// Host helper: the same decimal -> double routine the kernel uses, for bindings that must coerce a
// numeric string themselves (and for CPU-side validation of the algorithm).
struct HostBytes {
    const uint8_t *s;
    uint32_t byte(uint32_t i) const { return s[i]; }
};
extern "C" int32_t swgpu_parse_double(const char *text, uint32_t len, double *out) {
    if (!text || !out) return SWGPU_E_INVALID;
    HostBytes hb{(const uint8_t *)text};
    *out = sw_bits2d(sw_parse_double_bits(hb, 0, len));
    return SWGPU_OK;
}
