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
// the StdDateFormat rules.  Rows are placed by the ordered single-pass scan.  HBM-bound by design;
// what limits it today is instruction issue (byte-at-a-time scanning), see DESIGN.md.
#include "dparse.cuh"
#include "merge_table.cuh"
#include "swgpu_common.cuh"

#define JS_THREADS 128
#define JS_WARPS (JS_THREADS / 32)
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
    // fused decode + merge (json_decode_merge_kernel): the state table, per-RECORD event ids, sequence base
    MgTable mt;
    const uint64_t *id_hi, *id_lo;
    uint64_t seq_base;
    uint32_t fast;  // option "json.fast_path": try the fast structural pass first (default 1)
};

enum { JK_NONE = 0, JK_OBJECT, JK_ARRAY, JK_STRING, JK_INT, JK_FLOAT, JK_TRUE, JK_FALSE, JK_NULL };
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

    // ---- fast structural pass for the common shape ---------------------------------------------------------------
    // A request as clients send it: a root object whose members are the four envelope keys (each at most once, any
    // order), scalar values, "request" an object of known request properties with scalar values; strings are plain
    // printable ASCII (no escape, no control character, no byte >= 0x80).  For such a record this fills `slots`
    // exactly as run() does, in straight-line code: no nesting stack, keys taken from two 8-byte loads and matched
    // as words, strings closed by an 8-byte scan.  ANYTHING else — an unknown or repeated key, a container value, an
    // escape, odd white space, a grammar error — returns 3 and the caller takes run() from the start, so this
    // path never decides an error.  Only for records staged in shared memory (8-byte reads may run past the
    // record's end, into the tile's slack).
    static __device__ __forceinline__ uint64_t stop_mask(uint64_t w) {  // 0x80 in every byte that ends plain string content
        const uint64_t l = w & 0x7f7f7f7f7f7f7f7full;
        const uint64_t q = l ^ 0x2222222222222222ull, bs = l ^ 0x5c5c5c5c5c5c5c5cull;
        return (w | ~(l + 0x6060606060606060ull) | ~(q + 0x7f7f7f7f7f7f7f7full) | ~(bs + 0x7f7f7f7f7f7f7f7full)) &
               0x8080808080808080ull;
    }
    static __device__ __forceinline__ uint64_t low_bytes(uint64_t w, uint32_t k) {  // the first k (< 8) bytes of w
        return w & ((1ull << (8 * k)) - 1ull);
    }
    static constexpr __device__ uint64_t lit8(const char *lit, uint32_t l) {
        uint64_t v = 0;
        for (uint32_t k = 0; k < 8 && k < l; k++) v |= (uint64_t)(uint8_t)lit[k] << (8 * k);
        return v;
    }
    // plain string body from pos (just after the opening quote): false = not plain / not closed inside the record
    __device__ __forceinline__ bool fast_string(JSlot &s) {
        const uint32_t start = pos;
        for (;;) {
            if (pos >= n) return false;
            const uint64_t m = stop_mask(b.le64(pos));
            if (m == 0) {
                pos += 8;
                continue;
            }
            pos += (uint32_t)(__ffsll((long long)m) - 1) >> 3;
            break;
        }
        if (pos >= n || b.byte(pos) != '"') return false;
        s.off = start;
        s.len = pos - start;
        s.info = JK_STRING;
        pos++;
        return true;
    }
    __device__ __forceinline__ bool fast_scalar(JSlot &s) {
        if (pos >= n) return false;
        const uint32_t c = b.byte(pos), off = pos;
        if (c == '"') {
            pos++;
            return fast_string(s);
        }
        uint32_t info;
        if (c == '-' || (c - '0') <= 9u) {
            if (!scan_number(info)) return false;
        } else if (c == 't') {
            if (!scan_literal("true", 4)) return false;
            info = JK_TRUE;
        } else if (c == 'f') {
            if (!scan_literal("false", 5)) return false;
            info = JK_FALSE;
        } else if (c == 'n') {
            if (!scan_literal("null", 4)) return false;
            info = JK_NULL;
        } else
            return false;
        s.off = off;
        s.len = pos - off;
        s.info = info;
        return true;
    }
    // key at pos (just after its opening quote): slot id, or -1; pos moves past the closing quote
    __device__ __forceinline__ int fast_key(bool in_request) {
        if (pos + 2 > n) return -1;
        const uint64_t k0 = b.le64(pos), k1 = b.le64(pos + 8);
        const uint64_t m0 = stop_mask(k0);
        uint32_t kl;
        if (m0)
            kl = (uint32_t)(__ffsll((long long)m0) - 1) >> 3;
        else {
            const uint64_t m1 = stop_mask(k1);
            if (!m1) return -1;  // longer than any key we know
            kl = 8 + ((uint32_t)(__ffsll((long long)m1) - 1) >> 3);
        }
        if (pos + kl >= n) return -1;
        if ((uint32_t)(((kl < 8 ? k0 : k1) >> (8 * (kl & 7))) & 0xff) != '"') return -1;  // stopped on something else
        int slot = -1;
        const uint64_t a = kl < 8 ? low_bytes(k0, kl) : k0, c = kl > 8 ? low_bytes(k1, kl - 8) : 0ull;
        if (!in_request) {
            switch (kl) {
            case 4: if (a == lit8("type", 4)) slot = S_TYPE; break;
            case 7: if (a == lit8("request", 7)) slot = S_REQUEST; break;
            case 10: if (a == lit8("originat", 8) && c == lit8("or", 2)) slot = S_ORIG; break;
            case 11: if (a == lit8("deviceTo", 8) && c == lit8("ken", 3)) slot = S_TOKEN; break;
            }
        } else {
            switch (kl) {
            case 4: slot = a == lit8("name", 4) ? S_NAME : a == lit8("type", 4) ? S_ATYPE : -1; break;
            case 5: slot = a == lit8("value", 5) ? S_VALUE : a == lit8("level", 5) ? S_LEVEL : -1; break;
            case 6: if (a == lit8("source", 6)) slot = S_SOURCE; break;
            case 7: if (a == lit8("message", 7)) slot = S_MSG; break;
            case 8: if (a == lit8("latitude", 8)) slot = S_LAT; break;  // ("metadata": an object, not for this path)
            case 9:
                if (c == lit8("e", 1))
                    slot = a == lit8("eventDat", 8) ? S_EDATE : a == lit8("eventTyp", 8) ? S_ETYPE : a == lit8("longitud", 8) ? S_LON : -1;
                else if (c == lit8("n", 1) && a == lit8("elevatio", 8))
                    slot = S_ELEV;
                break;
            case 11:
                if (a == lit8("updateSt", 8) && c == lit8("ate", 3)) slot = S_UPD;
                else if (a == lit8("alternat", 8) && c == lit8("eId", 3)) slot = S_ALT;
                break;
            }
        }
        pos += kl + 1;
        return slot;
    }
    __device__ int run_fast(JSlot *slots) {
        esc_key = false;
        req_unknown = false;
        root_object = true;
#pragma unroll
        for (int k = 0; k < S_COUNT; k++) slots[k].info = JK_NONE;
        skip_ws();
        if (pos >= n || b.byte(pos) != '{') return 3;
        pos++;
        uint32_t seen = 0;
        for (;;) {  // members of the root object
            skip_ws();
            if (pos >= n || b.byte(pos) != '"') return 3;
            pos++;
            const int slot = fast_key(false);
            if (slot < 0 || ((seen >> slot) & 1u)) return 3;
            seen |= 1u << slot;
            skip_ws();
            if (pos >= n || b.byte(pos) != ':') return 3;
            pos++;
            skip_ws();
            if (pos >= n) return 3;
            if (slot == S_REQUEST && b.byte(pos) == '{') {
                const uint32_t obj = pos;
                pos++;
                for (;;) {  // members of the request object
                    skip_ws();
                    if (pos >= n || b.byte(pos) != '"') return 3;
                    pos++;
                    const int rs = fast_key(true);
                    if (rs < 0 || ((seen >> rs) & 1u)) return 3;
                    seen |= 1u << rs;
                    skip_ws();
                    if (pos >= n || b.byte(pos) != ':') return 3;
                    pos++;
                    skip_ws();
                    if (!fast_scalar(slots[rs])) return 3;
                    skip_ws();
                    if (pos >= n) return 3;
                    const uint32_t d = b.byte(pos);
                    pos++;
                    if (d == ',') continue;
                    if (d == '}') break;
                    return 3;
                }
                slots[S_REQUEST].off = obj;
                slots[S_REQUEST].len = pos - obj;
                slots[S_REQUEST].info = JK_OBJECT;
            } else if (!fast_scalar(slots[slot]))
                return 3;
            skip_ws();
            if (pos >= n) return 3;
            const uint32_t d = b.byte(pos);
            pos++;
            if (d == ',') continue;
            if (d == '}') return 0;
            return 3;
        }
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

// the fast structural pass reads 8 bytes at a time past the record's end: only where the record sits in a staged tile
template <typename B> struct JsFastOk { static constexpr bool value = false; };
template <> struct JsFastOk<SmemBytes> { static constexpr bool value = true; };

template <typename B>
__device__ void json_decode_record(B bytes, uint32_t n, uint32_t rec, JsRow &r, bool FAST = true) {
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
    int rc = 3;
    if (FAST && JsFastOk<B>::value) rc = sc.run_fast(slots);
    if (rc == 3) {
        sc.pos = start;
        rc = sc.run(slots);
    }
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

// one tile = JS_THREADS records, one per thread.  As in the protobuf decoder: each warp stages and decodes
// its own 32 records (no block barrier for staging).
template <bool DENSE>
__device__ __forceinline__ void json_decode_tile(const JsArgs &a, const uint32_t tile, uint8_t *smem, uint64_t *s_warp) {
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t rec0_64 = (uint64_t)tile * JS_THREADS + warp * 32;
    const uint32_t rec0 = rec0_64 < a.n ? (uint32_t)rec0_64 : a.n;
    const uint32_t nrec = min(32u, a.n - rec0);
    uint8_t *ws = smem + warp * (JS_WSTAGE + 32);
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
    JsRow r;
    r.status = 0;
    uint32_t count = 0;
    if (lane < nrec) {
        const uint64_t l64 = r1 - r0;
        const uint32_t len = l64 > 0x7fffffffull ? 0x7fffffffu : (uint32_t)l64;
        if (staged)
            json_decode_record(SmemBytes{ws + lead + (uint32_t)(r0 - b0)}, len, a.rec_base + rec0 + lane, r, a.fast != 0);
        else
            json_decode_record(GmemBytes{a.bytes + r0}, len, a.rec_base + rec0 + lane, r, a.fast != 0);
        count = r.status == SWGPU_ST_OK ? 1u : 0u;
    }
    uint64_t tile_total = 0, row;
    if (DENSE) {
        row = (uint64_t)rec0 + lane;
        if (lane < nrec && count != 1) atomicOr(a.not_dense, 1u);
    } else {
        row = sw_block_ordered_offset<JS_THREADS>(count, a.tile_state, tile, s_warp, &tile_total);
    }
    const uint64_t row_base = a.row_base ? (uint64_t)__ldg(a.row_base) : 0ull;
    const uint32_t tid_rec = lane;  // record index within the tile
    if (lane < nrec) {
        a.status[rec0 + tid_rec] = (uint8_t)r.status;
        a.first_row[rec0 + tid_rec] = (uint32_t)(row_base + row);
        if (count && row < a.max_rows) {
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
    }
    if ((uint64_t)tile * JS_THREADS + JS_THREADS >= a.n && tid == 0) {  // the tile holding the last record
        const uint64_t total = DENSE ? (uint64_t)a.n : row + tile_total;
        a.first_row[a.n] = (uint32_t)(row_base + total);
        *a.out_rows = (uint32_t)total;
        if (a.row_next) *a.row_next = (uint32_t)(row_base + total);
    }
}

// DENSE = true: optimistic pass, row index = record index, no inter-tile dependency; raises *not_dense
// when some record does not decode to exactly one row.  DENSE = false: ordered redo (ticketed tiles and
// look-back, a few CTAs per SM taking tiles in turn), which returns at once unless *not_dense was raised.
// (Same scheme as pb_decode_kernel.)
template <bool DENSE>
__global__ void __launch_bounds__(JS_THREADS) json_decode_kernel(JsArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];  // JS_WARPS * (JS_WSTAGE + 32)
    __shared__ uint64_t s_warp[JS_THREADS / 32 + 2];
    __shared__ uint32_t s_tile;
    if (DENSE) {
        json_decode_tile<true>(a, blockIdx.x, smem, s_warp);
        return;
    }
    if (__ldg(a.not_dense) == 0) return;
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
    extern __shared__ __align__(16) uint8_t smem[];  // JS_WARPS * (JS_WSTAGE + 32)
    __shared__ MgLocalCnt s_cnt[JS_WARPS];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (lane == 0) s_cnt[warp].keys = s_cnt[warp].bytes = 0;
    // newest records first (CTA 0 takes the end of the batch): see pb_decode_dense_kernel<true>
    const uint64_t rec0_64 = (uint64_t)(gridDim.x - 1 - blockIdx.x) * JS_THREADS + (JS_WARPS - 1 - warp) * 32;
    const uint32_t rec0 = rec0_64 < a.n ? (uint32_t)rec0_64 : a.n;
    const uint32_t nrec = min(32u, a.n - rec0);
    uint8_t *ws = smem + warp * (JS_WSTAGE + 32);
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
        if (staged)
            json_decode_record(SmemBytes{ws + lead + (uint32_t)(r0 - b0)}, len, a.rec_base + rec, r, a.fast != 0);
        else
            json_decode_record(GmemBytes{a.bytes + r0}, len, a.rec_base + rec, r, a.fast != 0);
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
    const int smem = JS_WARPS * (JS_WSTAGE + 32);
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
                          uint32_t rec_base, const uint32_t *row_base, uint32_t *row_next) {
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
    if (int rc = sw_ensure_scratch(ctx, sbytes)) return rc;
    SW_CUDA(ctx, cudaMemsetAsync(ctx->scratch, 0, sbytes, ctx->stream));
    JsArgs a;
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
    const int smem = JS_WARPS * (JS_WSTAGE + 32);
    {
        // per device, once: the attribute is a property of the loaded function
        static std::mutex mu;
        static bool done[64] = {false};
        std::lock_guard<std::mutex> lk(mu);
        const int d = ctx->device & 63;
        if (!done[d]) {
            SW_CUDA(ctx, cudaFuncSetAttribute(json_decode_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            SW_CUDA(ctx, cudaFuncSetAttribute(json_decode_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            done[d] = true;
        }
    }
    json_decode_kernel<true><<<ntiles, JS_THREADS, smem, ctx->stream>>>(a);   // optimistic: rows = records
    SW_CUDA(ctx, cudaGetLastError());
    const uint32_t nredo = ntiles < (uint32_t)ctx->sm_count * 4u ? ntiles : (uint32_t)ctx->sm_count * 4u;
    json_decode_kernel<false><<<nredo, JS_THREADS, smem, ctx->stream>>>(a);  // exact redo iff not dense
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}

extern "C" int32_t swgpu_decode_json_dev(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                                         uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                                         uint32_t *first_row, uint32_t *out_rows_dev) {
    return sw_decode_json_launch(ctx, bytes, rec_off, n, max_rows, out, status, first_row, out_rows_dev, 0, nullptr,
                                 nullptr);
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
