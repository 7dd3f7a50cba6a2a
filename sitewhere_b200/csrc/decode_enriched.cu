// decode_enriched.cu — batched parse + conversion of the enriched-event payload that the rule-processing and
// device-state consumers read off Kafka ("inbound-enriched-events").
// Reference: EventModelMarshaler.parseEnrichedEventPayloadMessage = GEnrichedEventPayload.parseFrom(byte[])
// (sitewhere-grpc-event-management/src/main/java/com/sitewhere/grpc/client/event/EventModelMarshaler.java:204-206)
// followed by EventModelConverter.asApiEnrichedEventPayload (.../EventModelConverter.java:1926-1931 ->
// asApiDeviceEventContext :1753-1764, asApiGenericDeviceEvent :1664-1690, asApiDeviceMeasurement :463-469,
// asApiDeviceAlert :689-697, asApiDeviceLocation :841-848, copyApiDeviceEvent :329-343); schema
// sitewhere-grpc-event-management/src/main/proto/device-event-model.proto:139-215,288-296,321-324 and
// sitewhere-grpc-model/src/main/proto/sitewhere-common.proto:12-33,146-151 (proto3, protobuf-java 3.3.0).
// Call sites: DeviceStateProcessingLogic.processRecord (service-device-state/.../processing/
// DeviceStateProcessingLogic.java:93-108), KafkaRuleProcessorHost (service-rule-processing/.../kafka/
// KafkaRuleProcessorHost.java:170-197).
//
// Same tiling as the other decoders: a warp stages the contiguous bytes of its 32 records in shared memory and
// each lane walks one record.  A record yields one row (measurement / location / alert), or none: a parse or
// conversion exception (status 1: the consumers log and skip the record) or an event type off the path
// (status 3).  Rows go to the same columns as the device-payload decoders — the row's "token" is the 16 bytes of
// the device assignment id, which is the reference's device-state key — plus the ids the payload carries; or,
// fused, straight into the device-state table.  Rules of protobuf-java 3.3.0 (full runtime, proto3; recalled)
// restated here: a message-typed field that occurs again MERGES into the value held so far (generated code:
// toBuilder().mergeFrom(readMessage())) — proto3 scalars of the later occurrence replace only when they are not
// the default, sub-messages merge recursively, a different oneof member replaces and the same member merges,
// while scalars repeated inside one occurrence take the last value; strings and map<string,string> entries are
// read with readStringRequireUtf8 (invalid UTF-8 is an InvalidProtocolBufferException); an enum number outside
// the schema becomes UNRECOGNIZED, for which SiteWhere's converters throw; parseFrom(byte[]) pushes the array
// length as the outermost limit.
#include "merge_table.cuh"
#include "pb_reader.cuh"
#include "swgpu_common.cuh"

#define EN_THREADS 128
#define EN_WARPS (EN_THREADS / 32)
#define EN_WSTAGE (8 * 1024)  // bytes staged per warp tile: 32 records of up to 256 B on average

struct EnArgs {
    const uint8_t *bytes;
    const uint64_t *rec_off;
    uint32_t n;
    uint32_t max_rows;
    swgpu_decoded_cols cols;
    swgpu_event_ids ids;  // all NULL when the caller does not want them
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
    MgTable mt;  // fused decode + merge
    uint64_t seq_base;
    uint32_t fast;      // option "enriched.fast_path": the forward-pass kernel ran first, this pass takes what it left
    uint32_t *pending;  // forward pass -> general pass: non-zero when records were left
};
#define EN_ST_PENDING 0xffu  // internal record status between the two passes

// ---- message values (defaults: all zero) and their proto3 merge ----------------------------------------------------
struct EnUuid {
    uint32_t has;
    uint64_t msb, lsb;
};
struct EnStr {
    uint32_t has, off, len;
};
struct EnDbl {
    uint32_t has;
    uint64_t bits;
};
struct EnEvent {
    uint32_t has, etype, md_first, md_count;
    EnUuid id, device, assignment;  // (customerId, areaId, assetId are validated and dropped: not state-table fields)
    EnStr alt;
    uint64_t date;
};
struct EnAny {
    uint32_t ocase;  // 0 none, 1 measurement, 2 alert, 3 location, 4 another event type
    uint32_t name_off, name_len, msg_off, msg_len, source, level;
    uint64_t value_bits;
    EnDbl lat, lon, elev;
    EnEvent ev;
};
__device__ __forceinline__ bool en_dbl_zero(uint64_t bits) { return (bits << 1) == 0; }  // +0.0 and -0.0 == 0D
// a message-typed field: the first occurrence is taken as parsed, a later one merges
__device__ __forceinline__ void en_merge(EnUuid &a, const EnUuid &o) {
    if (!o.has) return;
    if (!a.has) {
        a = o;
        return;
    }
    if (o.msb) a.msb = o.msb;
    if (o.lsb) a.lsb = o.lsb;
}
__device__ __forceinline__ void en_merge(EnStr &a, const EnStr &o) {
    if (!o.has) return;
    if (!a.has) {
        a = o;
        return;
    }
    if (o.len) {
        a.off = o.off;
        a.len = o.len;
    }
}
__device__ __forceinline__ void en_merge(EnDbl &a, const EnDbl &o) {
    if (!o.has) return;
    if (!a.has) {
        a = o;
        return;
    }
    if (!en_dbl_zero(o.bits)) a.bits = o.bits;
}
__device__ void en_merge(EnEvent &a, const EnEvent &o) {
    if (!o.has) return;
    if (!a.has) {
        a = o;
        return;
    }
    en_merge(a.id, o.id);
    en_merge(a.alt, o.alt);
    if (o.etype) a.etype = o.etype;
    en_merge(a.device, o.device);
    en_merge(a.assignment, o.assignment);
    if (o.date) a.date = o.date;
    if (o.md_count) {
        if (!a.md_count) a.md_first = o.md_first;
        a.md_count += o.md_count;
    }
}
// the oneof member held in `a` meets a new value `t`: another member replaces, the same member merges
__device__ void en_merge_member(EnAny &a, const EnAny &t) {
    if (a.ocase != t.ocase) {
        a = t;
        return;
    }
    if (t.name_len) {
        a.name_off = t.name_off;
        a.name_len = t.name_len;
    }
    if (t.msg_len) {
        a.msg_off = t.msg_off;
        a.msg_len = t.msg_len;
    }
    if (!en_dbl_zero(t.value_bits)) a.value_bits = t.value_bits;
    if (t.source) a.source = t.source;
    if (t.level) a.level = t.level;
    en_merge(a.lat, t.lat);
    en_merge(a.lon, t.lon);
    en_merge(a.elev, t.elev);
    en_merge(a.ev, t.ev);
}

template <typename B>
struct EnParser {
    PbReader<B> r;

    __device__ __forceinline__ void pop(uint32_t old) {  // readMessage epilogue
        if (!r.err && r.last_tag != 0) r.err = true;     // checkLastTagWas(0)
        r.limit = old;
    }
    // com.google.protobuf.Utf8.isValidUtf8 over [off, off+len)
    __device__ bool utf8_valid(uint32_t off, uint32_t len) {
        uint32_t i = off;
        const uint32_t e = off + len;
        while (i < e) {
            const uint32_t b1 = r.b.byte(i++);
            if (b1 < 0x80) continue;
            if (b1 < 0xC2) return false;
            if (b1 < 0xE0) {
                if (i >= e || (r.b.byte(i) & 0xC0) != 0x80) return false;
                i += 1;
            } else if (b1 < 0xF0) {
                if (i + 1 >= e) return false;
                const uint32_t b2 = r.b.byte(i), b3 = r.b.byte(i + 1);
                if ((b2 & 0xC0) != 0x80 || (b3 & 0xC0) != 0x80) return false;
                if (b1 == 0xE0 && b2 < 0xA0) return false;
                if (b1 == 0xED && b2 >= 0xA0) return false;
                i += 2;
            } else {
                if (b1 > 0xF4 || i + 2 >= e) return false;
                const uint32_t b2 = r.b.byte(i), b3 = r.b.byte(i + 1), b4 = r.b.byte(i + 2);
                if ((b2 & 0xC0) != 0x80 || (b3 & 0xC0) != 0x80 || (b4 & 0xC0) != 0x80) return false;
                if (b1 == 0xF0 && b2 < 0x90) return false;
                if (b1 == 0xF4 && b2 > 0x8F) return false;
                i += 3;
            }
        }
        return true;
    }
    __device__ void string(uint32_t &off, uint32_t &len) {  // readStringRequireUtf8
        r.lenfield(off, len);
        if (r.err) return;
        if (!utf8_valid(off, len)) {
            r.err = true;
            off = len = 0;
        }
    }
    __device__ void uuid(EnUuid &acc) {
        EnUuid t{1u, 0ull, 0ull};
        const uint32_t old = r.push_message();
        while (!r.err) {
            const uint32_t tag = r.read_tag();
            if (r.err || tag == 0) break;
            if (tag == 9)
                t.msb = r.fixed64();
            else if (tag == 17)
                t.lsb = r.fixed64();
            else if (!r.unknown(tag))
                break;
        }
        pop(old);
        if (!r.err) en_merge(acc, t);
    }
    __device__ void opt_string(EnStr &acc) {
        EnStr t{1u, 0u, 0u};
        const uint32_t old = r.push_message();
        while (!r.err) {
            const uint32_t tag = r.read_tag();
            if (r.err || tag == 0) break;
            if (tag == 10)
                string(t.off, t.len);
            else if (!r.unknown(tag))
                break;
        }
        pop(old);
        if (!r.err) en_merge(acc, t);
    }
    __device__ void opt_double(EnDbl &acc) {
        EnDbl t{1u, 0ull};
        const uint32_t old = r.push_message();
        while (!r.err) {
            const uint32_t tag = r.read_tag();
            if (r.err || tag == 0) break;
            if (tag == 9)
                t.bits = r.fixed64();
            else if (!r.unknown(tag))
                break;
        }
        pop(old);
        if (!r.err) en_merge(acc, t);
    }
    __device__ void map_entry() {  // map<string,string>: key = 1, value = 2, both validated
        const uint32_t old = r.push_message();
        uint32_t o, l;
        while (!r.err) {
            const uint32_t tag = r.read_tag();
            if (r.err || tag == 0) break;
            if (tag == 10 || tag == 18)
                string(o, l);
            else if (!r.unknown(tag))
                break;
        }
        pop(old);
    }
    __device__ void opaque() {  // an event type off the path: a well-formed message at the wire level
        const uint32_t old = r.push_message();
        while (!r.err) {
            const uint32_t tag = r.read_tag();
            if (r.err || tag == 0) break;
            if (!r.unknown(tag)) break;
        }
        pop(old);
    }
    __device__ void event(EnEvent &acc) {
        EnEvent t;
        memset(&t, 0, sizeof t);
        t.has = 1;
        EnUuid scratch;
        const uint32_t old = r.push_message();
        while (!r.err) {
            const uint32_t tag_pos = r.pos;
            const uint32_t tag = r.read_tag();
            if (r.err || tag == 0) break;
            scratch.has = 0;
            if (tag == 10)
                uuid(t.id);
            else if (tag == 18)
                opt_string(t.alt);
            else if (tag == 24)
                t.etype = r.varint32();
            else if (tag == 34)
                uuid(t.device);
            else if (tag == 42)
                uuid(t.assignment);
            else if (tag == 50 || tag == 58 || tag == 66)
                uuid(scratch);
            else if (tag == 72)
                t.date = r.varint64();
            else if (tag == 80)
                r.varint64();
            else if (tag == 90) {
                if (!t.md_count) t.md_first = tag_pos;
                t.md_count++;
                map_entry();
            } else if (!r.unknown(tag))
                break;
        }
        pop(old);
        if (!r.err) en_merge(acc, t);
    }
    __device__ void member(EnAny &a, uint32_t which) {
        EnAny t;
        memset(&t, 0, sizeof t);
        const uint32_t old = r.push_message();
        while (!r.err) {
            const uint32_t tag = r.read_tag();
            if (r.err || tag == 0) break;
            bool known = true;
            if (which == 1) {
                if (tag == 10)
                    string(t.name_off, t.name_len);
                else if (tag == 17)
                    t.value_bits = r.fixed64();
                else if (tag == 26)
                    event(t.ev);
                else
                    known = false;
            } else if (which == 2) {
                if (tag == 8)
                    t.source = r.varint32();
                else if (tag == 16)
                    t.level = r.varint32();
                else if (tag == 26)
                    string(t.name_off, t.name_len);
                else if (tag == 34)
                    string(t.msg_off, t.msg_len);
                else if (tag == 42)
                    event(t.ev);
                else
                    known = false;
            } else {
                if (tag == 10)
                    opt_double(t.lat);
                else if (tag == 18)
                    opt_double(t.lon);
                else if (tag == 26)
                    opt_double(t.elev);
                else if (tag == 34)
                    event(t.ev);
                else
                    known = false;
            }
            if (!known && !r.unknown(tag)) break;
        }
        pop(old);
        if (r.err) return;
        t.ocase = which;
        en_merge_member(a, t);
    }
    __device__ void any(EnAny &acc) {
        EnAny t;
        memset(&t, 0, sizeof t);
        const uint32_t old = r.push_message();
        while (!r.err) {
            const uint32_t tag = r.read_tag();
            if (r.err || tag == 0) break;
            if (tag == 10)
                member(t, 1);
            else if (tag == 18)
                member(t, 2);
            else if (tag == 26)
                member(t, 3);
            else if (tag == 34 || tag == 42 || tag == 50) {
                opaque();
                memset(&t, 0, sizeof t);
                t.ocase = 4;
            } else if (!r.unknown(tag))
                break;
        }
        pop(old);
        if (r.err || t.ocase == 0) return;
        en_merge_member(acc, t);
    }
    __device__ void context(uint32_t &status_acc) {
        EnUuid u;
        EnStr s;
        uint32_t st = 0;
        const uint32_t old = r.push_message();
        while (!r.err) {
            const uint32_t tag = r.read_tag();
            if (r.err || tag == 0) break;
            u.has = 0;
            s.has = 0;
            if (tag == 10 || tag == 18 || tag == 26)
                uuid(u);
            else if (tag == 34)
                opt_string(s);
            else if (tag == 42 || tag == 58)
                map_entry();
            else if (tag == 48)
                st = r.varint32();
            else if (!r.unknown(tag))
                break;
        }
        pop(old);
        if (!r.err && st) status_acc = st;
    }
};

struct EnRow {
    uint32_t status;
    uint64_t meta, th, nh, nspan, mspan, mdspan, aspan;
    int64_t date;
    uint64_t v0, v1, v2;
    uint64_t event_hi, event_lo, asg_hi, asg_lo, dev_hi, dev_lo;
};

// hash of the 16 key bytes (msb, lsb little-endian) with the library's row hash
struct EnKeyBytes {
    uint64_t msb, lsb;
    __device__ __forceinline__ uint32_t byte(uint32_t i) const { return (uint32_t)((i < 8 ? msb >> (8 * i) : lsb >> (8 * (i - 8))) & 0xff); }
    __device__ __forceinline__ uint64_t le64(uint32_t i) const { return i == 0 ? msb : lsb; }  // (only offsets 0 and 8 are asked for)
};

// what the converter makes of a parsed payload (EventModelConverter.asApiEnrichedEventPayload): enum checks, flags, row
template <typename B>
__device__ __forceinline__ void en_finish(const B &bytes, uint32_t rec, const EnAny &ev, uint32_t asg_status, EnRow &row) {
    if (ev.ocase == 4) {
        row.status = SWGPU_ST_PASSTHROUGH;
        return;
    }
    if (asg_status > 3 || ev.ocase == 0 || ev.ev.etype > 5 || (ev.ocase == 2 && (ev.source > 1 || ev.level > 3))) {
        row.status = SWGPU_ST_DECODE_EXCEPTION;  // SiteWhereException: unknown enum value / no event set
        return;
    }
    row.status = SWGPU_ST_OK;
    const uint32_t kind = ev.ocase == 1 ? SWGPU_KIND_MEASUREMENT : (ev.ocase == 2 ? SWGPU_KIND_ALERT : SWGPU_KIND_LOCATION);
    uint32_t fl = 0;
    if (ev.ev.date) fl |= SWGPU_F_HAS_DATE;
    if (ev.ev.md_count) fl |= SWGPU_F_HAS_METADATA;
    if (ev.ev.alt.has) fl |= SWGPU_F_HAS_ALT_ID;
    if (kind == SWGPU_KIND_MEASUREMENT) {
        fl |= SWGPU_F_HAS_NAME | SWGPU_F_HAS_V0;
        row.v0 = ev.value_bits;
    } else if (kind == SWGPU_KIND_ALERT) {
        fl |= SWGPU_F_HAS_NAME | SWGPU_F_HAS_MESSAGE;
        row.mspan = sw_mk_span(ev.msg_off, ev.msg_len);
    } else {
        if (ev.lat.has) fl |= SWGPU_F_HAS_V0, row.v0 = ev.lat.bits;
        if (ev.lon.has) fl |= SWGPU_F_HAS_V1, row.v1 = ev.lon.bits;
        if (ev.elev.has) fl |= SWGPU_F_HAS_V2, row.v2 = ev.elev.bits;
    }
    if (kind != SWGPU_KIND_LOCATION) {
        row.nspan = sw_mk_span(ev.name_off, ev.name_len);
        row.nh = sw_hash64(bytes, ev.name_off, ev.name_len);
    }
    row.th = sw_hash64(EnKeyBytes{ev.ev.assignment.msb, ev.ev.assignment.lsb}, 0, 16);
    row.meta = sw_mk_meta(kind, fl, kind == SWGPU_KIND_ALERT ? ev.level : 0, kind == SWGPU_KIND_ALERT ? ev.source : 0, 1,
                          ev.ev.etype, rec);
    row.date = (int64_t)ev.ev.date;
    row.mdspan = ev.ev.md_count ? sw_mk_span(ev.ev.md_first, ev.ev.md_count) : 0;
    row.aspan = ev.ev.alt.has ? sw_mk_span(ev.ev.alt.off, ev.ev.alt.len) : 0;
    row.event_hi = ev.ev.id.msb;
    row.event_lo = ev.ev.id.lsb;
    row.asg_hi = ev.ev.assignment.msb;
    row.asg_lo = ev.ev.assignment.lsb;
    row.dev_hi = ev.ev.device.msb;
    row.dev_lo = ev.ev.device.lsb;
}

template <typename B>
__device__ void en_decode_record(B bytes, uint32_t total, uint32_t rec, EnRow &row) {
    memset(&row, 0, sizeof row);
    EnParser<B> p;
    p.r.b = bytes;
    p.r.pos = 0;
    p.r.end = total;
    p.r.limit = total;  // parseFrom(byte[]): the array's length is the outermost limit
    p.r.err = false;
    p.r.last_tag = 0;
    EnAny ev;
    memset(&ev, 0, sizeof ev);
    uint32_t asg_status = 0;
    while (!p.r.err) {
        const uint32_t tag = p.r.read_tag();
        if (p.r.err || tag == 0) break;
        if (tag == 10)
            p.context(asg_status);
        else if (tag == 18)
            p.any(ev);
        else if (!p.r.unknown(tag))
            break;
    }
    if (!p.r.err && p.r.last_tag != 0) p.r.err = true;
    if (p.r.err) {
        row.status = SWGPU_ST_DECODE_EXCEPTION;
        return;
    }
    en_finish(bytes, rec, ev, asg_status, row);
}

// ---- the serialiser's own layout ---------------------------------------------------------------------------------
// protobuf-java writes a message's known fields in field-number order, each once (map entries one after another), and
// proto3 3.3.0 keeps no unknown fields: every payload the reference itself produces can be read in ONE forward pass with
// no merge bookkeeping.  en_fast_parse accepts exactly that — the expected tags in ascending order, lengths of at most
// two bytes, single-byte enum values, strings without a byte >= 0x80 — and fills the same EnAny the general parser
// fills (a field that occurs once is taken as parsed there too); anything else returns false and the general parser
// decodes the record from the start, so this path never decides an error.  Shared-memory records only.
struct EnFast {
    const uint8_t *s;
    uint32_t p, end;
    bool bad;
    __device__ __forceinline__ bool at(uint32_t tag) const { return p < end && s[p] == tag; }
    // the length after a tag at p (one or two bytes); p moves to the body
    __device__ __forceinline__ uint32_t len() {
        uint32_t l = p + 1 < end ? s[p + 1] : 0x80u;
        p += 2;
        if (l & 0x80) {
            const uint32_t h = p < end ? s[p] : 0x80u;
            p++;
            if (h & 0x80) bad = true;
            l = (l & 0x7f) | (h << 7);
        }
        if (p + l > end) bad = true;
        return l;
    }
    __device__ __forceinline__ uint64_t le64(uint32_t o) const { return SmemBytes{s}.le64(o); }
    __device__ __forceinline__ void ascii(uint32_t off, uint32_t l) {
        uint32_t acc = 0;
        for (uint32_t i = 0; i < l; i++) acc |= s[off + i];
        if (acc & 0x80) bad = true;
    }
    // GUUID {fixed64 msb = 1; fixed64 lsb = 2}
    __device__ __forceinline__ void uuid(uint32_t tag, EnUuid &u) {
        if (!at(tag)) return;
        const uint32_t l = len();
        if (bad) return;
        u.has = 1;
        uint32_t q = p;
        if (q < p + l && s[q] == 0x09 && q + 9 <= p + l) {
            u.msb = le64(q + 1);
            q += 9;
        }
        if (q < p + l && s[q] == 0x11 && q + 9 <= p + l) {
            u.lsb = le64(q + 1);
            q += 9;
        }
        if (q != p + l) bad = true;
        p += l;
    }
    __device__ __forceinline__ void str(uint32_t tag, uint32_t &off, uint32_t &l) {
        if (!at(tag)) return;
        l = len();
        if (bad) return;
        off = p;
        ascii(off, l);
        p += l;
    }
    // GOptionalString {string value = 1}
    __device__ __forceinline__ void opt_str(uint32_t tag, EnStr &t) {
        if (!at(tag)) return;
        const uint32_t l = len();
        if (bad) return;
        const uint32_t e = p + l;
        t.has = 1;
        EnFast in{s, p, e, false};
        in.str(0x0A, t.off, t.len);
        if (in.bad || in.p != e) bad = true;
        p = e;
    }
    // GOptionalDouble {double value = 1}
    __device__ __forceinline__ void opt_dbl(uint32_t tag, EnDbl &t) {
        if (!at(tag)) return;
        const uint32_t l = len();
        if (bad) return;
        t.has = 1;
        if (l == 9 && s[p] == 0x09)
            t.bits = le64(p + 1);
        else if (l != 0)
            bad = true;
        p += l;
    }
    // map<string,string> entries of one field, one after another; returns how many
    __device__ __forceinline__ uint32_t map(uint32_t tag, uint32_t &first_pos) {
        uint32_t cnt = 0;
        while (!bad && at(tag)) {
            if (!cnt) first_pos = p;
            cnt++;
            const uint32_t l = len();
            if (bad) break;
            const uint32_t e = p + l;
            EnFast in{s, p, e, false};
            uint32_t o, n;
            in.str(0x0A, o, n);
            in.str(0x12, o, n);
            if (in.bad || in.p != e) bad = true;
            p = e;
        }
        return cnt;
    }
    __device__ __forceinline__ void enum1(uint32_t tag, uint32_t &v) {  // single-byte varint
        if (!at(tag)) return;
        if (p + 1 >= end || (s[p + 1] & 0x80)) {
            bad = true;
            return;
        }
        v = s[p + 1];
        p += 2;
    }
    __device__ __forceinline__ void varint(uint32_t tag, uint64_t &v) {
        if (!at(tag)) return;
        p++;
        uint64_t r = 0;
        for (uint32_t k = 0; k < 10; k++) {
            if (p >= end) break;
            const uint32_t b = s[p++];
            r |= (uint64_t)(b & 0x7f) << (7 * k);
            if (!(b & 0x80)) {
                v = r;
                return;
            }
        }
        bad = true;
    }
    // GDeviceEvent
    __device__ __forceinline__ void event(uint32_t tag, EnEvent &t) {
        if (!at(tag)) return;
        const uint32_t l = len();
        if (bad) return;
        const uint32_t e = p + l;
        EnFast in{s, p, e, false};
        t.has = 1;
        in.uuid(0x0A, t.id);
        in.opt_str(0x12, t.alt);
        in.enum1(0x18, t.etype);
        in.uuid(0x22, t.device);
        in.uuid(0x2A, t.assignment);
        EnUuid drop{0u, 0ull, 0ull};
        in.uuid(0x32, drop);
        in.uuid(0x3A, drop);
        in.uuid(0x42, drop);
        in.varint(0x48, t.date);
        uint64_t received = 0;
        in.varint(0x50, received);
        t.md_count = in.map(0x5A, t.md_first);
        if (in.bad || in.p != e) bad = true;
        p = e;
    }
};
__device__ __forceinline__ bool en_fast_parse(const uint8_t *s, uint32_t total, EnAny &ev, uint32_t &asg_status) {
    EnFast f{s, 0u, total, false};
    if (f.at(0x0A)) {  // GDeviceEventContext
        const uint32_t l = f.len();
        if (f.bad) return false;
        const uint32_t e = f.p + l;
        EnFast c{s, f.p, e, false};
        EnUuid drop{0u, 0ull, 0ull};
        EnStr st{0u, 0u, 0u};
        uint32_t first;
        c.uuid(0x0A, drop);
        c.uuid(0x12, drop);
        c.uuid(0x1A, drop);
        c.opt_str(0x22, st);
        c.map(0x2A, first);
        c.enum1(0x30, asg_status);
        c.map(0x3A, first);
        if (c.bad || c.p != e) return false;
        f.p = e;
    }
    if (f.at(0x12)) {  // GAnyDeviceEvent: one member
        const uint32_t l = f.len();
        if (f.bad || f.p + l != total || l == 0) return false;
        const uint32_t which = s[f.p] == 0x0A ? 1u : s[f.p] == 0x12 ? 2u : s[f.p] == 0x1A ? 3u : 0u;
        if (!which) return false;  // (another event type, an unknown field: the general parser says what it is)
        const uint32_t lm = f.len();
        if (f.bad || f.p + lm != total) return false;
        EnFast m{s, f.p, total, false};
        ev.ocase = which;
        if (which == 1) {
            m.str(0x0A, ev.name_off, ev.name_len);
            if (m.at(0x11) && m.p + 9 <= m.end) {
                ev.value_bits = m.le64(m.p + 1);
                m.p += 9;
            }
            m.event(0x1A, ev.ev);
        } else if (which == 2) {
            m.enum1(0x08, ev.source);
            m.enum1(0x10, ev.level);
            m.str(0x1A, ev.name_off, ev.name_len);
            m.str(0x22, ev.msg_off, ev.msg_len);
            m.event(0x2A, ev.ev);
        } else {
            m.opt_dbl(0x0A, ev.lat);
            m.opt_dbl(0x12, ev.lon);
            m.opt_dbl(0x1A, ev.elev);
            m.event(0x22, ev.ev);
        }
        if (m.bad || m.p != total) return false;
        f.p = total;
    }
    return f.p == total;
}

// key words of an enriched row: the assignment id's two words, then the name from the payload
template <typename B>
struct EnMgKey {
    uint64_t msb, lsb;
    B b;
    uint32_t name_off, nl;
    __device__ __forceinline__ uint64_t word(uint32_t i) const {
        if (i == 1) return msb;
        if (i == 2) return lsb;
        const uint32_t j = i - 3, valid = min(8u, nl - 8 * j);
        if (valid >= 8) return b.le64(name_off + 8 * j);
        uint64_t v = 0;
        for (uint32_t k = 0; k < valid; k++) v |= (uint64_t)b.byte(name_off + 8 * j + k) << (8 * k);
        return v;
    }
};

__device__ __forceinline__ void en_write_row(const EnArgs &a, const uint64_t row, const EnRow &r) {
    a.cols.meta[row] = r.meta;
    a.cols.token_hash[row] = r.th;
    a.cols.name_hash[row] = r.nh;
    a.cols.event_date[row] = r.date;
    a.cols.v0[row] = __longlong_as_double((long long)r.v0);
    a.cols.v1[row] = __longlong_as_double((long long)r.v1);
    a.cols.v2[row] = __longlong_as_double((long long)r.v2);
    a.cols.token_span[row] = 0;
    a.cols.name_span[row] = r.nspan;
    if (a.cols.originator_span) a.cols.originator_span[row] = 0;
    if (a.cols.message_span) a.cols.message_span[row] = r.mspan;
    if (a.cols.metadata_span) a.cols.metadata_span[row] = r.mdspan;
    if (a.cols.alternate_id_span) a.cols.alternate_id_span[row] = r.aspan;
    if (a.ids.event_hi) {
        a.ids.event_hi[row] = r.event_hi;
        a.ids.event_lo[row] = r.event_lo;
        a.ids.assignment_hi[row] = r.asg_hi;
        a.ids.assignment_lo[row] = r.asg_lo;
        a.ids.device_hi[row] = r.dev_hi;
        a.ids.device_lo[row] = r.dev_lo;
    }
}
// one decoded row into the device-state table (fused decode + merge); key = assignment id + name
template <typename B>
__device__ __forceinline__ void en_merge_row(const EnArgs &a, MgLocalCnt *cnt, const B &rec_bytes, const EnRow &r, uint32_t rec) {
    const uint32_t kind = SWGPU_META_KIND(r.meta), fl = SWGPU_META_FLAGS(r.meta);
    const uint32_t nl = kind == SWGPU_KIND_LOCATION ? 0u : SWGPU_SPAN_LEN(r.nspan);
    if (nl > 0xffffu) {
        atomicAdd(&a.mt.ctr->key_too_long, 1ull);
        return;
    }
    const uint32_t kindbyte = mg_kindbyte(kind, fl);
    const uint64_t h = mg_mix(r.th, kind == SWGPU_KIND_LOCATION ? 0ull : r.nh, kindbyte);
    const uint64_t seq1 = a.seq_base + ((uint64_t)(a.rec_base + rec) << SWGPU_SEQ_RECORD_SHIFT) + 1;
    EnMgKey<B> key{r.asg_hi, r.asg_lo, rec_bytes, SWGPU_SPAN_OFF(r.nspan), nl};
    mg_upsert(a.mt, cnt, h, r.th, 16u, nl, kindbyte, key, seq1, r.event_hi, r.event_lo, (uint64_t)r.date, r.v0);
}

// MODE 0: optimistic pass, row index = record index (raises *not_dense when a record yields no row)
// MODE 1: ordered redo (ticketed tiles + look-back), returns at once unless *not_dense
// MODE 2: fused decode + merge (newest records first; no columns)
template <int MODE>
__device__ __forceinline__ void en_decode_tile(const EnArgs &a, const uint32_t tile, uint8_t *smem, uint64_t *s_warp,
                                               MgLocalCnt *s_cnt) {
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t rec0_64 = (uint64_t)tile * EN_THREADS + (MODE == 2 ? EN_WARPS - 1 - warp : warp) * 32;
    const uint32_t rec0 = rec0_64 < a.n ? (uint32_t)rec0_64 : a.n;
    const uint32_t nrec = min(32u, a.n - rec0);
    uint8_t *ws = smem + warp * (EN_WSTAGE + 32);
    uint64_t r0 = 0, r1 = 0;
    if (lane < nrec) {
        r0 = a.rec_off[rec0 + lane];
        r1 = a.rec_off[rec0 + lane + 1];
    }
    // after the forward pass (enriched_fast_kernel) modes 0 and 2 decode only the records that one left
    bool mine = lane < nrec;
    if (MODE != 1 && a.fast) {
        mine = mine && a.status[rec0 + lane] == EN_ST_PENDING;
        if (!__any_sync(SW_FULL, mine)) return;
    }
    __syncwarp();  // (modes 0 and 2 walk tiles in a loop: every lane is done with the previous tile's bytes)
    const uint64_t b0 = __shfl_sync(SW_FULL, r0, 0);
    const uint64_t b1 = __shfl_sync(SW_FULL, r1, nrec ? nrec - 1 : 0);
    const uintptr_t addr = (uintptr_t)(a.bytes + b0);
    const uint32_t lead = (uint32_t)(addr & 15);
    const uint64_t span = (uint64_t)lead + (b1 - b0);
    const bool staged = span <= EN_WSTAGE;
    if (staged && nrec) {
        const uint4 *src = (const uint4 *)(addr - lead);
        uint4 *dst = (uint4 *)ws;
        const uint32_t nvec = (uint32_t)((span + 15) >> 4);
        for (uint32_t i = lane; i < nvec; i += 32) dst[i] = sw_ld_stream_u4(src + i);
    }
    __syncwarp();
    EnRow r;
    r.status = 0;
    uint32_t count = 0;
    if (mine) {
        const uint64_t l64 = r1 - r0;
        const uint32_t len = l64 > 0x7fffffffull ? 0x7fffffffu : (uint32_t)l64;
        if (staged)
            en_decode_record(SmemBytes{ws + lead + (uint32_t)(r0 - b0)}, len, a.rec_base + rec0 + lane, r);
        else
            en_decode_record(GmemBytes{a.bytes + r0}, len, a.rec_base + rec0 + lane, r);
        count = r.status == SWGPU_ST_OK ? 1u : 0u;
        a.status[rec0 + lane] = (uint8_t)r.status;
    }
    if (MODE == 2) {
        if (count) {
            if (staged)
                en_merge_row(a, &s_cnt[warp], SmemBytes{ws + lead + (uint32_t)(r0 - b0)}, r, rec0 + lane);
            else
                en_merge_row(a, &s_cnt[warp], GmemBytes{a.bytes + r0}, r, rec0 + lane);
        }
        __syncwarp();
        count = __reduce_add_sync(SW_FULL, count);
        if (lane == 0) {
            if (count) atomicAdd(a.out_rows, count);
            mg_flush_counts(a.mt, &s_cnt[warp]);
        }
        return;
    }
    uint64_t tile_total = 0, row;
    if (MODE == 0) {
        row = (uint64_t)rec0 + lane;
        if (mine && count != 1) atomicOr(a.not_dense, 1u);
    } else {
        row = sw_block_ordered_offset<EN_THREADS>(count, a.tile_state, tile, s_warp, &tile_total);
    }
    const uint64_t row_base = a.row_base ? (uint64_t)__ldg(a.row_base) : 0ull;
    if (mine) {
        a.first_row[rec0 + lane] = (uint32_t)(row_base + row);
        if (count && row < a.max_rows) en_write_row(a, row, r);
    }
    if ((uint64_t)tile * EN_THREADS + EN_THREADS >= a.n && tid == 0) {  // the tile holding the last record
        const uint64_t total = MODE == 0 ? (uint64_t)a.n : row + tile_total;
        a.first_row[a.n] = (uint32_t)(row_base + total);
        *a.out_rows = (uint32_t)total;
        if (a.row_next) *a.row_next = (uint32_t)(row_base + total);
    }
}

// Forward pass (option "enriched.fast_path"): one warp per CTA (its staging goes back the moment its 32 records are
// done), en_fast_parse on each record; MERGE = false: row index = record index as in the dense pass; MERGE = true: the
// row goes straight into the device-state table, newest records first.  A record that is not in the serialiser's
// layout is marked EN_ST_PENDING for the general pass that follows.  None of the general parser is in this kernel.
template <bool MERGE>
__global__ void __launch_bounds__(32) enriched_fast_kernel(EnArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];  // EN_WSTAGE + 32
    __shared__ MgLocalCnt s_cnt[1];
    const uint32_t lane = threadIdx.x;
    const uint32_t tile = MERGE ? gridDim.x - 1 - blockIdx.x : blockIdx.x;
    const uint64_t rec0_64 = (uint64_t)tile * 32;
    const uint32_t rec0 = rec0_64 < a.n ? (uint32_t)rec0_64 : a.n;
    const uint32_t nrec = min(32u, a.n - rec0);
    if (MERGE && lane == 0) s_cnt[0].keys = s_cnt[0].bytes = 0;
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
    const bool staged = span <= EN_WSTAGE;
    if (staged && nrec) {
        const uint4 *src = (const uint4 *)(addr - lead);
        uint4 *dst = (uint4 *)smem;
        const uint32_t nvec = (uint32_t)((span + 15) >> 4);
        for (uint32_t i = lane; i < nvec; i += 32) dst[i] = sw_ld_stream_u4(src + i);
    }
    __syncwarp();
    const uint64_t row_base = (!MERGE && a.row_base) ? (uint64_t)__ldg(a.row_base) : 0ull;
    uint32_t ok = 0, left = 0;
    if (lane < nrec) {
        EnRow r;
        memset(&r, 0, sizeof r);
        bool hit = false;
        const SmemBytes rb{smem + lead + (uint32_t)(r0 - b0)};
        if (staged) {
            EnAny ev;
            memset(&ev, 0, sizeof ev);
            uint32_t asg_status = 0;
            hit = en_fast_parse(rb.s, (uint32_t)(r1 - r0), ev, asg_status);
            if (hit) en_finish(rb, a.rec_base + rec0 + lane, ev, asg_status, r);
        }
        a.status[rec0 + lane] = hit ? (uint8_t)r.status : (uint8_t)EN_ST_PENDING;
        left = hit ? 0u : 1u;
        ok = hit && r.status == SWGPU_ST_OK ? 1u : 0u;
        if (MERGE) {
            if (ok) en_merge_row(a, &s_cnt[0], rb, r, rec0 + lane);
        } else {
            const uint64_t row = (uint64_t)rec0 + lane;
            a.first_row[rec0 + lane] = (uint32_t)(row_base + row);
            if (hit && !ok) atomicOr(a.not_dense, 1u);  // decoded here, to no row: the ordered pass will place the rows
            if (ok && row < a.max_rows) en_write_row(a, row, r);
        }
    }
    __syncwarp();
    left = __reduce_add_sync(SW_FULL, left);
    if (MERGE) {
        ok = __reduce_add_sync(SW_FULL, ok);
        if (lane == 0) {
            if (ok) atomicAdd(a.out_rows, ok);
            mg_flush_counts(a.mt, &s_cnt[0]);
        }
    } else if ((uint64_t)tile * 32 + 32 >= a.n && lane == 0) {  // the tile holding the last record
        a.first_row[a.n] = (uint32_t)(row_base + a.n);
        *a.out_rows = a.n;
        if (a.row_next) *a.row_next = (uint32_t)(row_base + a.n);
    }
    if (lane == 0 && left) atomicOr(a.pending, 1u);
}

template <int MODE>
__global__ void __launch_bounds__(EN_THREADS) enriched_decode_kernel(EnArgs a) {
    extern __shared__ __align__(16) uint8_t smem[];  // EN_WARPS * (EN_WSTAGE + 32)
    __shared__ uint64_t s_warp[EN_THREADS / 32 + 2];
    __shared__ MgLocalCnt s_cnt[EN_WARPS];
    __shared__ uint32_t s_tile;
    const uint32_t ntiles = (uint32_t)(((uint64_t)a.n + EN_THREADS - 1) / EN_THREADS);
    if (MODE == 0) {
        if (a.fast && __ldg(a.pending) == 0) return;  // the forward pass took every record
        for (uint32_t t = blockIdx.x; t < ntiles; t += gridDim.x) en_decode_tile<0>(a, t, smem, s_warp, s_cnt);
        return;
    }
    if (MODE == 2) {
        if (a.fast && __ldg(a.pending) == 0) return;
        if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5].keys = s_cnt[threadIdx.x >> 5].bytes = 0;
        __syncwarp();
        for (uint32_t t = blockIdx.x; t < ntiles; t += gridDim.x)
            en_decode_tile<2>(a, ntiles - 1 - t, smem, s_warp, s_cnt);  // newest records first
        return;
    }
    if (!a.force_ordered && __ldg(a.not_dense) == 0) return;
    for (;;) {
        if (threadIdx.x == 0) s_tile = atomicAdd(a.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= ntiles) return;
        en_decode_tile<1>(a, tile, smem, s_warp, s_cnt);
        __syncthreads();
    }
}

static int en_attr(swgpu_ctx *ctx) {
    static std::mutex mu;
    static bool done[64] = {false};
    std::lock_guard<std::mutex> lk(mu);
    const int d = ctx->device & 63;
    if (!done[d]) {
        const int smem = EN_WARPS * (EN_WSTAGE + 32);
        SW_CUDA(ctx, cudaFuncSetAttribute(enriched_decode_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        SW_CUDA(ctx, cudaFuncSetAttribute(enriched_decode_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        SW_CUDA(ctx, cudaFuncSetAttribute(enriched_decode_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        SW_CUDA(ctx, cudaFuncSetAttribute(enriched_fast_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, EN_WSTAGE + 32));
        SW_CUDA(ctx, cudaFuncSetAttribute(enriched_fast_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, EN_WSTAGE + 32));
        done[d] = true;
    }
    return SWGPU_OK;
}

int sw_decode_enriched_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint32_t max_rows,
                              const swgpu_decoded_cols *out, const swgpu_event_ids *ids, uint8_t *status, uint32_t *first_row,
                              uint32_t *out_rows_dev, uint32_t rec_base, const uint32_t *row_base, uint32_t *row_next, bool small) {
    if (!ctx) return SWGPU_E_INVALID;
    if (!out || !out->meta || !out->token_hash || !out->name_hash || !out->event_date || !out->v0 || !out->v1 ||
        !out->v2 || !out->token_span || !out->name_span) {
        SW_SET_ERR(ctx, "decoded columns: the nine mandatory column pointers must be non-NULL");
        return SWGPU_E_INVALID;
    }
    if (ids && ids->event_hi &&
        (!ids->event_lo || !ids->assignment_hi || !ids->assignment_lo || !ids->device_hi || !ids->device_lo)) {
        SW_SET_ERR(ctx, "swgpu_event_ids: give all six columns or none");
        return SWGPU_E_INVALID;
    }
    if (!rec_off || !first_row || !out_rows_dev || (n && !status)) {
        SW_SET_ERR(ctx, "swgpu_decode_enriched_dev: NULL argument");
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    if (n == 0) {
        SW_CUDA(ctx, cudaMemsetAsync(first_row, 0, sizeof(uint32_t), ctx->stream));
        SW_CUDA(ctx, cudaMemsetAsync(out_rows_dev, 0, sizeof(uint32_t), ctx->stream));
        return SWGPU_OK;
    }
    const uint32_t ntiles = (n + EN_THREADS - 1) / EN_THREADS;
    const size_t sbytes = SW_SCRATCH_HDR + (size_t)ntiles * 8;
    if (!small) {  // (small: the caller has sized and cleared the scratch; see sw_decode_pb_launch)
        if (int rc = sw_ensure_scratch(ctx, sbytes)) return rc;
        SW_CUDA(ctx, cudaMemsetAsync(ctx->scratch, 0, sbytes, ctx->stream));
    }
    EnArgs a;
    memset(&a, 0, sizeof a);
    a.force_ordered = small ? 1u : 0u;
    a.bytes = bytes;
    a.rec_off = rec_off;
    a.n = n;
    a.max_rows = max_rows;
    a.cols = *out;
    if (ids && ids->event_hi) a.ids = *ids;
    a.status = status;
    a.first_row = first_row;
    a.out_rows = out_rows_dev;
    a.ticket = (uint32_t *)ctx->scratch;
    a.tile_state = (uint64_t *)((uint8_t *)ctx->scratch + SW_SCRATCH_HDR);
    a.rec_base = rec_base;
    a.row_base = row_base;
    a.row_next = row_next;
    a.not_dense = (uint32_t *)ctx->scratch + 1;
    if (int rc = en_attr(ctx)) return rc;
    const int smem = EN_WARPS * (EN_WSTAGE + 32);
    const uint32_t nredo = ntiles < (uint32_t)ctx->sm_count * 4u ? ntiles : (uint32_t)ctx->sm_count * 4u;
    if (small) {
        enriched_decode_kernel<1><<<nredo, EN_THREADS, smem, ctx->stream>>>(a);  // the ordered pass decodes the batch
        SW_CUDA(ctx, cudaGetLastError());
        return SWGPU_OK;
    }
    a.fast = ctx->opt_enriched_fast;
    a.pending = (uint32_t *)ctx->scratch + 3;
    if (a.fast) {
        enriched_fast_kernel<false><<<(n + 31) / 32, 32, EN_WSTAGE + 32, ctx->stream>>>(a);
        SW_CUDA(ctx, cudaGetLastError());
    }
    enriched_decode_kernel<0><<<a.fast ? nredo : ntiles, EN_THREADS, smem, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    enriched_decode_kernel<1><<<nredo, EN_THREADS, smem, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}

int sw_decode_merge_enriched_launch(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, const MgTable &t,
                                    uint64_t seq_base, uint8_t *status, uint32_t *out_rows_dev) {
    EnArgs a;
    memset(&a, 0, sizeof a);
    a.bytes = bytes;
    a.rec_off = rec_off;
    a.n = n;
    a.status = status;
    a.out_rows = out_rows_dev;
    a.mt = t;
    a.seq_base = seq_base;
    if (int rc = en_attr(ctx)) return rc;
    const int smem = EN_WARPS * (EN_WSTAGE + 32);
    const uint32_t ntiles = (n + EN_THREADS - 1) / EN_THREADS;
    a.fast = ctx->opt_enriched_fast;
    if (a.fast) {
        if (int rc = sw_ensure_scratch(ctx, SW_SCRATCH_HDR)) return rc;
        a.pending = (uint32_t *)ctx->scratch + 3;
        SW_CUDA(ctx, cudaMemsetAsync(a.pending, 0, sizeof(uint32_t), ctx->stream));
        enriched_fast_kernel<true><<<(n + 31) / 32, 32, EN_WSTAGE + 32, ctx->stream>>>(a);
        SW_CUDA(ctx, cudaGetLastError());
    }
    const uint32_t ngen = a.fast ? std::min<uint32_t>(ntiles, (uint32_t)ctx->sm_count * 4u) : ntiles;
    enriched_decode_kernel<2><<<ngen, EN_THREADS, smem, ctx->stream>>>(a);
    SW_CUDA(ctx, cudaGetLastError());
    return SWGPU_OK;
}

extern "C" int32_t swgpu_decode_enriched_dev(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                                             uint32_t max_rows, const swgpu_decoded_cols *out, const swgpu_event_ids *ids,
                                             uint8_t *status, uint32_t *first_row, uint32_t *out_rows_dev) {
    return sw_decode_enriched_launch(ctx, bytes, rec_off, n, max_rows, out, ids, status, first_row, out_rows_dev, 0, nullptr,
                                     nullptr, false);
}

// host-pointer entry point: stage in, decode, stage out
extern "C" int32_t swgpu_decode_enriched(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                                         uint32_t max_rows, const swgpu_decoded_cols *out, const swgpu_event_ids *ids,
                                         uint8_t *status, uint32_t *first_row, uint32_t *out_rows) {
    if (!ctx) return SWGPU_E_INVALID;
    if (!out || !rec_off || !first_row || !out_rows || (n && !status) || (n && rec_off[n] && !bytes)) {
        SW_SET_ERR(ctx, "swgpu_decode_enriched: NULL argument");
        return SWGPU_E_INVALID;
    }
    SW_CUDA(ctx, cudaSetDevice(ctx->device));
    const uint64_t total = rec_off[n];
    const int NCOL = 13, NID = 6;
    const bool want_ids = ids && ids->event_hi;
    if (int rc = sw_ensure(ctx, ctx->d_bytes, total + 64)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_off, (size_t)(n + 1) * 8)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_status, (size_t)n + 16)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_first, (size_t)(n + 2) * 4 + 16)) return rc;
    if (int rc = sw_ensure(ctx, ctx->d_cols, (size_t)max_rows * 8 * (NCOL + NID) + 64)) return rc;
    cudaStream_t st = ctx->stream;
    if (total) SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_bytes.p, bytes, total, cudaMemcpyHostToDevice, st));
    SW_CUDA(ctx, cudaMemcpyAsync(ctx->d_off.p, rec_off, (size_t)(n + 1) * 8, cudaMemcpyHostToDevice, st));
    swgpu_decoded_cols d;
    swgpu_event_ids di;
    uint64_t *base = (uint64_t *)ctx->d_cols.p;
    uint64_t **dp = (uint64_t **)&d, **ip = (uint64_t **)&di;
    uint64_t *const *hp = (uint64_t *const *)out;
    for (int c = 0; c < NCOL; c++) dp[c] = (c < 9 || hp[c]) ? base + (size_t)c * max_rows : nullptr;
    for (int c = 0; c < NID; c++) ip[c] = want_ids ? base + (size_t)(NCOL + c) * max_rows : nullptr;
    uint32_t *d_first = (uint32_t *)ctx->d_first.p;
    uint32_t *d_rows = d_first + (n + 1);
    if (int rc = swgpu_decode_enriched_dev(ctx, (const uint8_t *)ctx->d_bytes.p, (const uint64_t *)ctx->d_off.p, n, max_rows, &d,
                                           want_ids ? &di : nullptr, (uint8_t *)ctx->d_status.p, d_first, d_rows))
        return rc;
    SW_CUDA(ctx, cudaMemcpyAsync(first_row, d_first, (size_t)(n + 1) * 4, cudaMemcpyDeviceToHost, st));
    if (n) SW_CUDA(ctx, cudaMemcpyAsync(status, ctx->d_status.p, n, cudaMemcpyDeviceToHost, st));
    SW_CUDA(ctx, cudaStreamSynchronize(st));
    const uint32_t rows = first_row[n];
    *out_rows = rows;
    const uint32_t live = rows < max_rows ? rows : max_rows;
    if (live) {
        for (int c = 0; c < NCOL; c++)
            if (dp[c] && hp[c]) SW_CUDA(ctx, cudaMemcpyAsync(hp[c], dp[c], (size_t)live * 8, cudaMemcpyDeviceToHost, st));
        if (want_ids) {
            uint64_t *const *hi = (uint64_t *const *)ids;
            for (int c = 0; c < NID; c++) SW_CUDA(ctx, cudaMemcpyAsync(hi[c], ip[c], (size_t)live * 8, cudaMemcpyDeviceToHost, st));
        }
        SW_CUDA(ctx, cudaStreamSynchronize(st));
    }
    if (rows > max_rows) {
        SW_SET_ERR(ctx, "decode produced %u rows but max_rows is %u", rows, max_rows);
        return SWGPU_E_CAPACITY;
    }
    return SWGPU_OK;
}
