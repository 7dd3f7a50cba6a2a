/*
 * sw_oracle_enriched.c — CPU oracle, part 3: the enriched-event payload the rule-processing and device-state
 * consumers read off Kafka.  TEST INFRASTRUCTURE ONLY (see sw_oracle.h).  PARITY UNPINNED BY THE REFERENCE.
 *
 * Reference (paths relative to the SiteWhere tree; GEV = sitewhere-grpc-event-management/src/main):
 *   - parse: EventModelMarshaler.parseEnrichedEventPayloadMessage = GEnrichedEventPayload.parseFrom(byte[])
 *     (GEV/java/com/sitewhere/grpc/client/event/EventModelMarshaler.java:204-206); called from
 *     DeviceStateProcessingLogic.processRecord (service-device-state/.../processing/DeviceStateProcessingLogic.java:93-108)
 *     and KafkaRuleProcessorHost (service-rule-processing/.../kafka/KafkaRuleProcessorHost.java:170-197)
 *   - schema: GEV/proto/device-event-model.proto:139-215 (GDeviceEvent, GDeviceMeasurement, GDeviceAlert,
 *     GDeviceLocation, GAnyDeviceEvent), :288-296 (GDeviceEventContext), :321-324 (GEnrichedEventPayload);
 *     sitewhere-grpc-model/src/main/proto/sitewhere-common.proto:12-33 (GOptionalDouble/Boolean/String, GUUID),
 *     :146-151 (GDeviceAssignmentStatus)
 *   - conversion: EventModelConverter.asApiEnrichedEventPayload :1926-1931 -> asApiDeviceEventContext :1753-1764,
 *     asApiGenericDeviceEvent :1664-1690 -> asApiDeviceMeasurement :463-469 / asApiDeviceAlert :689-697 /
 *     asApiDeviceLocation :841-848, copyApiDeviceEvent :329-343, enum conversions :277-295, :525-535, :561-575;
 *     CommonModelConverter.asApiUuid :345-350, asApiDate :57-62 (0 <=> null), asApiDeviceAssignmentStatus :406-421
 *
 * protobuf-java 3.3.0 (full runtime, proto3; third party, recalled) semantics restated here:
 *   - readTag: 0 at end of input; field number 0 is invalidTag.  The generated loop switches on the whole tag;
 *     any other tag goes to parseUnknownFieldProto3 = CodedInputStream.skipField (3.3.0 discards proto3 unknown
 *     fields): varint / fixed / length-delimited skipped, groups skipped recursively with the end tag checked,
 *     END_GROUP stops the message, wire types 6 and 7 are invalidWireType.  After the loop the caller checks
 *     checkLastTagWas(0).
 *   - readMessage: length varint (negative -> negativeSize), pushLimit (beyond the enclosing limit ->
 *     truncatedMessage; the outermost limit is the length of the byte array), parse, checkLastTagWas(0).
 *   - a message-typed field that occurs again MERGES into the value held so far (toBuilder().mergeFrom()):
 *     proto3 scalars of the later occurrence replace only when they are not the default (0, 0.0, ""), sub-messages
 *     merge recursively, maps take the union.  For the oneof in GAnyDeviceEvent a different member replaces, the
 *     same member merges.  Scalars repeated inside ONE occurrence: the last one wins.
 *   - proto3 strings (and map<string,string> entries) are read with readStringRequireUtf8: invalid UTF-8 is an
 *     InvalidProtocolBufferException.
 *   - enum fields hold the raw number; a number outside the schema converts to UNRECOGNIZED, for which the
 *     SiteWhere converters throw.
 * Any exception makes the consumer log and skip the record: status 1 (decode exception).  An event of another
 * type (command invocation / response, state change) is status 3 (pass-through: not on the measured path).
 */
#include <stdlib.h>
#include <string.h>

#include "sw_oracle.h"

#define KIND_MEASUREMENT 0u
#define KIND_LOCATION 1u
#define KIND_ALERT 2u
#define F_HAS_DATE (1u << 1)
#define F_HAS_METADATA (1u << 3)
#define F_HAS_ALT_ID (1u << 4)
#define F_HAS_V0 (1u << 6)
#define F_HAS_V1 (1u << 7)
#define F_HAS_V2 (1u << 8)
#define F_HAS_NAME (1u << 10)
#define F_HAS_MESSAGE (1u << 11)
#define ST_OK 0
#define ST_DECODE_EXCEPTION 1
#define ST_PASSTHROUGH 3
#define EN_MAX_GROUP_DEPTH 64 /* the JVM recurses until its stack ends; documented bound */

typedef struct {
    const uint8_t *p;
    uint64_t pos, end, limit;
    int err;
    uint32_t last_tag;
} enr;

static int en_at_end(enr *r) { return r->pos >= r->end || r->pos >= r->limit; }
static int en_byte(enr *r, uint8_t *b) {
    if (en_at_end(r)) {
        r->err = 1;
        return 0;
    }
    *b = r->p[r->pos++];
    return 1;
}
static uint64_t en_varint64(enr *r) {
    uint64_t v = 0;
    for (int shift = 0; shift < 64; shift += 7) {
        uint8_t b;
        if (!en_byte(r, &b)) return 0;
        v |= (uint64_t)(b & 0x7f) << shift;
        if (!(b & 0x80)) return v;
    }
    r->err = 1;
    return 0;
}
static uint32_t en_varint32(enr *r) { return (uint32_t)en_varint64(r); }
static uint64_t en_fixed64(enr *r) {
    uint64_t v = 0;
    for (int i = 0; i < 8; i++) {
        uint8_t b;
        if (!en_byte(r, &b)) return 0;
        v |= (uint64_t)b << (8 * i);
    }
    return v;
}
static void en_skip(enr *r, uint64_t n) {
    uint64_t lim = r->end < r->limit ? r->end : r->limit;
    if (n > lim - r->pos) {
        r->pos = lim;
        r->err = 1;
        return;
    }
    r->pos += n;
}
static uint32_t en_read_tag(enr *r) {
    if (en_at_end(r)) {
        r->last_tag = 0;
        return 0;
    }
    uint32_t t = en_varint32(r);
    if (r->err) return 0;
    r->last_tag = t;
    if ((t >> 3) == 0) {
        r->err = 1;
        return 0;
    }
    return t;
}
/* CodedInputStream.skipField; 0 = END_GROUP (the message stops) */
static int en_skip_field(enr *r, uint32_t tag, int depth) {
    switch (tag & 7) {
    case 0: en_varint64(r); return 1;
    case 1: en_skip(r, 8); return 1;
    case 5: en_skip(r, 4); return 1;
    case 2: {
        int32_t size = (int32_t)en_varint32(r);
        if (r->err) return 1;
        if (size < 0) {
            r->err = 1;
            return 1;
        }
        en_skip(r, (uint64_t)size);
        return 1;
    }
    case 3:
        if (depth >= EN_MAX_GROUP_DEPTH) {
            r->err = 1;
            return 1;
        }
        for (;;) {
            uint32_t t = en_read_tag(r);
            if (r->err) return 1;
            if (t == 0) break;
            if (!en_skip_field(r, t, depth + 1)) break;
            if (r->err) return 1;
        }
        if (r->last_tag != (((tag >> 3) << 3) | 4)) r->err = 1;
        return 1;
    case 4: return 0;
    default: r->err = 1; return 1;
    }
}
/* readMessage prologue / epilogue */
static uint64_t en_push(enr *r) {
    int32_t length = (int32_t)en_varint32(r);
    uint64_t old = r->limit;
    if (r->err) return old;
    if (length < 0) {
        r->err = 1;
        return old;
    }
    uint64_t nl = r->pos + (uint64_t)length;
    if (nl > old) {
        r->err = 1;
        return old;
    }
    r->limit = nl;
    return old;
}
static void en_pop(enr *r, uint64_t old) {
    if (!r->err && r->last_tag != 0) r->err = 1; /* checkLastTagWas(0) */
    r->limit = old;
}
/* com.google.protobuf.Utf8.isValidUtf8 */
static int utf8_valid(const uint8_t *s, uint64_t n) {
    uint64_t i = 0;
    while (i < n) {
        uint32_t b1 = s[i++];
        if (b1 < 0x80) continue;
        if (b1 < 0xC2) return 0;
        if (b1 < 0xE0) {
            if (i >= n || (s[i] & 0xC0) != 0x80) return 0;
            i += 1;
        } else if (b1 < 0xF0) {
            if (i + 1 >= n) return 0;
            uint32_t b2 = s[i], b3 = s[i + 1];
            if ((b2 & 0xC0) != 0x80 || (b3 & 0xC0) != 0x80) return 0;
            if (b1 == 0xE0 && b2 < 0xA0) return 0;
            if (b1 == 0xED && b2 >= 0xA0) return 0;
            i += 2;
        } else {
            if (b1 > 0xF4 || i + 2 >= n) return 0;
            uint32_t b2 = s[i], b3 = s[i + 1], b4 = s[i + 2];
            if ((b2 & 0xC0) != 0x80 || (b3 & 0xC0) != 0x80 || (b4 & 0xC0) != 0x80) return 0;
            if (b1 == 0xF0 && b2 < 0x90) return 0;
            if (b1 == 0xF4 && b2 > 0x8F) return 0;
            i += 3;
        }
    }
    return 1;
}
/* readStringRequireUtf8: span of the string */
static void en_string(enr *r, uint64_t *off, uint64_t *len) {
    int32_t size = (int32_t)en_varint32(r);
    *off = *len = 0;
    if (r->err) return;
    if (size < 0) {
        r->err = 1;
        return;
    }
    uint64_t start = r->pos;
    en_skip(r, (uint64_t)size);
    if (r->err) return;
    if (!utf8_valid(r->p + start, (uint64_t)size)) {
        r->err = 1;
        return;
    }
    *off = start;
    *len = (uint64_t)size;
}

/* ---- message values (defaults = all zero) and their proto3 merge ---- */
typedef struct { int has; uint64_t msb, lsb; } g_uuid;
typedef struct { int has; uint64_t off, len; } g_ostr;
typedef struct { int has; uint64_t bits; } g_odbl;
typedef struct {
    int has;
    g_uuid id, device, assignment, customer, area, asset;
    g_ostr alt;
    uint32_t etype;
    uint64_t date, received;
    uint64_t md_first, md_count;
} g_event;
typedef struct {
    int ocase; /* 0 none, 1 measurement, 2 alert, 3 location, 4..6 other */
    /* measurement / alert */
    uint64_t name_off, name_len, msg_off, msg_len;
    uint64_t value_bits;
    uint32_t source, level;
    g_odbl lat, lon, elev;
    g_event ev;
} g_any;

static int dbl_is_zero(uint64_t bits) { return (bits << 1) == 0; /* +0.0 and -0.0 compare equal to 0D */ }

/* a message-typed field: the first occurrence is taken as parsed, a later one merges (generated parsing code:
 * subBuilder = existing.toBuilder(); value = readMessage(); subBuilder.mergeFrom(value)) */
static void merge_uuid(g_uuid *a, const g_uuid *o) {
    if (!o->has) return;
    if (!a->has) {
        *a = *o;
        return;
    }
    if (o->msb) a->msb = o->msb;
    if (o->lsb) a->lsb = o->lsb;
}
static void merge_ostr(g_ostr *a, const g_ostr *o) {
    if (!o->has) return;
    if (!a->has) {
        *a = *o;
        return;
    }
    if (o->len) {
        a->off = o->off;
        a->len = o->len;
    }
}
static void merge_odbl(g_odbl *a, const g_odbl *o) {
    if (!o->has) return;
    if (!a->has) {
        *a = *o;
        return;
    }
    if (!dbl_is_zero(o->bits)) a->bits = o->bits;
}
static void merge_event(g_event *a, const g_event *o) {
    if (!o->has) return;
    if (!a->has) {
        *a = *o;
        return;
    }
    merge_uuid(&a->id, &o->id);
    merge_ostr(&a->alt, &o->alt);
    if (o->etype) a->etype = o->etype;
    merge_uuid(&a->device, &o->device);
    merge_uuid(&a->assignment, &o->assignment);
    merge_uuid(&a->customer, &o->customer);
    merge_uuid(&a->area, &o->area);
    merge_uuid(&a->asset, &o->asset);
    if (o->date) a->date = o->date;
    if (o->received) a->received = o->received;
    if (o->md_count) {
        if (!a->md_count) a->md_first = o->md_first;
        a->md_count += o->md_count;
    }
}

static void parse_uuid(enr *r, g_uuid *acc) {
    g_uuid t = {1, 0, 0};
    uint64_t old = en_push(r);
    while (!r->err) {
        uint32_t tag = en_read_tag(r);
        if (r->err || tag == 0) break;
        if (tag == 9) t.msb = en_fixed64(r);
        else if (tag == 17) t.lsb = en_fixed64(r);
        else if (!en_skip_field(r, tag, 0)) break;
    }
    en_pop(r, old);
    if (!r->err) merge_uuid(acc, &t);
}
static void parse_ostr(enr *r, g_ostr *acc) {
    g_ostr t = {1, 0, 0};
    uint64_t old = en_push(r);
    while (!r->err) {
        uint32_t tag = en_read_tag(r);
        if (r->err || tag == 0) break;
        if (tag == 10) en_string(r, &t.off, &t.len);
        else if (!en_skip_field(r, tag, 0)) break;
    }
    en_pop(r, old);
    if (!r->err) merge_ostr(acc, &t);
}
static void parse_odbl(enr *r, g_odbl *acc) {
    g_odbl t = {1, 0};
    uint64_t old = en_push(r);
    while (!r->err) {
        uint32_t tag = en_read_tag(r);
        if (r->err || tag == 0) break;
        if (tag == 9) t.bits = en_fixed64(r);
        else if (!en_skip_field(r, tag, 0)) break;
    }
    en_pop(r, old);
    if (!r->err) merge_odbl(acc, &t);
}
/* map<string,string> entry: key = 1, value = 2, both validated */
static void parse_map_entry(enr *r) {
    uint64_t old = en_push(r), o, l;
    while (!r->err) {
        uint32_t tag = en_read_tag(r);
        if (r->err || tag == 0) break;
        if (tag == 10 || tag == 18) en_string(r, &o, &l);
        else if (!en_skip_field(r, tag, 0)) break;
    }
    en_pop(r, old);
}
static void parse_event(enr *r, g_event *acc) {
    g_event t;
    memset(&t, 0, sizeof t);
    t.has = 1;
    uint64_t old = en_push(r);
    while (!r->err) {
        uint64_t tag_pos = r->pos;
        uint32_t tag = en_read_tag(r);
        if (r->err || tag == 0) break;
        switch (tag) {
        case 10: parse_uuid(r, &t.id); break;
        case 18: parse_ostr(r, &t.alt); break;
        case 24: t.etype = en_varint32(r); break;
        case 34: parse_uuid(r, &t.device); break;
        case 42: parse_uuid(r, &t.assignment); break;
        case 50: parse_uuid(r, &t.customer); break;
        case 58: parse_uuid(r, &t.area); break;
        case 66: parse_uuid(r, &t.asset); break;
        case 72: t.date = en_varint64(r); break;
        case 80: t.received = en_varint64(r); break;
        case 90:
            if (!t.md_count) t.md_first = tag_pos;
            t.md_count++;
            parse_map_entry(r);
            break;
        default:
            if (!en_skip_field(r, tag, 0)) goto done;
        }
    }
done:
    en_pop(r, old);
    if (!r->err) merge_event(acc, &t);
}
/* the oneof member held in `a` meets a new value `t`: another member replaces, the same member merges */
static void merge_member(g_any *a, const g_any *t) {
    if (a->ocase != t->ocase) {
        *a = *t;
        return;
    }
    if (t->name_len) {
        a->name_off = t->name_off;
        a->name_len = t->name_len;
    }
    if (t->msg_len) {
        a->msg_off = t->msg_off;
        a->msg_len = t->msg_len;
    }
    if (!dbl_is_zero(t->value_bits)) a->value_bits = t->value_bits;
    if (t->source) a->source = t->source;
    if (t->level) a->level = t->level;
    merge_odbl(&a->lat, &t->lat);
    merge_odbl(&a->lon, &t->lon);
    merge_odbl(&a->elev, &t->elev);
    merge_event(&a->ev, &t->ev);
}
/* one occurrence of a oneof member (1 measurement, 2 alert, 3 location) into `a` */
static void parse_member(enr *r, g_any *a, int which) {
    g_any t;
    memset(&t, 0, sizeof t);
    uint64_t old = en_push(r);
    while (!r->err) {
        uint32_t tag = en_read_tag(r);
        if (r->err || tag == 0) break;
        int known = 1;
        if (which == 1) {
            if (tag == 10) en_string(r, &t.name_off, &t.name_len);
            else if (tag == 17) t.value_bits = en_fixed64(r);
            else if (tag == 26) parse_event(r, &t.ev);
            else known = 0;
        } else if (which == 2) {
            if (tag == 8) t.source = en_varint32(r);
            else if (tag == 16) t.level = en_varint32(r);
            else if (tag == 26) en_string(r, &t.name_off, &t.name_len);
            else if (tag == 34) en_string(r, &t.msg_off, &t.msg_len);
            else if (tag == 42) parse_event(r, &t.ev);
            else known = 0;
        } else {
            if (tag == 10) parse_odbl(r, &t.lat);
            else if (tag == 18) parse_odbl(r, &t.lon);
            else if (tag == 26) parse_odbl(r, &t.elev);
            else if (tag == 34) parse_event(r, &t.ev);
            else known = 0;
        }
        if (!known && !en_skip_field(r, tag, 0)) break;
    }
    en_pop(r, old);
    if (r->err) return;
    t.ocase = which;
    merge_member(a, &t);
}
/* GAnyDeviceEvent, one occurrence merged into `acc` */
static void parse_any(enr *r, g_any *acc) {
    g_any t;
    memset(&t, 0, sizeof t);
    uint64_t old = en_push(r);
    while (!r->err) {
        uint32_t tag = en_read_tag(r);
        if (r->err || tag == 0) break;
        if (tag == 10) parse_member(r, &t, 1);
        else if (tag == 18) parse_member(r, &t, 2);
        else if (tag == 26) parse_member(r, &t, 3);
        else if (tag == 34 || tag == 42 || tag == 50) {
            /* another event type (command invocation / response, state change): checked as a well-formed
             * message at the wire level only (its own strings and sub-messages are not validated here); if it
             * is still the member set at the end, the record is handed back to the stock consumer */
            uint64_t o2 = en_push(r);
            while (!r->err) {
                uint32_t t2 = en_read_tag(r);
                if (r->err || t2 == 0) break;
                if (!en_skip_field(r, t2, 0)) break;
            }
            en_pop(r, o2);
            memset(&t, 0, sizeof t);
            t.ocase = 4;
        } else if (!en_skip_field(r, tag, 0))
            break;
    }
    en_pop(r, old);
    if (r->err || t.ocase == 0) return; /* first occurrence without a member, or mergeFrom(EVENT_NOT_SET): nothing */
    merge_member(acc, &t);
}
/* GDeviceEventContext: validated; only the assignment status can make the conversion throw */
static void parse_context(enr *r, uint32_t *status_acc) {
    g_uuid u;
    g_ostr s;
    uint32_t st = 0;
    uint64_t old = en_push(r);
    while (!r->err) {
        uint32_t tag = en_read_tag(r);
        if (r->err || tag == 0) break;
        memset(&u, 0, sizeof u);
        memset(&s, 0, sizeof s);
        if (tag == 10 || tag == 18 || tag == 26) parse_uuid(r, &u);
        else if (tag == 34) parse_ostr(r, &s);
        else if (tag == 42 || tag == 58) parse_map_entry(r);
        else if (tag == 48) st = en_varint32(r);
        else if (!en_skip_field(r, tag, 0)) break;
    }
    en_pop(r, old);
    if (!r->err && st) *status_acc = st; /* merge: non-default replaces */
}

static double en_bits2d(uint64_t b) {
    double d;
    memcpy(&d, &b, 8);
    return d;
}
static uint64_t en_span(uint64_t off, uint64_t len) { return (off & 0xffffffffu) | (len << 32); }

uint64_t orc_decode_enriched(const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint64_t max_rows,
                             const orc_cols *out, const orc_event_ids *ids, uint8_t *status, uint32_t *first_row) {
    uint64_t row = 0;
    for (uint32_t i = 0; i < n; i++) {
        first_row[i] = (uint32_t)row;
        const uint8_t *p = bytes + rec_off[i];
        uint64_t total = rec_off[i + 1] - rec_off[i];
        /* parseFrom(byte[]): CodedInputStream.newInstance(buf, off, len) pushes `len` as the outermost limit, so
         * a nested length that reaches past the end of the record is truncatedMessage at its pushLimit */
        enr r = {p, 0, total, total, 0, 0};
        g_any ev;
        memset(&ev, 0, sizeof ev);
        uint32_t asg_status = 0;
        while (!r.err) {
            uint32_t tag = en_read_tag(&r);
            if (r.err || tag == 0) break;
            if (tag == 10) parse_context(&r, &asg_status);
            else if (tag == 18) parse_any(&r, &ev);
            else if (!en_skip_field(&r, tag, 0)) break;
        }
        if (!r.err && r.last_tag != 0) r.err = 1; /* top level: checkLastTagWas(0) */
        if (r.err) {
            status[i] = ST_DECODE_EXCEPTION;
            continue;
        }
        if (ev.ocase == 4) { /* a command invocation / response or state change: hand the record back */
            status[i] = ST_PASSTHROUGH;
            continue;
        }
        /* conversion: context first (assignment status), then the event */
        if (asg_status > 3 || ev.ocase == 0 || ev.ev.etype > 5 || (ev.ocase == 2 && (ev.source > 1 || ev.level > 3))) {
            status[i] = ST_DECODE_EXCEPTION; /* SiteWhereException: unknown enum value / no event set */
            continue;
        }
        status[i] = ST_OK;
        uint32_t kind = ev.ocase == 1 ? KIND_MEASUREMENT : (ev.ocase == 2 ? KIND_ALERT : KIND_LOCATION);
        uint32_t fl = 0;
        if (ev.ev.date) fl |= F_HAS_DATE;
        if (ev.ev.md_count) fl |= F_HAS_METADATA;
        if (ev.ev.alt.has) fl |= F_HAS_ALT_ID;
        double v0 = 0, v1 = 0, v2 = 0;
        uint64_t nspan = 0, nh = 0, mspan = 0;
        if (kind == KIND_MEASUREMENT) {
            fl |= F_HAS_NAME | F_HAS_V0;
            v0 = en_bits2d(ev.value_bits);
        } else if (kind == KIND_ALERT) {
            fl |= F_HAS_NAME | F_HAS_MESSAGE;
            mspan = en_span(ev.msg_off, ev.msg_len);
        } else {
            if (ev.lat.has) fl |= F_HAS_V0, v0 = en_bits2d(ev.lat.bits);
            if (ev.lon.has) fl |= F_HAS_V1, v1 = en_bits2d(ev.lon.bits);
            if (ev.elev.has) fl |= F_HAS_V2, v2 = en_bits2d(ev.elev.bits);
        }
        if (kind != KIND_LOCATION) {
            nspan = en_span(ev.name_off, ev.name_len);
            nh = orc_hash64(p + ev.name_off, (uint32_t)ev.name_len);
        }
        /* the state key's "token" is the 16 bytes of the device assignment id (msb, lsb little-endian) */
        uint8_t key[16];
        for (int k = 0; k < 8; k++) {
            key[k] = (uint8_t)(ev.ev.assignment.msb >> (8 * k));
            key[8 + k] = (uint8_t)(ev.ev.assignment.lsb >> (8 * k));
        }
        if (row < max_rows) {
            out->meta[row] = (uint64_t)(kind & 0xff) | ((uint64_t)(fl & 0xffff) << 8) |
                             ((uint64_t)(kind == KIND_ALERT ? ev.level & 3 : 0) << 24) |
                             ((uint64_t)(kind == KIND_ALERT ? ev.source & 1 : 0) << 26) | (1ull << 27) |
                             ((uint64_t)(ev.ev.etype & 7) << 28) | ((uint64_t)i << 32);
            out->token_hash[row] = orc_hash64(key, 16);
            out->name_hash[row] = nh;
            out->event_date[row] = (int64_t)ev.ev.date;
            out->v0[row] = v0;
            out->v1[row] = v1;
            out->v2[row] = v2;
            out->token_span[row] = 0;
            out->name_span[row] = nspan;
            if (out->originator_span) out->originator_span[row] = 0;
            if (out->message_span) out->message_span[row] = mspan;
            if (out->metadata_span) out->metadata_span[row] = ev.ev.md_count ? en_span(ev.ev.md_first, ev.ev.md_count) : 0;
            if (out->alternate_id_span) out->alternate_id_span[row] = ev.ev.alt.has ? en_span(ev.ev.alt.off, ev.ev.alt.len) : 0;
            if (ids) {
                ids->event_hi[row] = ev.ev.id.msb;
                ids->event_lo[row] = ev.ev.id.lsb;
                ids->assignment_hi[row] = ev.ev.assignment.msb;
                ids->assignment_lo[row] = ev.ev.assignment.lsb;
                ids->device_hi[row] = ev.ev.device.msb;
                ids->device_lo[row] = ev.ev.device.lsb;
            }
        }
        row++;
    }
    first_row[n] = (uint32_t)row;
    return row;
}
