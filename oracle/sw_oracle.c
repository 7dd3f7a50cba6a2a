/*
 * sw_oracle.c — CPU oracle, part 1: hashing/partitioning, protobuf decode, zone test, state fold.
 * TEST INFRASTRUCTURE ONLY (see sw_oracle.h).  PARITY UNPINNED BY THE REFERENCE (see sw_oracle.h).
 *
 * Reference paths (relative to the SiteWhere tree):
 *   ES/ = service-event-sources/src/main/java/com/sitewhere/sources/
 *   RP/ = service-rule-processing/src/main/java/com/sitewhere/rules/
 *   DS/ = service-device-state/src/main/java/com/sitewhere/devicestate/
 */
#include "sw_oracle.h"

#include <stdlib.h>
#include <string.h>

/* ---- meta word layout (mirrors include/swgpu.h) ---- */
#define KIND_MEASUREMENT 0u
#define KIND_LOCATION 1u
#define KIND_ALERT 2u
#define F_UPDATE_STATE (1u << 0)
#define F_HAS_DATE (1u << 1)
#define F_HAS_ORIGINATOR (1u << 2)
#define F_HAS_METADATA (1u << 3)
#define F_NULL_TOKEN (1u << 5)
#define F_HAS_V0 (1u << 6)
#define F_HAS_V1 (1u << 7)
#define F_HAS_V2 (1u << 8)
#define F_HAS_NAME (1u << 10)
#define F_HAS_MESSAGE (1u << 11)
#define F_NULL_REQUEST (1u << 13)

#define ST_OK 0
#define ST_DECODE_EXCEPTION 1
#define ST_RUNTIME_EXCEPTION 2
#define ST_PASSTHROUGH 3

static uint64_t mk_meta(uint32_t kind, uint32_t flags, uint32_t level, uint32_t source, int has_etype,
                        uint32_t etype, uint32_t rec) {
    return (uint64_t)(kind & 0xff) | ((uint64_t)(flags & 0xffff) << 8) | ((uint64_t)(level & 3) << 24) |
           ((uint64_t)(source & 1) << 26) | ((uint64_t)(has_etype ? 1 : 0) << 27) |
           ((uint64_t)(etype & 7) << 28) | ((uint64_t)rec << 32);
}
static uint64_t mk_span(uint64_t off, uint64_t len) { return (off & 0xffffffffu) | (len << 32); }

/* ====================================================================================== */
/* Hashing.  Kafka's DefaultPartitioner hashes the serialized key with Utils.murmur2 and takes
 * toPositive(h) % numPartitions (kafka-clients 0.11.0.0, third-party, recalled); the reference
 * depends on it by sending keyed records without a partition
 * (sitewhere-microservice/.../kafka/MicroserviceKafkaProducer.java:89-92, key = deviceToken at
 * ES/EventSourcesManager.java:163). */
uint32_t orc_murmur2(const uint8_t *data, uint32_t len) {
    const uint32_t seed = 0x9747b28cu, m = 0x5bd1e995u;
    uint32_t h = seed ^ len;
    uint32_t n4 = len / 4;
    for (uint32_t i = 0; i < n4; i++) {
        uint32_t k = (uint32_t)data[4 * i] | ((uint32_t)data[4 * i + 1] << 8) |
                     ((uint32_t)data[4 * i + 2] << 16) | ((uint32_t)data[4 * i + 3] << 24);
        k *= m;
        k ^= k >> 24;
        k *= m;
        h *= m;
        h ^= k;
    }
    uint32_t base = len & ~3u;
    switch (len & 3u) {
    case 3: h ^= (uint32_t)data[base + 2] << 16; /* fallthrough */
    case 2: h ^= (uint32_t)data[base + 1] << 8;  /* fallthrough */
    case 1:
        h ^= (uint32_t)data[base];
        h *= m;
    }
    h ^= h >> 13;
    h *= m;
    h ^= h >> 15;
    return h;
}
uint32_t orc_partition(const uint8_t *key, uint32_t len, uint32_t n_partitions) {
    return (orc_murmur2(key, len) & 0x7fffffffu) % n_partitions;
}
/* 64-bit row hash: low half is the Kafka hash (so a row's shard is derivable from it); high half an
 * independent hash over the same little-endian 4-byte words (xor, multiply, rotate per word; tail bytes and
 * length folded in last; fmix32 finish).  Not part of the reference; a new column of the columnar form. */
static uint32_t hash_hi_step(uint32_t f, uint32_t word) {
    f = (f ^ word) * 0x9E3779B1u;
    return (f << 15) | (f >> 17);
}
uint64_t orc_hash64(const uint8_t *data, uint32_t len) {
    uint32_t h = 0x811c9dc5u, i = 0, tail = 0;
    for (; i + 4 <= len; i += 4)
        h = hash_hi_step(h, (uint32_t)data[i] | ((uint32_t)data[i + 1] << 8) | ((uint32_t)data[i + 2] << 16) |
                                ((uint32_t)data[i + 3] << 24));
    for (uint32_t k = 0; i + k < len; k++) tail |= (uint32_t)data[i + k] << (8 * k);
    h = hash_hi_step(h, tail ^ (len << 24));
    h ^= h >> 16;
    h *= 0x85ebca6bu;
    h ^= h >> 13;
    h *= 0xc2b2ae35u;
    h ^= h >> 16;
    return ((uint64_t)h << 32) | orc_murmur2(data, len);
}

/* ====================================================================================== */
/* Protobuf decode: ProtobufDeviceEventDecoder.decode, ES/decoder/protobuf/ProtobufDeviceEventDecoder.java:75-304,
 * schema sitewhere-communication/src/main/proto/sitewhere.proto:1-109, parsed with the semantics of
 * protobuf-java 3.3.0 lite (third-party, recalled):
 *   - parseDelimitedFrom(InputStream): null at EOF; length = readRawVarint32(firstByte, stream);
 *     the message is parsed from a LimitedInputStream(stream, length) — a declared length larger
 *     than what is left simply ends at EOF; a negative length is an empty stream.
 *   - readTag(): 0 at end of input; a tag whose field number is 0 is invalidTag.
 *   - varints: at most 10 bytes, upper bits dropped; 32-bit reads are the low 32 bits.
 *   - generated parse loop switches on the whole tag (field number AND wire type); anything else
 *     goes to parseUnknownField: END_GROUP stops the message (then checkLastTagWas(0) fails),
 *     other wire types are skipped, groups recursively, wire types 6/7 are invalidWireType.
 *   - nested messages: readMessage = readRawVarint32 length, pushLimit (negative -> negativeSize;
 *     the top-level limit is Integer.MAX_VALUE so an over-long nested length is only bounded by
 *     the stream), parse to limit/EOF, checkLastTagWas(0).
 *   - last occurrence of a scalar wins; required fields are checked after the parse
 *     (UninitializedMessageException -> InvalidProtocolBufferException).
 */
typedef struct {
    const uint8_t *p; /* record base */
    uint64_t pos;     /* absolute position in record */
    uint64_t end;     /* end of the (limited) stream */
    uint64_t limit;   /* current pushLimit (<= end or beyond it) */
    int err;          /* sticky: any InvalidProtocolBufferException */
    uint32_t last_tag;
} pbr;

static int pb_at_end(pbr *r) { return r->pos >= r->end || r->pos >= r->limit; }

static int pb_byte(pbr *r, uint8_t *b) {
    if (r->pos >= r->end || r->pos >= r->limit) {
        r->err = 1; /* truncatedMessage */
        return 0;
    }
    *b = r->p[r->pos++];
    return 1;
}
static uint64_t pb_varint64(pbr *r) {
    uint64_t v = 0;
    for (int shift = 0; shift < 64; shift += 7) {
        uint8_t b;
        if (!pb_byte(r, &b)) return 0;
        v |= (uint64_t)(b & 0x7f) << shift;
        if (!(b & 0x80)) return v;
    }
    r->err = 1; /* malformedVarint */
    return 0;
}
static uint32_t pb_varint32(pbr *r) { return (uint32_t)pb_varint64(r); }
static uint64_t pb_fixed64(pbr *r) {
    uint64_t v = 0;
    for (int i = 0; i < 8; i++) {
        uint8_t b;
        if (!pb_byte(r, &b)) return 0;
        v |= (uint64_t)b << (8 * i);
    }
    return v;
}
static void pb_skip(pbr *r, uint64_t n) {
    uint64_t lim = r->end < r->limit ? r->end : r->limit;
    if (n > lim - r->pos) {
        r->pos = lim;
        r->err = 1;
        return;
    }
    r->pos += n;
}
/* readString / readBytes: returns span start, sets *len */
static uint64_t pb_lenfield(pbr *r, uint64_t *len) {
    int32_t size = (int32_t)pb_varint32(r);
    *len = 0;
    if (r->err) return 0;
    if (size < 0) {
        r->err = 1; /* negativeSize */
        return 0;
    }
    uint64_t start = r->pos;
    pb_skip(r, (uint64_t)size);
    if (r->err) return 0;
    *len = (uint64_t)size;
    return start;
}
static uint32_t pb_read_tag(pbr *r) {
    if (pb_at_end(r)) {
        r->last_tag = 0;
        return 0;
    }
    uint32_t t = pb_varint32(r);
    if (r->err) return 0;
    r->last_tag = t;
    if ((t >> 3) == 0) {
        r->err = 1; /* invalidTag */
        return 0;
    }
    return t;
}
#define PB_MAX_GROUP_DEPTH 64
/* UnknownFieldSetLite.mergeFieldFrom; returns 0 if the field was END_GROUP (caller stops) */
static int pb_skip_unknown(pbr *r, uint32_t tag, int depth) {
    switch (tag & 7) {
    case 0: pb_varint64(r); return 1;
    case 5: pb_skip(r, 4); return 1;
    case 1: pb_skip(r, 8); return 1;
    case 2: {
        uint64_t len;
        pb_lenfield(r, &len);
        return 1;
    }
    case 3: {
        if (depth >= PB_MAX_GROUP_DEPTH) { /* the JVM would recurse further; documented bound */
            r->err = 1;
            return 1;
        }
        for (;;) {
            uint32_t t = pb_read_tag(r);
            if (r->err) return 1;
            if (t == 0) break;
            if (!pb_skip_unknown(r, t, depth + 1)) break;
            if (r->err) return 1;
        }
        /* checkLastTagWas(makeTag(fieldNumber, END_GROUP)) */
        if (r->last_tag != (((tag >> 3) << 3) | 4)) r->err = 1;
        return 1;
    }
    case 4: return 0;
    default: r->err = 1; return 1; /* invalidWireType */
    }
}
/* generated-code default branch: parseUnknownField; returns 0 when the message must stop */
static int pb_unknown(pbr *r, uint32_t tag) {
    if ((tag & 7) == 4) return 0;
    return pb_skip_unknown(r, tag, 0);
}

/* Model.Metadata { required string name = 1; required string value = 2; }  sitewhere.proto:9-12 */
static void pb_parse_metadata(pbr *r) {
    int32_t length = (int32_t)pb_varint32(r);
    if (r->err) return;
    if (length < 0) {
        r->err = 1;
        return;
    }
    uint64_t old = r->limit;
    uint64_t nl = r->pos + (uint64_t)length;
    if (nl > old) { /* pushLimit: beyond the enclosing limit (Integer.MAX_VALUE at top level) */
        r->err = 1;
        return;
    }
    r->limit = nl;
    int has = 0;
    for (;;) {
        uint32_t tag = pb_read_tag(r);
        if (r->err) return;
        if (tag == 0) break;
        uint64_t len;
        if (tag == 10) {
            pb_lenfield(r, &len);
            has |= 1;
        } else if (tag == 18) {
            pb_lenfield(r, &len);
            has |= 2;
        } else if (!pb_unknown(r, tag))
            break;
        if (r->err) return;
    }
    if (r->last_tag != 0) r->err = 1; /* checkLastTagWas(0) */
    if (has != 3) r->err = 1;         /* required */
    r->limit = old;
}

typedef struct {
    int present;
    uint64_t off, len;
} pbstr;

typedef struct {
    pbstr name;
    double value;
    int has_value;
} pbmx;

/* parseDelimitedFrom: returns 0 -> null (stream exhausted), 1 -> a limited reader in *r */
static int pb_open_delimited(const uint8_t *p, uint64_t *pos, uint64_t total, pbr *r) {
    if (*pos >= total) return 0;
    pbr hdr = {p, *pos, total, UINT64_MAX, 0, 0};
    /* readRawVarint32(firstByte, input): same 10-byte rule, truncation is an IOException */
    int32_t size = (int32_t)pb_varint32(&hdr);
    r->p = p;
    r->err = hdr.err;
    r->last_tag = 0;
    r->limit = 0x7fffffffull + hdr.pos; /* top-level currentLimit = Integer.MAX_VALUE */
    r->pos = hdr.pos;
    if (hdr.err) {
        r->end = hdr.pos;
        *pos = total;
        return 1;
    }
    uint64_t avail = total - hdr.pos;
    uint64_t take = size < 0 ? 0 : ((uint64_t)size < avail ? (uint64_t)size : avail);
    r->end = hdr.pos + take;
    *pos = r->end; /* a successful parse consumes the whole limited stream */
    return 1;
}

static void emit_row(const orc_cols *o, uint64_t row, uint64_t max_rows, uint64_t meta, uint64_t th,
                     uint64_t nh, int64_t date, double v0, double v1, double v2, uint64_t tspan,
                     uint64_t nspan, uint64_t ospan, uint64_t mspan, uint64_t mdspan, uint64_t aspan) {
    if (row >= max_rows) return;
    o->meta[row] = meta;
    o->token_hash[row] = th;
    o->name_hash[row] = nh;
    o->event_date[row] = date;
    o->v0[row] = v0;
    o->v1[row] = v1;
    o->v2[row] = v2;
    o->token_span[row] = tspan;
    o->name_span[row] = nspan;
    if (o->originator_span) o->originator_span[row] = ospan;
    if (o->message_span) o->message_span[row] = mspan;
    if (o->metadata_span) o->metadata_span[row] = mdspan;
    if (o->alternate_id_span) o->alternate_id_span[row] = aspan;
}

static double bits2d(uint64_t b) {
    double d;
    memcpy(&d, &b, 8);
    return d;
}

#define PB_MAX_MX 4096 /* oracle-side bound on measurements per message (heap-grown) */

uint64_t orc_decode_pb(const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint64_t max_rows,
                       const orc_cols *out, uint8_t *status, uint32_t *first_row) {
    uint64_t row = 0;
    pbmx *mx = NULL;
    size_t mx_cap = 0;
    for (uint32_t i = 0; i < n; i++) {
        const uint8_t *p = bytes + rec_off[i];
        uint64_t total = rec_off[i + 1] - rec_off[i];
        first_row[i] = (uint32_t)row;
        uint64_t pos = 0;
        pbr r;
        /* Header header = SiteWhere.Header.parseDelimitedFrom(stream);   :79 */
        if (!pb_open_delimited(p, &pos, total, &r)) {
            status[i] = ST_RUNTIME_EXCEPTION; /* header == null -> NPE at header.getCommand() :81 */
            continue;
        }
        int has_cmd = 0;
        uint32_t cmd = 0;
        pbstr orig = {0, 0, 0};
        while (!r.err) {
            uint32_t tag = pb_read_tag(&r);
            if (r.err || tag == 0) break;
            if (tag == 8) { /* required Command command = 1  (sitewhere.proto:85) */
                uint32_t raw = pb_varint32(&r);
                if (!r.err && raw >= 1 && raw <= 8) { /* Command.forNumber; unknown -> unknown field */
                    has_cmd = 1;
                    cmd = raw;
                }
            } else if (tag == 18) { /* optional string originator = 2 */
                orig.off = pb_lenfield(&r, &orig.len);
                orig.present = 1;
            } else if (!pb_unknown(&r, tag))
                break;
        }
        if (!r.err && r.last_tag != 0) r.err = 1;
        if (!r.err && !has_cmd) r.err = 1;
        if (r.err) {
            status[i] = ST_DECODE_EXCEPTION;
            continue;
        }
        if (cmd != 3 && cmd != 4 && cmd != 5) { /* registration, ack, streams: not events */
            status[i] = ST_PASSTHROUGH;
            continue;
        }
        /* Xxx.parseDelimitedFrom(stream)   :124 / :160 / :194 */
        if (!pb_open_delimited(p, &pos, total, &r)) {
            status[i] = ST_RUNTIME_EXCEPTION; /* message == null -> NPE at getHardwareId() */
            continue;
        }
        pbstr hw = {0, 0, 0}, atype = {0, 0, 0}, amsg = {0, 0, 0};
        int has_lat = 0, has_lon = 0, has_date = 0, has_us = 0, us = 0;
        double lat = 0, lon = 0, elev = 0;
        int64_t date = 0;
        uint64_t md_first = 0, md_count = 0;
        size_t nmx = 0;
        /* field tags per message (sitewhere.proto:15-48) */
        uint32_t t_date, t_md, t_us;
        if (cmd == 3) {
            t_date = 41;
            t_md = 50;
            t_us = 56;
        } else if (cmd == 4) {
            t_date = 33;
            t_md = 42;
            t_us = 48;
        } else {
            t_date = 25;
            t_md = 34;
            t_us = 40;
        }
        while (!r.err) {
            uint64_t tag_pos = r.pos;
            uint32_t tag = pb_read_tag(&r);
            if (r.err || tag == 0) break;
            if (tag == 10) {
                hw.off = pb_lenfield(&r, &hw.len);
                hw.present = 1;
            } else if (tag == t_date) {
                date = (int64_t)pb_fixed64(&r);
                has_date = 1;
            } else if (tag == t_md) {
                if (md_count == 0) md_first = tag_pos;
                md_count++;
                pb_parse_metadata(&r);
            } else if (tag == t_us) {
                us = pb_varint64(&r) != 0;
                has_us = 1;
            } else if (cmd == 3 && tag == 17) {
                lat = bits2d(pb_fixed64(&r));
                has_lat = 1;
            } else if (cmd == 3 && tag == 25) {
                lon = bits2d(pb_fixed64(&r));
                has_lon = 1;
            } else if (cmd == 3 && tag == 33) {
                elev = bits2d(pb_fixed64(&r));
            } else if (cmd == 4 && tag == 18) {
                atype.off = pb_lenfield(&r, &atype.len);
                atype.present = 1;
            } else if (cmd == 4 && tag == 26) {
                amsg.off = pb_lenfield(&r, &amsg.len);
                amsg.present = 1;
            } else if (cmd == 5 && tag == 18) {
                /* repeated Measurement { required string measurementId = 1; required double measurementValue = 2; } */
                int32_t length = (int32_t)pb_varint32(&r);
                if (r.err) break;
                if (length < 0) {
                    r.err = 1;
                    break;
                }
                uint64_t old = r.limit, nl = r.pos + (uint64_t)length;
                if (nl > old) {
                    r.err = 1;
                    break;
                }
                r.limit = nl;
                pbmx cur;
                memset(&cur, 0, sizeof cur);
                while (!r.err) {
                    uint32_t t2 = pb_read_tag(&r);
                    if (r.err || t2 == 0) break;
                    if (t2 == 10) {
                        cur.name.off = pb_lenfield(&r, &cur.name.len);
                        cur.name.present = 1;
                    } else if (t2 == 17) {
                        cur.value = bits2d(pb_fixed64(&r));
                        cur.has_value = 1;
                    } else if (!pb_unknown(&r, t2))
                        break;
                }
                if (!r.err && r.last_tag != 0) r.err = 1;
                if (!r.err && (!cur.name.present || !cur.has_value)) r.err = 1;
                r.limit = old;
                if (r.err) break;
                if (nmx == mx_cap) {
                    mx_cap = mx_cap ? mx_cap * 2 : 16;
                    mx = (pbmx *)realloc(mx, mx_cap * sizeof(pbmx));
                }
                mx[nmx++] = cur;
            } else if (!pb_unknown(&r, tag))
                break;
        }
        if (!r.err && r.last_tag != 0) r.err = 1;
        if (!r.err) {
            if (!hw.present) r.err = 1;
            if (cmd == 3 && (!has_lat || !has_lon)) r.err = 1;
            if (cmd == 4 && (!atype.present || !amsg.present)) r.err = 1;
        }
        if (r.err) {
            status[i] = ST_DECODE_EXCEPTION;
            continue;
        }
        status[i] = ST_OK;
        uint32_t flags = 0;
        if (has_us && us) flags |= F_UPDATE_STATE; /* :131-133 / :167-169 / :201-203 */
        if (has_date) flags |= F_HAS_DATE;         /* else new Date() :143-147 */
        if (orig.present) flags |= F_HAS_ORIGINATOR;
        if (md_count) flags |= F_HAS_METADATA;
        uint64_t th = orc_hash64(p + hw.off, (uint32_t)hw.len);
        uint64_t tspan = mk_span(hw.off, hw.len);
        uint64_t ospan = orig.present ? mk_span(orig.off, orig.len) : 0;
        uint64_t mdspan = md_count ? mk_span(md_first, md_count) : 0;
        if (cmd == 3) {
            /* setElevation(location.getElevation()): absent -> proto default 0.0   :165 */
            emit_row(out, row++, max_rows,
                     mk_meta(KIND_LOCATION, flags | F_HAS_V0 | F_HAS_V1 | F_HAS_V2, 0, 0, 1, KIND_LOCATION, i),
                     th, 0, date, lat, lon, elev, tspan, 0, ospan, 0, mdspan, 0);
        } else if (cmd == 4) {
            /* level forced to Info :199; source stays Device */
            emit_row(out, row++, max_rows,
                     mk_meta(KIND_ALERT, flags | F_HAS_NAME | F_HAS_MESSAGE, 0, 0, 1, KIND_ALERT, i), th,
                     orc_hash64(p + atype.off, (uint32_t)atype.len), date, 0, 0, 0, tspan,
                     mk_span(atype.off, atype.len), ospan, mk_span(amsg.off, amsg.len), mdspan, 0);
        } else {
            for (size_t k = 0; k < nmx; k++) /* one request per Measurement :126-157 */
                emit_row(out, row++, max_rows,
                         mk_meta(KIND_MEASUREMENT, flags | F_HAS_NAME | F_HAS_V0, 0, 0, 1, KIND_MEASUREMENT, i),
                         th, orc_hash64(p + mx[k].name.off, (uint32_t)mx[k].name.len), date, mx[k].value,
                         0, 0, tspan, mk_span(mx[k].name.off, mx[k].name.len), ospan, 0, mdspan, 0);
        }
    }
    first_row[n] = (uint32_t)row;
    free(mx);
    return row;
}

/* ====================================================================================== */
/* Zone test.  ZoneTestRuleProcessor.onLocation, RP/processors/geospatial/ZoneTestRuleProcessor.java:47-63;
 * polygon/point construction GeoUtils.java:34-37,55-67 (x = longitude, y = latitude, ring closed by
 * repeating the first vertex, no holes).  poly.contains(point) is JTS 1.13 (third-party, recalled):
 *   Geometry.contains: envelope short-circuit; isRectangle() -> RectangleContains; else
 *   relate(g).isContains(), which for a point is PointLocator -> RayCrossingCounter.locatePointInRing
 *   == INTERIOR, with the crossing side decided by RobustDeterminant.signOfDet2x2 on coordinates
 *   translated (in double) so that the point is the origin.
 */

/* exact sign of x1*y2 - y1*x2 for finite doubles (what RobustDeterminant.signOfDet2x2 returns),
 * here by integer arithmetic on the mantissas */
typedef unsigned __int128 u128;
static void dsplit(double d, int *sgn, uint64_t *mant, int *exp) {
    uint64_t b;
    memcpy(&b, &d, 8);
    *sgn = (int)(b >> 63);
    int e = (int)((b >> 52) & 0x7ff);
    uint64_t m = b & 0xfffffffffffffull;
    if (e == 0) {
        *mant = m;
        *exp = -1074;
    } else {
        *mant = m | (1ull << 52);
        *exp = e - 1075;
    }
}
static int bitlen128(u128 v) {
    int n = 0;
    while (v) {
        n++;
        v >>= 1;
    }
    return n;
}
/* compare a*2^ea with b*2^eb, a,b < 2^106 */
static int cmp_scaled(u128 a, int ea, u128 b, int eb) {
    if (a == 0 || b == 0) return (a != 0) - (b != 0);
    int la = bitlen128(a) + ea, lb = bitlen128(b) + eb;
    if (la != lb) return la > lb ? 1 : -1;
    if (ea >= eb)
        a <<= (ea - eb);
    else
        b <<= (eb - ea);
    return a > b ? 1 : (a < b ? -1 : 0);
}
int orc_sign_det2x2(double x1, double y1, double x2, double y2) {
    int s1, s2, s3, s4, e1, e2, e3, e4;
    uint64_t m1, m2, m3, m4;
    dsplit(x1, &s1, &m1, &e1);
    dsplit(y2, &s2, &m2, &e2);
    dsplit(y1, &s3, &m3, &e3);
    dsplit(x2, &s4, &m4, &e4);
    u128 A = (u128)m1 * m2, B = (u128)m3 * m4;
    int sa = A == 0 ? 0 : ((s1 ^ s2) ? -1 : 1);
    int sb = B == 0 ? 0 : ((s3 ^ s4) ? -1 : 1);
    if (sa != sb) return sa > sb ? 1 : -1; /* also covers one of them zero */
    if (sa == 0) return 0;
    int c = cmp_scaled(A, e1 + e2, B, e3 + e4);
    return sa > 0 ? c : -c;
}

int orc_polygon_is_rectangle(const double *xs, const double *ys, uint32_t npts) {
    /* Polygon.isRectangle (JTS 1.13): 5 points, every vertex on envelope corners, x and y
     * change alternately */
    if (npts != 5) return 0;
    double minx = xs[0], maxx = xs[0], miny = ys[0], maxy = ys[0];
    for (uint32_t i = 1; i < 5; i++) {
        if (xs[i] < minx) minx = xs[i];
        if (xs[i] > maxx) maxx = xs[i];
        if (ys[i] < miny) miny = ys[i];
        if (ys[i] > maxy) maxy = ys[i];
    }
    for (uint32_t i = 0; i < 5; i++) {
        if (!(xs[i] == minx || xs[i] == maxx)) return 0;
        if (!(ys[i] == miny || ys[i] == maxy)) return 0;
    }
    double px = xs[0], py = ys[0];
    for (uint32_t i = 1; i <= 4; i++) {
        int xc = xs[i] != px, yc = ys[i] != py;
        if (xc == yc) return 0;
        px = xs[i];
        py = ys[i];
    }
    return 1;
}

static void ring_envelope(const double *xs, const double *ys, uint32_t npts, double *minx, double *maxx,
                          double *miny, double *maxy) {
    /* Envelope.expandToInclude over the coordinate sequence */
    *minx = *maxx = xs[0];
    *miny = *maxy = ys[0];
    for (uint32_t i = 1; i < npts; i++) {
        if (xs[i] < *minx) *minx = xs[i];
        if (xs[i] > *maxx) *maxx = xs[i];
        if (ys[i] < *miny) *miny = ys[i];
        if (ys[i] > *maxy) *maxy = ys[i];
    }
}

/* RayCrossingCounter.locatePointInRing == INTERIOR */
static int ring_interior(const double *xs, const double *ys, uint32_t npts, double px, double py) {
    int crossings = 0;
    for (uint32_t i = 1; i < npts; i++) {
        double p1x = xs[i], p1y = ys[i], p2x = xs[i - 1], p2y = ys[i - 1];
        if (p1x < px && p2x < px) continue;
        if (px == p2x && py == p2y) return 0; /* on a vertex -> BOUNDARY */
        if (p1y == py && p2y == py) {
            double mn = p1x, mxx = p2x;
            if (mn > mxx) {
                mn = p2x;
                mxx = p1x;
            }
            if (px >= mn && px <= mxx) return 0; /* on a horizontal edge -> BOUNDARY */
            continue;
        }
        if ((p1y > py && p2y <= py) || (p2y > py && p1y <= py)) {
            double x1 = p1x - px, y1 = p1y - py, x2 = p2x - px, y2 = p2y - py;
            int s = orc_sign_det2x2(x1, y1, x2, y2);
            if (s == 0) return 0; /* on the edge -> BOUNDARY */
            if (y2 < y1) s = -s;
            if (s > 0) crossings++;
        }
    }
    return (crossings % 2) == 1;
}

static int contains_prepared(const double *xs, const double *ys, uint32_t npts, double minx, double maxx,
                             double miny, double maxy, int is_rect, double px, double py) {
    /* Geometry.contains: Envelope.contains(point envelope) */
    if (!(px >= minx && px <= maxx && py >= miny && py <= maxy)) return 0;
    if (is_rect) /* RectangleContains: inside the envelope and not on its border */
        return !(px == minx || px == maxx || py == miny || py == maxy);
    return ring_interior(xs, ys, npts, px, py);
}

int orc_polygon_contains(const double *xs, const double *ys, uint32_t npts, double px, double py) {
    double minx, maxx, miny, maxy;
    ring_envelope(xs, ys, npts, &minx, &maxx, &miny, &maxy);
    return contains_prepared(xs, ys, npts, minx, maxx, miny, maxy, orc_polygon_is_rectangle(xs, ys, npts),
                             px, py);
}

uint64_t orc_zone_test(uint32_t n_tests, const uint32_t *vtx_off, const double *lon, const double *lat,
                       const uint8_t *cond, const double *plon, const double *plat, const uint64_t *meta,
                       uint32_t n, uint64_t max_pairs, uint64_t *pairs) {
    /* closed rings + per-test envelope/isRectangle, as the cached Polygon objects hold them
     * (ZoneTestRuleProcessor.java:72-84) */
    uint64_t total_pts = 0;
    for (uint32_t t = 0; t < n_tests; t++) total_pts += (uint64_t)(vtx_off[t + 1] - vtx_off[t]) + 1;
    double *cx = (double *)malloc(sizeof(double) * (total_pts + 1));
    double *cy = (double *)malloc(sizeof(double) * (total_pts + 1));
    uint64_t *coff = (uint64_t *)malloc(sizeof(uint64_t) * ((uint64_t)n_tests + 1));
    double *env = (double *)malloc(sizeof(double) * 4 * ((uint64_t)n_tests + 1));
    int *rect = (int *)malloc(sizeof(int) * ((uint64_t)n_tests + 1));
    uint64_t w = 0;
    for (uint32_t t = 0; t < n_tests; t++) {
        coff[t] = w;
        uint32_t a = vtx_off[t], b = vtx_off[t + 1];
        for (uint32_t k = a; k < b; k++) {
            cx[w] = lon[k];
            cy[w] = lat[k];
            w++;
        }
        cx[w] = lon[a];
        cy[w] = lat[a];
        w++;
        uint32_t np = b - a + 1;
        ring_envelope(cx + coff[t], cy + coff[t], np, &env[4 * t], &env[4 * t + 1], &env[4 * t + 2],
                      &env[4 * t + 3]);
        rect[t] = orc_polygon_is_rectangle(cx + coff[t], cy + coff[t], np);
    }
    coff[n_tests] = w;
    uint64_t np_out = 0;
    for (uint32_t i = 0; i < n; i++) {
        if (meta) {
            uint64_t m = meta[i];
            uint32_t fl = (uint32_t)((m >> 8) & 0xffff);
            if ((m & 0xff) != KIND_LOCATION || !(fl & F_HAS_V0) || !(fl & F_HAS_V1)) continue;
        }
        for (uint32_t t = 0; t < n_tests; t++) {
            uint32_t npts = (uint32_t)(coff[t + 1] - coff[t]);
            /* 1 or 2 bounds close to a ring of 2 or 3 points: GeometryFactory.createLinearRing throws
             * IllegalArgumentException (JTS 1.13 LinearRing.validateConstruction: 0 or >= 4 points) from
             * getZonePolygon (ZoneTestRuleProcessor.java:49,72-84; GeoUtils.java:55-67), the polygon is never
             * cached, and onLocation ends there for every location: tests after it never fire.  (The GPU
             * library refuses such a table in swgpu_set_zones.) */
            if (npts < 4) break;
            int inside = contains_prepared(cx + coff[t], cy + coff[t], npts, env[4 * t], env[4 * t + 1],
                                           env[4 * t + 2], env[4 * t + 3], rect[t], plon[i], plat[i]);
            /* containment == test.getCondition()  (Inside = 0, Outside = 1)   :50-53 */
            if ((cond[t] == 0) == (inside != 0)) {
                if (np_out < max_pairs) pairs[np_out] = ((uint64_t)i << 32) | t;
                np_out++;
            }
        }
    }
    free(cx);
    free(cy);
    free(coff);
    free(env);
    free(rect);
    return np_out;
}

/* ====================================================================================== */
/* Device-state fold.  DeviceStateProcessingLogic.processDeviceStateEvent + merge*
 * (DS/processing/DeviceStateProcessingLogic.java:116-201) with deviceStateUpdateLogic
 * (DS/persistence/DeviceStatePersistence.java:60-85): a sequential loop in arrival order;
 * lastLocationEventId := id, lastMeasurementEventIds[name] := id, lastAlertEventIds[type] := id;
 * maps only grow.  The oracle keeps, per (token, kind, name), the last writer's sequence number,
 * event id, event date and value. */
typedef struct {
    uint8_t *key;
    uint32_t klen;
    uint64_t hash, seq, id_hi, id_lo;
    int64_t date;
    double value;
} st_entry;
struct orc_state {
    uint64_t presence_now1; /* 0 = presence not tracked; else the timestamp (ms) of the events being folded + 1 */
    st_entry *e;
    size_t n, cap;
    int64_t *table; /* open addressing -> index into e, -1 empty */
    size_t tcap;
};
orc_state *orc_state_new(void) {
    orc_state *s = (orc_state *)calloc(1, sizeof *s);
    s->tcap = 1024;
    s->table = (int64_t *)malloc(sizeof(int64_t) * s->tcap);
    for (size_t i = 0; i < s->tcap; i++) s->table[i] = -1;
    return s;
}
void orc_state_free(orc_state *s) {
    if (!s) return;
    for (size_t i = 0; i < s->n; i++) free(s->e[i].key);
    free(s->e);
    free(s->table);
    free(s);
}
static uint64_t fnv64(const uint8_t *p, uint32_t n) {
    uint64_t h = 1469598103934665603ull;
    for (uint32_t i = 0; i < n; i++) {
        h ^= p[i];
        h *= 1099511628211ull;
    }
    return h;
}
static void st_grow(orc_state *s) {
    size_t nc = s->tcap * 2;
    int64_t *nt = (int64_t *)malloc(sizeof(int64_t) * nc);
    for (size_t i = 0; i < nc; i++) nt[i] = -1;
    for (size_t i = 0; i < s->n; i++) {
        size_t j = s->e[i].hash & (nc - 1);
        while (nt[j] >= 0) j = (j + 1) & (nc - 1);
        nt[j] = (int64_t)i;
    }
    free(s->table);
    s->table = nt;
    s->tcap = nc;
}
static st_entry *st_get(orc_state *s, const uint8_t *key, uint32_t klen) {
    uint64_t h = fnv64(key, klen);
    size_t j = h & (s->tcap - 1);
    while (s->table[j] >= 0) {
        st_entry *e = &s->e[s->table[j]];
        if (e->hash == h && e->klen == klen && memcmp(e->key, key, klen) == 0) return e;
        j = (j + 1) & (s->tcap - 1);
    }
    if (s->n == s->cap) {
        s->cap = s->cap ? s->cap * 2 : 1024;
        s->e = (st_entry *)realloc(s->e, s->cap * sizeof(st_entry));
    }
    st_entry *e = &s->e[s->n];
    memset(e, 0, sizeof *e);
    e->key = (uint8_t *)malloc(klen ? klen : 1);
    memcpy(e->key, key, klen);
    e->klen = klen;
    e->hash = h;
    s->table[j] = (int64_t)s->n;
    s->n++;
    if (s->n * 2 > s->tcap) st_grow(s);
    return &s->e[s->n - 1];
}
void orc_state_merge(orc_state *s, const uint8_t *bytes, const uint64_t *rec_off, const orc_cols *rows,
                     uint32_t n_rows, uint64_t seq_base, const uint64_t *id_hi, const uint64_t *id_lo) {
    orc_state_merge_seq(s, bytes, rec_off, rows, n_rows, seq_base, NULL, id_hi, id_lo);
}
/* seq == NULL: row r has sequence seq_base + r; else seq_base + seq[r] (rows are in arrival order either way).
 * ids != NULL: enriched rows — the token is the assignment id's 16 bytes and the event id comes from `ids`. */
static void state_merge_any(orc_state *s, const uint8_t *bytes, const uint64_t *rec_off, const orc_cols *rows,
                            const orc_event_ids *ids, uint32_t n_rows, uint64_t seq_base, const uint64_t *seq,
                            const uint64_t *id_hi, const uint64_t *id_lo) {
    uint8_t *kb = NULL;
    size_t kcap = 0;
    for (uint32_t r = 0; r < n_rows; r++) {
        uint64_t m = rows->meta[r];
        uint32_t kind = (uint32_t)(m & 0xff), fl = (uint32_t)((m >> 8) & 0xffff);
        /* only Alert / Location / Measurement touch these fields  :122-127 */
        if (kind > KIND_ALERT) continue;
        if (fl & F_NULL_TOKEN) continue; /* no device -> never reaches device state */
        /* "request": null decodes to a DecodedDeviceRequest with a null request, which EventSourcesManager
         * drops before it is sent on (ES/EventSourcesManager.java:155): it never reaches device state */
        if (fl & F_NULL_REQUEST) continue;
        const uint8_t *rec = bytes + rec_off[m >> 32];
        uint64_t ts = rows->token_span[r], ns = rows->name_span[r];
        uint8_t idkey[16];
        const uint8_t *tok = rec + (uint32_t)ts;
        uint32_t tl = (uint32_t)(ts >> 32), nl = (kind == KIND_LOCATION) ? 0 : (uint32_t)(ns >> 32);
        if (ids) {
            for (int k = 0; k < 8; k++) {
                idkey[k] = (uint8_t)(ids->assignment_hi[r] >> (8 * k));
                idkey[8 + k] = (uint8_t)(ids->assignment_lo[r] >> (8 * k));
            }
            tok = idkey;
            tl = 16;
        }
        if (tl > 0xffff || nl > 0xffff) continue; /* documented bound of the state key */
        size_t kl = (size_t)tl + 3 + nl;
        if (kl > kcap) {
            kcap = kl * 2;
            kb = (uint8_t *)realloc(kb, kcap);
        }
        /* key = token length (2 B LE) | kind | token | name */
        kb[0] = (uint8_t)(tl & 0xff);
        kb[1] = (uint8_t)(tl >> 8);
        kb[2] = (uint8_t)(kind | ((kind != KIND_LOCATION && !(fl & F_HAS_NAME)) ? 0x80 : 0));
        memcpy(kb + 3, tok, tl);
        if (nl) memcpy(kb + 3 + tl, rec + (uint32_t)ns, nl);
        st_entry *e = st_get(s, kb, (uint32_t)kl);
        e->seq = seq_base + (seq ? seq[r] : r); /* sequential: the later event overwrites */
        e->id_hi = ids ? ids->event_hi[r] : id_hi[r];
        e->id_lo = ids ? ids->event_lo[r] : id_lo[r];
        e->date = rows->event_date[r];
        e->value = rows->v0[r];
        if (s->presence_now1) {
            /* every merged event: lastInteractionDate := now, presenceMissingDate := null
             * (DeviceStateProcessingLogic.java:142-143).  The device's row: key = token, kind byte 0x7f, no name;
             * seq holds lastInteractionDate, date holds presenceMissingDate (0 = null). */
            kb[2] = 0x7f;
            st_entry *d = st_get(s, kb, (uint32_t)(3 + tl));
            if (d->seq <= s->presence_now1 - 1) { /* (a batch stamped earlier than one already folded changes nothing) */
                d->seq = s->presence_now1 - 1;
                d->date = 0;
            }
        }
    }
    free(kb);
}
void orc_state_set_presence_now(orc_state *s, uint64_t now_ms_plus1) { s->presence_now1 = now_ms_plus1; }
/* DevicePresenceManager.PresenceChecker (DS/presence/DevicePresenceManager.java:139-167): device states with
 * lastInteractionDate <= now - interval (query: MongoDeviceStateManagement.java:170-193, ascending by that date)
 * for which SendOnceNotificationStrategy permits an event (presenceMissingDate == null,
 * PresenceNotificationStrategies.java:24-31) get presenceMissingDate := now and are reported.  Returns the count;
 * writes up to max_out rows (lastInteractionDate + the device's key bytes), ascending by (date, key). */
static int pres_cmp(const void *a, const void *b) {
    const st_entry *x = *(st_entry *const *)a, *y = *(st_entry *const *)b;
    if (x->seq != y->seq) return x->seq < y->seq ? -1 : 1;
    uint32_t ml = x->klen < y->klen ? x->klen : y->klen;
    int c = memcmp(x->key, y->key, ml);
    if (c) return c;
    return (x->klen > y->klen) - (x->klen < y->klen);
}
uint64_t orc_presence_scan(orc_state *s, uint64_t now_ms, uint64_t interval_ms, uint64_t max_out, uint64_t *last_interaction,
                           uint64_t *key_off, uint8_t *key_bytes) {
    st_entry **hit = (st_entry **)malloc(sizeof(st_entry *) * (s->n ? s->n : 1));
    uint64_t nh = 0;
    const uint64_t cutoff = now_ms - interval_ms;
    for (size_t i = 0; i < s->n; i++) {
        st_entry *e = &s->e[i];
        if (e->klen < 3 || e->key[2] != 0x7f) continue;
        if (e->seq <= cutoff && e->date == 0) hit[nh++] = e;
    }
    qsort(hit, nh, sizeof(st_entry *), pres_cmp);
    uint64_t off = 0;
    for (uint64_t k = 0; k < nh; k++) {
        hit[k]->date = (int64_t)now_ms;
        if (k < max_out) {
            last_interaction[k] = hit[k]->seq;
            key_off[k] = off;
            memcpy(key_bytes + off, hit[k]->key, hit[k]->klen);
            off += hit[k]->klen;
        }
    }
    if (nh <= max_out) key_off[nh] = off;
    else key_off[max_out] = off;
    free(hit);
    return nh;
}
void orc_state_merge_seq(orc_state *s, const uint8_t *bytes, const uint64_t *rec_off, const orc_cols *rows,
                         uint32_t n_rows, uint64_t seq_base, const uint64_t *seq, const uint64_t *id_hi,
                         const uint64_t *id_lo) {
    state_merge_any(s, bytes, rec_off, rows, NULL, n_rows, seq_base, seq, id_hi, id_lo);
}
void orc_state_merge_enriched(orc_state *s, const uint8_t *bytes, const uint64_t *rec_off, const orc_cols *rows,
                              const orc_event_ids *ids, uint32_t n_rows, uint64_t seq_base, const uint64_t *seq) {
    state_merge_any(s, bytes, rec_off, rows, ids, n_rows, seq_base, seq, NULL, NULL);
}
uint64_t orc_state_size(orc_state *s, uint64_t *key_bytes) {
    uint64_t kb = 0;
    for (size_t i = 0; i < s->n; i++) kb += s->e[i].klen;
    if (key_bytes) *key_bytes = kb;
    return s->n;
}
static int st_cmp(const void *a, const void *b) {
    const st_entry *x = (const st_entry *)a, *y = (const st_entry *)b;
    uint32_t ml = x->klen < y->klen ? x->klen : y->klen;
    int c = memcmp(x->key, y->key, ml);
    if (c) return c;
    return (x->klen > y->klen) - (x->klen < y->klen);
}
void orc_state_snapshot(orc_state *s, uint64_t *seq, uint64_t *id_hi, uint64_t *id_lo, int64_t *event_date,
                        double *value, uint64_t *key_off, uint8_t *key_bytes) {
    st_entry *tmp = (st_entry *)malloc(sizeof(st_entry) * (s->n ? s->n : 1));
    memcpy(tmp, s->e, sizeof(st_entry) * s->n);
    qsort(tmp, s->n, sizeof(st_entry), st_cmp);
    uint64_t off = 0;
    for (size_t i = 0; i < s->n; i++) {
        seq[i] = tmp[i].seq;
        id_hi[i] = tmp[i].id_hi;
        id_lo[i] = tmp[i].id_lo;
        event_date[i] = tmp[i].date;
        value[i] = tmp[i].value;
        key_off[i] = off;
        memcpy(key_bytes + off, tmp[i].key, tmp[i].klen);
        off += tmp[i].klen;
    }
    key_off[s->n] = off;
    free(tmp);
}
