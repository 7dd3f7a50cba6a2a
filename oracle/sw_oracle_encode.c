/*
 * sw_oracle_encode.c — CPU oracle, part 4: the outbound payload a decoded request is forwarded as.
 * TEST INFRASTRUCTURE ONLY (see sw_oracle.h).  PARITY UNPINNED BY THE REFERENCE.
 *
 * Reference (paths relative to the SiteWhere tree; GEV = sitewhere-grpc-event-management/src/main):
 *   - EventSourcesManager.handleDecodedEvent (service-event-sources/.../sources/EventSourcesManager.java:152-166): for a
 *     decoded request that is an IDeviceEventCreateRequest, an InboundEventPayload {sourceId, deviceToken, originator,
 *     eventCreateRequest} is marshalled by EventModelMarshaler.buildInboundEventPayloadMessage
 *     (GEV/java/com/sitewhere/grpc/client/event/EventModelMarshaler.java:93-115) and sent to the decoded-events topic,
 *     keyed by the device token.  A null request is not an instance of anything: nothing is sent.
 *   - conversion: EventModelConverter.asGrpcInboundEventPayload :1878-1887, asGrpcDeviceEventCreateRequest :1621-1653
 *     (switch on getEventType(), then a cast), asGrpcDeviceMeasurementCreateRequest :399-406,
 *     asGrpcDeviceAlertCreateRequest :623-632, asGrpcDeviceLocationCreateRequest :770-784,
 *     createGrpcDeviceEventCreateRequest :196-211, asGrpcAlertSource :544-552, asGrpcAlertLevel :584-596;
 *     CommonModelConverter.asGrpcDate :71-76 (null -> 0)
 *   - schema: GEV/proto/device-event-model.proto:34-42 (GDeviceEventCreateRequest), :45-49, :66-72, :75-80,
 *     :111-120 (GAnyDeviceEventCreateRequest), :307-312 (GInboundEventPayload)
 *
 * protobuf-java 3.3.0 (third party, recalled) writeTo for proto3: fields in field-number order; a scalar equal to its
 * default is not written (string empty, integer / enum 0, double `value != 0D` false — so -0.0 is NOT written); a
 * message-typed field that was set is written even when it is empty (tag, length 0); a builder setter given null
 * throws NullPointerException, and so does unboxing a null Double or a switch on a null enum.  Strings are written as
 * the UTF-8 of the Java String.
 *
 * What is restated: the byte string for a row whose strings are plain ASCII taken verbatim from the payload.  Rows the
 * bytes of which depend on Java-side state are handed back (status 3): a metadata map (HashMap iteration order), a
 * string with JSON escapes or coerced from a non-string token, any string byte >= 0x80 (the String round trip rewrites
 * invalid UTF-8).
 */
#include <string.h>

#include "sw_oracle.h"

#define F_UPDATE_STATE (1u << 0)
#define F_HAS_DATE (1u << 1)
#define F_HAS_ORIGINATOR (1u << 2)
#define F_HAS_METADATA (1u << 3)
#define F_HAS_ALT_ID (1u << 4)
#define F_NULL_TOKEN (1u << 5)
#define F_HAS_V0 (1u << 6)
#define F_HAS_V1 (1u << 7)
#define F_HAS_V2 (1u << 8)
#define F_STR_ESCAPED (1u << 9)
#define F_HAS_NAME (1u << 10)
#define F_HAS_MESSAGE (1u << 11)
#define F_NAME_COERCED (1u << 12)
#define F_NULL_REQUEST (1u << 13)
#define F_NULL_LEVEL (1u << 14)
#define F_NULL_SOURCE (1u << 15)

typedef struct {
    uint8_t *p;     /* NULL = only measure */
    uint64_t n;
} wr;
static void put(wr *w, uint8_t b) {
    if (w->p) w->p[w->n] = b;
    w->n++;
}
static void put_varint(wr *w, uint64_t v) {
    while (v >= 0x80) {
        put(w, (uint8_t)(v | 0x80));
        v >>= 7;
    }
    put(w, (uint8_t)v);
}
static void put_bytes(wr *w, const uint8_t *s, uint64_t len) {
    if (w->p && len) memcpy(w->p + w->n, s, len);
    w->n += len;
}
static void put_f64(wr *w, double v) {
    uint64_t b;
    memcpy(&b, &v, 8);
    for (int k = 0; k < 8; k++) put(w, (uint8_t)(b >> (8 * k)));
}
static uint32_t varint_len(uint64_t v) {
    uint32_t n = 1;
    while (v >= 0x80) {
        v >>= 7;
        n++;
    }
    return n;
}
/* a length-delimited field holding a string */
static void put_str(wr *w, uint8_t tag, const uint8_t *s, uint64_t len) {
    if (!len) return; /* proto3: the empty string is the default */
    put(w, tag);
    put_varint(w, len);
    put_bytes(w, s, len);
}
/* GOptionalString{value = s} as field `tag` (written even when s is empty: the message was set) */
static void put_opt_str(wr *w, uint8_t tag, const uint8_t *s, uint64_t len) {
    put(w, tag);
    put_varint(w, len ? 1 + varint_len(len) + len : 0);
    put_str(w, 0x0A, s, len);
}
/* GOptionalDouble{value = v} as field `tag` */
static void put_opt_f64(wr *w, uint8_t tag, double v) {
    put(w, tag);
    if (v != 0.0) { /* (-0.0 != 0D is false too) */
        put(w, 9);
        put(w, 0x09);
        put_f64(w, v);
    } else
        put(w, 0);
}

typedef struct {
    const uint8_t *tok, *name, *msg, *orig, *alt;
    uint64_t tok_len, name_len, msg_len, orig_len, alt_len;
    uint32_t kind, flags, level, source;
    int64_t date; /* what asGrpcDate gives */
    double v0, v1, v2;
} row;

/* GDeviceEventCreateRequest */
static void put_event(wr *w, const row *r) {
    if (r->flags & F_HAS_ALT_ID) put_opt_str(w, 0x0A, r->alt, r->alt_len);
    if (r->date != 0) {
        put(w, 0x28);
        put_varint(w, (uint64_t)r->date);
    }
    if (r->flags & F_UPDATE_STATE) {
        put(w, 0x32);
        put(w, 2);
        put(w, 0x08);
        put(w, 1);
    }
}
static void put_msg(wr *w, uint8_t tag, void (*body)(wr *, const row *), const row *r) {
    wr m = {NULL, 0};
    body(&m, r);
    put(w, tag);
    put_varint(w, m.n);
    body(w, r);
}
static void body_measurement(wr *w, const row *r) {
    put_str(w, 0x0A, r->name, r->name_len);
    if (r->v0 != 0.0) {
        put(w, 0x11);
        put_f64(w, r->v0);
    }
    put_msg(w, 0x1A, put_event, r);
}
static void body_alert(wr *w, const row *r) {
    if (r->source) {
        put(w, 0x08);
        put(w, (uint8_t)r->source);
    }
    if (r->level) {
        put(w, 0x10);
        put(w, (uint8_t)r->level);
    }
    put_str(w, 0x1A, r->name, r->name_len);
    put_str(w, 0x22, r->msg, r->msg_len);
    put_msg(w, 0x2A, put_event, r);
}
static void body_location(wr *w, const row *r) {
    if (r->flags & F_HAS_V0) put_opt_f64(w, 0x0A, r->v0);
    if (r->flags & F_HAS_V1) put_opt_f64(w, 0x12, r->v1);
    if (r->flags & F_HAS_V2) put_opt_f64(w, 0x1A, r->v2);
    put_msg(w, 0x22, put_event, r);
}
static void body_any(wr *w, const row *r) {
    if (r->kind == 0)
        put_msg(w, 0x0A, body_measurement, r);
    else if (r->kind == 2)
        put_msg(w, 0x12, body_alert, r);
    else
        put_msg(w, 0x1A, body_location, r);
}

static int ascii(const uint8_t *s, uint64_t n) {
    for (uint64_t i = 0; i < n; i++)
        if (s[i] >= 0x80) return 0;
    return 1;
}

/* format: 0 = rows of orc_decode_pb (a missing event date is "now"), 1 = rows of orc_decode_json (it is null -> 0).
 * enc_status: 0 written, 1 the reference throws (NullPointerException / ClassCastException), 3 handed back,
 * 4 nothing to send (the request is null).  out_off[n_rows + 1]; returns the total length (written only while it
 * fits out_cap; a row that does not fit leaves the rest unwritten but counted). */
uint64_t orc_encode_inbound(int format, const uint8_t *bytes, const uint64_t *rec_off, const orc_cols *c, uint32_t n_rows,
                            const uint8_t *source_id, uint32_t source_len, int64_t now_ms, uint8_t *out, uint64_t out_cap,
                            uint64_t *out_off, uint8_t *enc_status) {
    uint64_t pos = 0;
    for (uint32_t i = 0; i < n_rows; i++) {
        out_off[i] = pos;
        const uint64_t m = c->meta[i];
        row r;
        memset(&r, 0, sizeof r);
        r.kind = (uint32_t)(m & 0xff);
        r.flags = (uint32_t)((m >> 8) & 0xffff);
        r.level = (uint32_t)((m >> 24) & 3);
        r.source = (uint32_t)((m >> 26) & 1);
        const uint32_t has_etype = (uint32_t)((m >> 27) & 1), etype = (uint32_t)((m >> 28) & 7);
        const uint8_t *rec = bytes + rec_off[m >> 32];
#define SPAN(col, ptr, len)                                 \
    do {                                                    \
        const uint64_t s_ = c->col ? c->col[i] : 0;         \
        ptr = rec + (uint32_t)s_;                           \
        len = s_ >> 32;                                     \
    } while (0)
        SPAN(token_span, r.tok, r.tok_len);
        SPAN(name_span, r.name, r.name_len);
        SPAN(message_span, r.msg, r.msg_len);
        SPAN(originator_span, r.orig, r.orig_len);
        SPAN(alternate_id_span, r.alt, r.alt_len);
#undef SPAN
        r.v0 = c->v0[i];
        r.v1 = c->v1[i];
        r.v2 = c->v2[i];
        r.date = (r.flags & F_HAS_DATE) ? c->event_date[i] : (format == 0 ? now_ms : 0);
        uint8_t st = 0;
        if (r.flags & F_NULL_REQUEST)
            st = 4;
        else if (!has_etype || etype != r.kind || (r.flags & F_NULL_TOKEN))
            st = 1; /* switch on null / cast to another request class / setDeviceToken(null) */
        else if (r.kind == 0 && (!(r.flags & F_HAS_NAME) || !(r.flags & F_HAS_V0)))
            st = 1; /* setName(null) / unboxing a null Double */
        else if (r.kind == 2 && (!(r.flags & F_HAS_NAME) || !(r.flags & F_HAS_MESSAGE) ||
                                 (r.flags & (F_NULL_LEVEL | F_NULL_SOURCE))))
            st = 1;
        else if ((r.flags & (F_HAS_METADATA | F_STR_ESCAPED | F_NAME_COERCED)) || !ascii(r.tok, r.tok_len) ||
                 ((r.flags & F_HAS_NAME) && r.kind != 1 && !ascii(r.name, r.name_len)) ||
                 ((r.flags & F_HAS_MESSAGE) && r.kind == 2 && !ascii(r.msg, r.msg_len)) ||
                 ((r.flags & F_HAS_ORIGINATOR) && !ascii(r.orig, r.orig_len)) ||
                 ((r.flags & F_HAS_ALT_ID) && !ascii(r.alt, r.alt_len)))
            st = 3;
        enc_status[i] = st;
        if (st) continue;
        if (r.kind == 1) r.name_len = 0;
        wr any = {NULL, 0};
        body_any(&any, &r);
        wr w = {NULL, 0};
        for (int pass = 0; pass < 2; pass++) {
            w.n = 0;
            put_str(&w, 0x0A, source_id, source_len);
            put_str(&w, 0x12, r.tok, r.tok_len);
            if (r.flags & F_HAS_ORIGINATOR) put_opt_str(&w, 0x1A, r.orig, r.orig_len);
            put(&w, 0x22);
            put_varint(&w, any.n);
            body_any(&w, &r);
            if (pass == 0) {
                if (pos + w.n > out_cap) break; /* measured only */
                w.p = out + pos;
            }
        }
        pos += w.n;
    }
    out_off[n_rows] = pos;
    return pos;
}
