/*
 * swgpu.h — C ABI of libswgpu.so, the B200 (sm_100a) implementation of SiteWhere's
 * inbound event hot path: payload decode -> geofence zone test -> device-state merge.
 *
 * This is the boundary a JNI / Panama binding in a `sitewhere-gpu` Java module binds to
 * (see INTEGRATION.md).  Plain pointers and sizes only; no C++ or torch types.
 *
 * Reference paths below are relative to the SiteWhere 2.0.rc2 tree (waffcloud/sitewhere):
 *   ES/ = service-event-sources/src/main/java/com/sitewhere/sources/
 *   RP/ = service-rule-processing/src/main/java/com/sitewhere/rules/
 *   DS/ = service-device-state/src/main/java/com/sitewhere/devicestate/
 *   MS/ = sitewhere-microservice/src/main/java/com/sitewhere/microservice/
 *
 * Conventions
 *   - every function returns int32_t: 0 = ok, <0 = call-level error (text: swgpu_last_error).
 *   - a bad RECORD never fails a call: each record gets a status byte (SWGPU_ST_*), mirroring
 *     the reference's per-record try/catch (ES/InboundEventSource.java:208-215,
 *     DS/processing/DeviceStateProcessingLogic.java:75-85).
 *   - "_dev" entry points take DEVICE pointers, enqueue on the context's stream and do not
 *     synchronise; the others take HOST pointers, copy in/out and return when results are
 *     in the caller's buffers.
 *   - there is no CPU implementation behind this ABI: without a CUDA device every call fails.
 */
#ifndef SWGPU_H
#define SWGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SWGPU_ABI_VERSION 2

/* ---- call-level error codes ------------------------------------------------------------ */
#define SWGPU_OK 0
#define SWGPU_E_INVALID (-1)  /* bad argument */
#define SWGPU_E_CUDA (-2)     /* CUDA runtime error (text has the CUDA error string) */
#define SWGPU_E_NOMEM (-3)    /* allocation failed */
#define SWGPU_E_CAPACITY (-4) /* an output did not fit the caller's capacity; counts are valid */
#define SWGPU_E_STATE (-5)    /* call not valid in this state (e.g. zone test before set_zones) */

/* ---- per-record decode status (uint8) --------------------------------------------------- */
/* 0: decoded, rows written.                                                                  */
#define SWGPU_ST_OK 0
/* 1: the reference throws EventDecodeException (JsonParseException / JsonMappingException /
 *    IOException in ES/decoder/json/JsonDeviceRequestDecoder.java:55-61; IOException incl.
 *    InvalidProtocolBufferException in ES/decoder/protobuf/ProtobufDeviceEventDecoder.java:301-303)
 *    -> the record goes to the failed-decode topic (ES/InboundEventSource.java:211-213).       */
#define SWGPU_ST_DECODE_EXCEPTION 1
/* 2: the reference throws an unchecked exception that is NOT routed to the failed-decode
 *    topic (NullPointerException when parseDelimitedFrom returns null on an exhausted stream,
 *    ProtobufDeviceEventDecoder.java:79-81,124,160,194).                                        */
#define SWGPU_ST_RUNTIME_EXCEPTION 2
/* 3: valid envelope of a non-event request (registration, acknowledge, stream ...): outside the
 *    measured path; the binding hands the raw record to the stock decoder.                      */
#define SWGPU_ST_PASSTHROUGH 3

/* ---- decoded row: `kind` (low byte of meta) ------------------------------------------------ */
/* numbering = com.sitewhere.spi.device.event.DeviceEventType ordinals                          */
#define SWGPU_KIND_MEASUREMENT 0
#define SWGPU_KIND_LOCATION 1
#define SWGPU_KIND_ALERT 2

/* meta word:  bits 0-7 kind | 8-23 flags | 24-25 alert level | 26 alert source |
 *             27 eventType present | 28-30 eventType ordinal | 32-63 source record index      */
#define SWGPU_META_KIND(m) ((uint32_t)((m) & 0xff))
#define SWGPU_META_FLAGS(m) ((uint32_t)(((m) >> 8) & 0xffff))
#define SWGPU_META_LEVEL(m) ((uint32_t)(((m) >> 24) & 0x3))
#define SWGPU_META_SOURCE(m) ((uint32_t)(((m) >> 26) & 0x1))
#define SWGPU_META_HAS_ETYPE(m) ((uint32_t)(((m) >> 27) & 0x1))
#define SWGPU_META_ETYPE(m) ((uint32_t)(((m) >> 28) & 0x7))
#define SWGPU_META_REC(m) ((uint32_t)((m) >> 32))

#define SWGPU_F_UPDATE_STATE (1u << 0)  /* request.updateState                                 */
#define SWGPU_F_HAS_DATE (1u << 1)      /* event_date valid; else reference uses null / now()  */
#define SWGPU_F_HAS_ORIGINATOR (1u << 2)
#define SWGPU_F_HAS_METADATA (1u << 3)
#define SWGPU_F_HAS_ALT_ID (1u << 4)
#define SWGPU_F_NULL_TOKEN (1u << 5)    /* JSON deviceToken was not a string -> null token     */
#define SWGPU_F_HAS_V0 (1u << 6)        /* value / latitude  is non-null                       */
#define SWGPU_F_HAS_V1 (1u << 7)        /* longitude is non-null                               */
#define SWGPU_F_HAS_V2 (1u << 8)        /* elevation is non-null                               */
#define SWGPU_F_STR_ESCAPED (1u << 9)   /* some JSON string span holds backslash escapes       */
#define SWGPU_F_HAS_NAME (1u << 10)     /* measurement name / alert type is non-null           */
#define SWGPU_F_HAS_MESSAGE (1u << 11)  /* alert message is non-null                           */
#define SWGPU_F_NAME_COERCED (1u << 12) /* JSON message/alternateId given as number/bool token:
                                           the span is the raw token, Jackson re-renders it        */
#define SWGPU_F_NULL_REQUEST (1u << 13) /* JSON "request":null -> DecodedDeviceRequest with a null
                                           request (dropped later at ES/EventSourcesManager.java:155) */
#define SWGPU_F_NULL_LEVEL (1u << 14)   /* JSON alert "level":null                              */
#define SWGPU_F_NULL_SOURCE (1u << 15)  /* JSON alert "source":null                             */

/* A span is  (offset_in_record | length << 32); bytes are the caller's payload bytes.        */
#define SWGPU_SPAN_OFF(s) ((uint32_t)((s) & 0xffffffffu))
#define SWGPU_SPAN_LEN(s) ((uint32_t)((s) >> 32))

/*
 * Decoded rows, structure-of-arrays, one element per DECODED EVENT (not per record: a protobuf
 * DeviceMeasurements message yields one row per Measurement, ProtobufDeviceEventDecoder.java:126-157;
 * a failed record yields none).  Rows are in reference order: record order, then the order in
 * which the reference appends to its result list.  The first nine columns are mandatory (72 B
 * per row); the last four may be NULL.
 *
 * This is the columnar form of ES/DecodedDeviceRequest.java:17-29 + the request POJOs
 * (sitewhere-core-api/.../rest/model/device/event/request/Device{Event,Measurement,Location,
 * Alert}CreateRequest.java).
 */
typedef struct swgpu_decoded_cols {
    uint64_t *meta;        /* see SWGPU_META_*                                               */
    uint64_t *token_hash;  /* low 32: Kafka murmur2(token bytes); high 32: second hash        */
    uint64_t *name_hash;   /* same hash of measurement name / alert type (0 if none)          */
    int64_t *event_date;   /* ms since epoch (valid iff SWGPU_F_HAS_DATE)                      */
    double *v0;            /* measurement value | latitude                                    */
    double *v1;            /* longitude                                                       */
    double *v2;            /* elevation                                                       */
    uint64_t *token_span;  /* deviceToken / hardwareId bytes                                  */
    uint64_t *name_span;   /* measurement name / alert type bytes                             */
    uint64_t *originator_span;   /* optional */
    uint64_t *message_span;      /* optional: alert message */
    uint64_t *metadata_span;     /* optional: JSON: text of the metadata object; protobuf:
                                    offset of first Metadata field | entry count << 32 */
    uint64_t *alternate_id_span; /* optional (JSON only) */
} swgpu_decoded_cols;

/* Ids an enriched event carries (GDeviceEvent.id, .deviceAssignmentId, .deviceId: GUUID{msb, lsb},
 * sitewhere-grpc-model/src/main/proto/sitewhere-common.proto:30-33), one element per decoded row.  Either all six
 * pointers or none.                                                                                            */
typedef struct swgpu_event_ids {
    uint64_t *event_hi, *event_lo;           /* the persisted event's id            */
    uint64_t *assignment_hi, *assignment_lo; /* the reference's device-state key    */
    uint64_t *device_hi, *device_lo;
} swgpu_event_ids;

typedef struct swgpu_ctx swgpu_ctx;

/* ---- context ---------------------------------------------------------------------------- */
/* One context per GPU and per calling thread.  `stream` is a cudaStream_t to enqueue on, or
 * NULL to let the library create its own non-blocking stream.                                 */
int32_t swgpu_create(int32_t device_ordinal, void *stream, swgpu_ctx **out_ctx);
void swgpu_destroy(swgpu_ctx *ctx);
const char *swgpu_last_error(swgpu_ctx *ctx); /* ctx may be NULL: error of the last failed create */
int32_t swgpu_abi_version(void);
void *swgpu_stream(swgpu_ctx *ctx);
int32_t swgpu_sync(swgpu_ctx *ctx);
/* Named integer options of a context (no counterpart in the reference: they pick between code paths that
 * give identical results, so that tests can cover each path, or switch on work counters):
 *   "zone.force_ray"        0 = the zone walk picks its ray per point (default); 1..4 = always walk the
 *                           +x, -x, +y or -y copy of a ring, even for a zone whose table entry restricts the
 *                           choice (JTS 1.13 casts +x: RayCrossingCounter, called from the zone test at
 *                           RP/processors/geospatial/ZoneTestRuleProcessor.java:50)
 *   "zone.scan_sequential"  1 = scan the per-tile alert counts on one CTA
 *   "state.presence_now_ms" t > 0: every event folded into the device state from now on also touches its device's
 *                           presence row with lastInteractionDate = t and presenceMissingDate = null
 *                           (DS/processing/DeviceStateProcessingLogic.java:142-143; the reference takes new Date(),
 *                           here the caller supplies the batch's time so that results are reproducible); 0 = off
 *                           (default).  Set it before each batch.  Costs one more table access per event.
 *   "json.fast_path"        0 = every JSON record takes the general structural pass; 1 (default) = a record in
 *                           exactly the shape the reference's own serialiser writes is first tried against that
 *                           template (same result, far fewer instructions)
 *   "enriched.fast_path"    0 = every enriched record takes the general proto3 parser; 1 (default) = a payload in
 *                           the serialiser's own field order is read in one forward pass first (same result)
 *   "json.stats"            1 = JSON decodes count the records the template took (swgpu_json_stats)
 *   "zone.stats"            1 = zone tests also count their work (read it with swgpu_zone_stats)        */
int32_t swgpu_set_option(swgpu_ctx *ctx, const char *name, int64_t value);
/* Pinned host staging the Java side wraps as a direct ByteBuffer / MemorySegment.             */
void *swgpu_alloc_pinned(size_t bytes);
void swgpu_free_pinned(void *p);

/* ---- zone table --------------------------------------------------------------------------
 * Replaces the per-token polygon cache of RP/processors/geospatial/ZoneTestRuleProcessor.java:36,72-84
 * and GeoUtils.createPolygonForLocations (sitewhere-core/.../geospatial/GeoUtils.java:55-67).
 * Test t uses bounds[vtx_off[t] .. vtx_off[t+1]) as the OPEN ring in zone order (x = longitude,
 * y = latitude); the library closes it with the first vertex as GeoUtils does.  cond[t] is
 * ZoneContainment: 0 = Inside, 1 = Outside (RP/processors/geospatial/ZoneTest.java:25).
 * Host pointers.  Rebuilds the device table and its spatial index; call on configuration change. */
int32_t swgpu_set_zones(swgpu_ctx *ctx, uint32_t n_tests, const uint32_t *vtx_off,
                        const double *lon, const double *lat, const uint8_t *cond);
/* Packed table (one relocatable blob) for replication to other GPUs: export on the rank that
 * built it, ncclBroadcast the bytes, import on the others.  Device pointers.                   */
int32_t swgpu_zone_table_bytes(swgpu_ctx *ctx, uint64_t *out_bytes);
int32_t swgpu_zone_table_export_dev(swgpu_ctx *ctx, void *dst_dev, uint64_t capacity);
int32_t swgpu_zone_table_import_dev(swgpu_ctx *ctx, const void *src_dev, uint64_t bytes);

/* ---- decode -------------------------------------------------------------------------------
 * n records; record i is bytes[rec_off[i] .. rec_off[i+1]).  status[i] = SWGPU_ST_*;
 * first_row[i] = index of record i's first row (rows of record i are first_row[i] ..
 * first_row[i+1]); first_row has n+1 entries and first_row[n] = total rows.
 * Rows beyond max_rows are counted but not written and the call returns SWGPU_E_CAPACITY.
 *
 * swgpu_decode_pb   replaces ProtobufDeviceEventDecoder.decode(byte[], Map)
 *                   (ES/decoder/protobuf/ProtobufDeviceEventDecoder.java:75-304) applied to each record.
 * swgpu_decode_json replaces JsonDeviceRequestDecoder.decode(byte[], Map)
 *                   (ES/decoder/json/JsonDeviceRequestDecoder.java:48-62 with
 *                   JsonDeviceRequestMarshaler.java:57-160) applied to each record.
 * Both implement IDeviceEventDecoder<byte[]>.decode (ES/spi/IDeviceEventDecoder.java:39).       */
int32_t swgpu_decode_pb(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                        uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                        uint32_t *first_row, uint32_t *out_rows);
int32_t swgpu_decode_json(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                          uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                          uint32_t *first_row, uint32_t *out_rows);
/* Device-resident variants.  The payload is staged in 16-byte-aligned windows: the 16-byte block that holds
 * byte rec_off[0] and the one that holds byte rec_off[n]-1 must be readable in full (true for any buffer
 * from cudaMalloc; give a sub-buffer 16 bytes of slack after the last record).
 * out_rows_dev: one uint32 on the device.                                                       */
int32_t swgpu_decode_pb_dev(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                            uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                            uint32_t *first_row, uint32_t *out_rows_dev);
int32_t swgpu_decode_json_dev(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off,
                              uint32_t n, uint32_t max_rows, const swgpu_decoded_cols *out,
                              uint8_t *status, uint32_t *first_row, uint32_t *out_rows_dev);

/* Enriched events: what the rule-processing and device-state consumers read off the "inbound-enriched-events"
 * topic.  Record i is one serialised GEnrichedEventPayload (proto3;
 * sitewhere-grpc-event-management/src/main/proto/device-event-model.proto:321-324 with :139-215, :288-296).
 * Replaces EventModelMarshaler.parseEnrichedEventPayloadMessage (.../grpc/client/event/EventModelMarshaler.java:204-206)
 * + EventModelConverter.asApiEnrichedEventPayload (.../EventModelConverter.java:1926-1931) as called per record by
 * DeviceStateProcessingLogic.processRecord (DS/processing/DeviceStateProcessingLogic.java:93-108) and
 * KafkaRuleProcessorHost (RP/kafka/KafkaRuleProcessorHost.java:170-197).
 * A record yields one row — kind = the oneof member (measurement / location / alert), eventType ordinal =
 * GDeviceEvent.eventType, event_date valid iff eventDate != 0, v0/v1/v2 with SWGPU_F_HAS_V* = has{Latitude,
 * Longitude,Elevation}, name = measurement name / alert type, message, alternateId; token_hash = hash of the 16
 * bytes (msb, lsb little-endian) of deviceAssignmentId and token_span = 0 — or none: status
 * SWGPU_ST_DECODE_EXCEPTION when the reference's parse or conversion throws (malformed message, invalid UTF-8 in a
 * string, an enum number outside the schema, no event set: the consumers log and skip the record), or
 * SWGPU_ST_PASSTHROUGH for the other event types.  `ids` (may be NULL) receives the ids per row.
 * The rows feed swgpu_zone_test* unchanged; for the device state use swgpu_decode_merge* with
 * SWGPU_FMT_ENRICHED (key = assignment id, event id = the payload's own: id_hi / id_lo are ignored).          */
int32_t swgpu_decode_enriched(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                              uint32_t max_rows, const swgpu_decoded_cols *out, const swgpu_event_ids *ids,
                              uint8_t *status, uint32_t *first_row, uint32_t *out_rows);
int32_t swgpu_decode_enriched_dev(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                                  uint32_t max_rows, const swgpu_decoded_cols *out, const swgpu_event_ids *ids,
                                  uint8_t *status, uint32_t *first_row, uint32_t *out_rows_dev);

/* ---- zone test ------------------------------------------------------------------------------
 * Replaces ZoneTestRuleProcessor.onLocation (RP/processors/geospatial/ZoneTestRuleProcessor.java:47-63)
 * applied to each location: for each test t in list order, inside = JTS-1.13
 * Polygon.contains(Point(lon,lat)); an alert fires iff (cond[t] == Inside) == inside.
 * Output: pairs (location index << 32 | test index), ordered by location then test — the order
 * in which the reference calls addDeviceAlerts.  *n_pairs gets the total count even when it
 * exceeds max_pairs (then SWGPU_E_CAPACITY).  `meta` may be NULL; if given, rows whose kind is
 * not SWGPU_KIND_LOCATION or that lack latitude/longitude are skipped (the reference only routes
 * Location events to onLocation, RP/kafka/KafkaRuleProcessorHost.java:205-237).                  */
int32_t swgpu_zone_test(swgpu_ctx *ctx, const double *lon, const double *lat, const uint64_t *meta,
                        uint32_t n, uint64_t max_pairs, uint64_t *pairs, uint64_t *n_pairs);
/* Device variant: n_rows_dev (may be NULL -> use n) holds the live row count (<= n), as written
 * by a *_dev decode; n_pairs_dev is one uint64 on the device.                                    */
int32_t swgpu_zone_test_dev(swgpu_ctx *ctx, const double *lon, const double *lat,
                            const uint64_t *meta, uint32_t n, const uint32_t *n_rows_dev,
                            uint64_t max_pairs, uint64_t *pairs, uint64_t *n_pairs_dev);

/* Measurement aid (no counterpart in the reference): work done by the last swgpu_zone_test[_dev] that ran
 * with option "zone.stats" = 1.  out[0] = grid-cell zone records read, out[1] = edge records read,
 * out[2] = straddle tests (an edge whose ends lie on either side of the ray: the exact-sign determinant of
 * RayCrossingCounter.countSegment), out[3] = locations.  Waits for the context's stream.                  */
int32_t swgpu_zone_stats(swgpu_ctx *ctx, uint64_t out[4]);

/* Measurement aid: of the last swgpu_decode_json_dev that ran with option "json.stats" = 1 (and nothing
 * else on the context since), out[0] = records the canonical-layout recogniser decoded (the rest went through
 * the general parser; counted again by the ordered second pass when there was one), out[1] = non-zero when
 * the ordered second pass ran.  Waits for the context's stream.                                            */
int32_t swgpu_json_stats(swgpu_ctx *ctx, uint64_t out[2]);

/* ---- alternate-id set --------------------------------------------------
 * AlternateIdDeduplicator.isDuplicate (ES/deduplicator/AlternateIdDeduplicator.java:42-57) asks the event
 * store, for every decoded request with an alternateId, whether an event with that id exists.  The context
 * keeps the set of alternate ids it has been shown (bytes decide, hashes only pick the bucket); for each decoded
 * row, in row order:                                                                                       */
#define SWGPU_DEDUP_NEW 0  /* no alternate id, or an id the set did not hold: it is in the set now          */
#define SWGPU_DEDUP_SEEN 1 /* the id was shown before: by an earlier call or by an earlier row of this one  */
#define SWGPU_DEDUP_ASK 2  /* the span's bytes are not the id's characters (JSON escapes, a coerced token),
                            * the id is longer than SWGPU_DEDUP_MAX_ID_BYTES, or it found no room: the store
                            * decides, as in the reference                                                   */
#define SWGPU_DEDUP_MAX_ID_BYTES 72u /* (a UUID string is 36) */
/* SEEN is the reference's "duplicate" whenever every id the set holds is an id of a stored event; NEW is
 * authoritative only for a set that was seeded with the store's ids (probe them once: any batch whose rows
 * carry them) — otherwise the caller still asks the store for NEW rows and saves the round trip for SEEN ones.
 * meta / alternate_id_span: the decoded columns over the same bytes / rec_off.  Asynchronous on the context's
 * stream.  swgpu_dedup_reserve sizes the set beforehand (it grows by itself otherwise), swgpu_dedup_size reads
 * the number of distinct ids (waits for the stream), swgpu_dedup_clear empties it.                           */
int32_t swgpu_dedup_probe_dev(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, const uint64_t *meta,
                              const uint64_t *alternate_id_span, uint32_t n_rows, const uint32_t *n_rows_dev,
                              uint8_t *dup);
/* host form: the alternate ids back to back (id_off[n + 1]), one per row; waits for the result.  This is what a
 * deduplicator bean calls per request or per micro-batch (IDeviceEventDeduplicator.isDuplicate,
 * ES/spi/IDeviceEventDeduplicator.java:29, called at ES/InboundEventSource.java:226-227).                   */
int32_t swgpu_dedup_probe(swgpu_ctx *ctx, const uint8_t *ids, const uint64_t *id_off, uint32_t n, uint8_t *dup);
int32_t swgpu_dedup_reserve(swgpu_ctx *ctx, uint64_t max_ids);
int32_t swgpu_dedup_size(swgpu_ctx *ctx, uint64_t *n_ids);
int32_t swgpu_dedup_clear(swgpu_ctx *ctx);

/* ---- outbound payload -------------------------------------------------
 * The GInboundEventPayload every decoded request is forwarded as, built on the device from the decoded rows.
 * Replaces, per decoded request, EventSourcesManager.handleDecodedEvent's
 * EventModelMarshaler.buildInboundEventPayloadMessage(payload) (ES/EventSourcesManager.java:152-166;
 * sitewhere-grpc-event-management/.../EventModelMarshaler.java:93-115, EventModelConverter.java:1878-1887,
 * 1621-1653, 399-406, 623-632, 770-784, 196-211).  `rows` are the columns of swgpu_decode_{pb,json}_dev over
 * the same `bytes` / `rec_off` (strings are read through the spans; the four optional span columns may be NULL =
 * no such property in any row), `format` says which decoder produced them (SWGPU_FMT_PROTOBUF: a row without
 * event date gets `now_ms`, as ProtobufDeviceEventDecoder substitutes new Date(); SWGPU_FMT_JSON: it stays null
 * -> 0).  Output: the records back to back in row order in `out`, out_off[n_rows + 1] (row r is
 * out[out_off[r] .. out_off[r+1])), *out_bytes_dev = total length, and per row enc_status:                  */
#define SWGPU_ENC_OK 0       /* bytes written                                                               */
#define SWGPU_ENC_THROWS 1   /* the reference throws here (NullPointerException from a builder setter or an
                              * unboxing, ClassCastException from an eventType that is not the request's):
                              * nothing is sent for this request                                             */
#define SWGPU_ENC_HANDBACK 3 /* the bytes depend on Java-side state (metadata map order, escaped / coerced /
                              * non-ASCII strings): the caller builds this one with the stock marshaler      */
#define SWGPU_ENC_NOTHING 4  /* the request is null: handleDecodedEvent sends nothing                        */
/* A 256-row tile whose bytes do not fit `out_cap` is not written (out_off and *out_bytes_dev are still exact:
 * call again with room).  source_id: at most 256 bytes, host memory.  Asynchronous on the context's stream. */
int32_t swgpu_encode_inbound_dev(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off,
                                 const swgpu_decoded_cols *rows, uint32_t n_rows, const uint32_t *n_rows_dev,
                                 const uint8_t *source_id, uint32_t source_id_len, int64_t now_ms, uint8_t *out,
                                 uint64_t out_cap, uint64_t *out_off, uint8_t *enc_status, uint64_t *out_bytes_dev);

/* ---- fused host-buffer path ------------------------------------------------------------------------
 * Decode + zone test of one micro-batch held in HOST memory, in one call: what the Java binding's
 * micro-batcher issues per batch.  Replaces IDeviceEventDecoder.decode per payload
 * (ES/spi/IDeviceEventDecoder.java:39) followed by ZoneTestRuleProcessor.onLocation per decoded location
 * (RP/processors/geospatial/ZoneTestRuleProcessor.java:47-63).  The batch is processed in chunks of
 * `chunk_records` records (0 = about a dozen chunks per call, between 131,072 and 1,048,576 records each) through a three-slot device ring so that the copy-in of
 * the next chunk, the kernels of the current one and the copy-out of the previous one overlap (a batch of at
 * most 16,384 records with chunk_records = 0 instead takes one staging block each way: one upload, the
 * kernels, one download — a 64 KB poll completes in ~75 us on the calling thread); for
 * that the caller's buffers should come from swgpu_alloc_pinned (pageable memory works, slower).
 * Outputs are exactly those of swgpu_decode_{pb,json} followed by swgpu_zone_test on the decoded
 * lon/lat/meta columns: rows in reference order, `first_row` / `status` per record, alert pairs
 * (row << 32 | test) in the order the reference would call addDeviceAlerts.  In `out` EVERY column
 * pointer may be NULL (not copied back); `status` and `first_row` may be NULL too.                  */
#define SWGPU_FMT_PROTOBUF 0
#define SWGPU_FMT_JSON 1
#define SWGPU_FMT_ENRICHED 2 /* records are GEnrichedEventPayload messages (swgpu_decode_enriched) */
int32_t swgpu_decode_zone_test(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off,
                               uint32_t n, uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                               uint32_t *first_row, uint32_t *out_rows, uint64_t max_pairs, uint64_t *pairs,
                               uint64_t *n_pairs, uint32_t chunk_records);

/* The same two steps on DEVICE buffers, asynchronous on the context's stream, in one call:
 * swgpu_decode_{pb,json,enriched}_dev followed by swgpu_zone_test_dev(lon = out->v1, lat = out->v0, meta =
 * out->meta, n = max_rows, n_rows_dev = out_rows_dev), with the same outputs
 * (RP/kafka/KafkaRuleProcessorHost.java:144-153 hands each decoded location to
 * ZoneTestRuleProcessor.onLocation).  All nine mandatory columns of `out` are required.                   */
int32_t swgpu_decode_zone_test_dev(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off,
                                   uint32_t n, uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                                   uint32_t *first_row, uint32_t *out_rows_dev, uint64_t max_pairs, uint64_t *pairs,
                                   uint64_t *n_pairs_dev);

/* Start-up sizing (no counterpart in the reference): allocates every device and pinned buffer that
 * swgpu_decode_zone_test needs for batches of up to max_records records / max_bytes payload bytes, so that no
 * cudaMalloc / cudaFree — which stall every stream of the device, other contexts' included — happens once batches
 * are flowing.  (A batch whose rows or alerts outnumber its records still grows its buffers.)                    */
int32_t swgpu_reserve_batch(swgpu_ctx *ctx, uint32_t max_records, uint64_t max_bytes);
/* Asynchronous form of swgpu_decode_zone_test (no counterpart in the reference, whose decoders are called
 * synchronously per payload): the call is queued on a worker thread owned by the context and *ticket identifies
 * it; swgpu_wait(ticket) blocks until it is done and returns ITS status (outputs valid then; swgpu_last_error has
 * the text).  Tickets complete in submission order.  Until every ticket has been waited for, the buffers passed in
 * belong to the library and no other call may be made on the context.                                             */
int32_t swgpu_submit_decode_zone_test(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off,
                                      uint32_t n, uint32_t max_rows, const swgpu_decoded_cols *out, uint8_t *status,
                                      uint32_t *first_row, uint32_t *out_rows, uint64_t max_pairs, uint64_t *pairs,
                                      uint64_t *n_pairs, uint32_t chunk_records, uint64_t *ticket);
int32_t swgpu_wait(swgpu_ctx *ctx, uint64_t ticket);

/* Allocation-free form of swgpu_zone_test_dev.  It has no counterpart in the reference: it exists
 * for callers of this library that must not allocate device memory while a batch is in flight —
 * the rule-processing consumer thread when other contexts of the same process have work on the GPU
 * (cudaMalloc / cudaFree stall every stream of the device), or a deployment whose memory budget is
 * already spent on the state table, where a late scratch growth could fail mid-partition.
 * Same arguments, outputs and ordering as swgpu_zone_test_dev; the caller still owns every input
 * and output buffer.  The scan scratch is a fixed array inside the library image (512 KiB + 64 B
 * per GPU, held until the process exits), so:
 *   - n must not exceed SWGPU_NOALLOC_MAX_LOCATIONS, else SWGPU_E_CAPACITY and nothing is launched;
 *   - the first call on each GPU resolves the array's device address (one-time set-up): make that
 *     first call before starting any CUDA stream capture;
 *   - all calls on a GPU share the one array (it is cleared at the start of each call): two calls
 *     whose kernels can overlap on the GPU — different contexts or streams — race.  Serialise them.
 * The call itself performs no cudaMalloc/cudaFree; it needs swgpu_set_zones to have been called.   */
#define SWGPU_NOALLOC_MAX_LOCATIONS 16777216u
int32_t swgpu_zone_test_noalloc_dev(swgpu_ctx *ctx, const double *lon, const double *lat,
                                    const uint64_t *meta, uint32_t n, const uint32_t *n_rows_dev,
                                    uint64_t max_pairs, uint64_t *pairs, uint64_t *n_pairs_dev);

/* ---- device-state merge -------------------------------------------------------------------------
 * Replaces the fold DeviceStateProcessingLogic.processDeviceStateEvent + merge*
 * (DS/processing/DeviceStateProcessingLogic.java:116-201) with DeviceStatePersistence.deviceStateUpdateLogic
 * (DS/persistence/DeviceStatePersistence.java:60-85) over a batch, behind
 * IDeviceStateProcessingLogic.process(List<ConsumerRecord>) (DS/spi/processing/IDeviceStateProcessingLogic.java:30).
 *
 * State key = (device token bytes, kind, name bytes) where name is the measurement name, the
 * alert type, or empty for a location (lastLocationEventId / lastMeasurementEventIds[name] /
 * lastAlertEventIds[type]).  Per key the winner is the event with the largest sequence number
 * (arrival order, as the reference's sequential loop).  The table keeps, per key, the winner's sequence,
 * 128-bit event id, event date and value.  Keys are compared by BYTES (hashes only pick the bucket).
 * Rows without a device token (SWGPU_F_NULL_TOKEN) or with a null request (SWGPU_F_NULL_REQUEST: the
 * reference drops them at ES/EventSourcesManager.java:155) do not reach the state.
 *
 * swgpu_merge / swgpu_merge_dev fold rows that a decode call produced on the same `bytes`/`rec_off`:
 * row r has sequence seq_base + r; id_hi/id_lo[r] are the event ids the persistence layer assigned
 * (EM/persistence/DeviceEventManagementPersistence.java:82).  Advance seq_base by n_rows.
 *
 * Capacity.  The host-buffer calls (swgpu_merge, swgpu_decode_merge) grow the table as needed and return
 * when the batch is folded.  The *_dev calls only enqueue: they grow the table from a host-side upper bound
 * before launching, but rows they cannot place (table full after a fan-out of new keys, keys longer than 72
 * bytes when the long-key arena is full, token or name beyond 65,535 bytes) are dropped and counted; the next
 * swgpu_state_size reports that as SWGPU_E_CAPACITY (once).  After swgpu_state_reserve the dropped batches can
 * be replayed: folding a batch again with the same sequence numbers changes nothing.                    */
int32_t swgpu_state_reserve(swgpu_ctx *ctx, uint64_t max_keys, uint64_t max_long_key_bytes);
int32_t swgpu_merge(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n_records,
                    const swgpu_decoded_cols *rows, uint32_t n_rows, uint64_t seq_base,
                    const uint64_t *id_hi, const uint64_t *id_lo);
int32_t swgpu_merge_dev(swgpu_ctx *ctx, const uint8_t *bytes, const uint64_t *rec_off,
                        uint32_t n_records, const swgpu_decoded_cols *rows, uint32_t n_rows,
                        const uint32_t *n_rows_dev, uint64_t seq_base, const uint64_t *id_hi,
                        const uint64_t *id_lo);
/* Fused call of the device-state consumer: decode the batch and fold every decoded row into the state in ONE
 * kernel — the row's key bytes and payload go from the decoder's registers and shared memory into the table;
 * no decoded columns are written or read back.  Replaces, for one poll() batch,
 * DeviceStateProcessingLogic.process (DS/processing/DeviceStateProcessingLogic.java:65-67: unmarshal each
 * record, then processDeviceStateEvent :116-160).  The state afterwards is what swgpu_decode_{pb,json} followed
 * by swgpu_merge leaves, with these sequence numbers: row k of record i (k counts the rows of a
 * DeviceMeasurements fan-out) has  seq_base + (i << SWGPU_SEQ_RECORD_SHIFT) + min(k, 4095)  — arrival order,
 * independent of how many rows other records yield, so nothing has to be placed in order first.  Advance
 * seq_base by SWGPU_SEQ_STRIDE(n) per batch.
 * id_hi/id_lo[n]: the persisted event id of each RECORD; a record that decodes to several rows gives all of
 * them its id.  status[n] as in swgpu_decode_* (host variant: may be NULL); *out_rows = rows decoded.
 * swgpu_decode_merge takes HOST pointers and returns when the batch is folded; swgpu_decode_merge_dev takes
 * DEVICE pointers (payload readable as for swgpu_decode_*_dev), enqueues on the context's stream, and
 * out_rows_dev is one uint32 on the device.                                                               */
#define SWGPU_SEQ_RECORD_SHIFT 12
#define SWGPU_SEQ_STRIDE(n_records) ((uint64_t)(n_records) << SWGPU_SEQ_RECORD_SHIFT)
int32_t swgpu_decode_merge(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                           uint64_t seq_base, const uint64_t *id_hi, const uint64_t *id_lo, uint8_t *status,
                           uint32_t *out_rows);
int32_t swgpu_decode_merge_dev(swgpu_ctx *ctx, int32_t format, const uint8_t *bytes, const uint64_t *rec_off, uint32_t n,
                               uint64_t seq_base, const uint64_t *id_hi, const uint64_t *id_lo, uint8_t *status,
                               uint32_t *out_rows_dev);
/* Number of live keys and the bytes of their public form (see swgpu_state_rows).  Returns SWGPU_E_CAPACITY
 * (once; the counts are still valid) when a *_dev merge since the last check had to drop rows.          */
int32_t swgpu_state_size(swgpu_ctx *ctx, uint64_t *n_keys, uint64_t *key_bytes);
/* Presence scan.  Replaces one round of DevicePresenceManager.PresenceChecker (DS/presence/DevicePresenceManager.java:
 * 139-167; query DS/persistence/mongodb/MongoDeviceStateManagement.java:170-193): every device whose presence row
 * has lastInteractionDate <= now_ms - interval_ms and presenceMissingDate == null (SendOnceNotificationStrategy,
 * DS/presence/PresenceNotificationStrategies.java:24-31) is reported — the caller sends the "presence" state change —
 * and gets presenceMissingDate := now_ms; an event folded later (with "state.presence_now_ms" set) clears it again.
 * Host pointers.  Output ascending by (lastInteractionDate, key): last_interaction[k], and the device's key
 * key_bytes[key_off[k] .. key_off[k+1]) = token length (2 bytes LE) | 0x7f | token bytes (for enriched events the
 * token is the 16 bytes of the assignment id).  *n_found = all devices found missing; when they do not fit
 * (max_devices, max_key_bytes) the ones that fit are reported and marked, the call returns SWGPU_E_CAPACITY and the
 * next scan finds the rest.  A presence row also appears in swgpu_state_snapshot: kind byte 0x7f, seq =
 * lastInteractionDate, event_date = presenceMissingDate (0 = null).                                                 */
int32_t swgpu_presence_scan(swgpu_ctx *ctx, int64_t now_ms, int64_t interval_ms, uint64_t max_devices,
                            uint64_t max_key_bytes, int64_t *last_interaction, uint64_t *key_off, uint8_t *key_bytes,
                            uint64_t *n_reported, uint64_t *n_found);
/* Snapshot (host pointers): one row per live key, in unspecified order.  key_off has n_keys+1
 * entries into key_bytes; a key is  token length (2 bytes LE) | kind byte | token bytes | name bytes.               */
typedef struct swgpu_state_rows {
    uint64_t *seq;
    uint64_t *id_hi;
    uint64_t *id_lo;
    int64_t *event_date;
    double *value;
    uint64_t *key_off;
    uint8_t *key_bytes;
} swgpu_state_rows;
int32_t swgpu_state_snapshot(swgpu_ctx *ctx, uint64_t max_keys, uint64_t max_key_bytes,
                             const swgpu_state_rows *out, uint64_t *n_keys);
int32_t swgpu_state_restore(swgpu_ctx *ctx, const swgpu_state_rows *in, uint64_t n_keys);
int32_t swgpu_state_clear(swgpu_ctx *ctx);

/* ---- partitioning ----------------------------------------------------------------------------
 * Kafka DefaultPartitioner for keyed records, which the reference relies on by sending
 * (key = deviceToken) without a partition (MS/kafka/MicroserviceKafkaProducer.java:89-92,
 * ES/EventSourcesManager.java:163):  toPositive(murmur2(utf8(key))) % n_partitions.
 * Host helper (no GPU): the shard a token belongs to.                                          */
uint32_t swgpu_murmur2(const uint8_t *key, uint32_t len);
/* Host helper (no GPU): decimal text -> double with the routine the JSON kernel uses, bit-exact with
 * java.lang.Double.parseDouble on [+-]?digits[.digits][(e|E)[+-]digits] (what Jackson hands it for a
 * number token, JsonDeviceRequestMarshaler.java:97-160 via treeToValue).  For bindings that coerce a
 * numeric string themselves.                                                                      */
int32_t swgpu_parse_double(const char *text, uint32_t len, double *out);
uint32_t swgpu_partition(const uint8_t *key, uint32_t len, uint32_t n_partitions);

#ifdef __cplusplus
}
#endif
#endif /* SWGPU_H */
