/*
 * sw_oracle.h — CPU ORACLE for the SiteWhere inbound event hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under sitewhere_b200/ may include, link or call this.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it,
 * as the checker (and as the timed CPU baseline), never as the product path.
 *
 * It restates, in plain C, what the reference (SiteWhere 2.0.rc2, Java) computes on this path.
 * The reference cannot be compiled or run here (no JDK, no jars), and the arithmetic lives in
 * third-party libraries that are not in the reference tree:
 *     protobuf-java 3.3.0 lite   (build.gradle:25)
 *     Jackson 2.9.x              (platform-bom Cairo-SR1, build.gradle:64)
 *     JTS 1.13                   (sitewhere-core/build.gradle:20)
 *     kafka-clients 0.11.0.0     (build.gradle:21)
 *     JDK 8 Double.parseDouble
 * whose published algorithms are restated from recollection.  The reference has no test that
 * pins results for this path, so:   PARITY UNPINNED BY THE REFERENCE.
 * What pins this oracle instead (tests/test_oracle_*.py): google.protobuf's runtime on a
 * hand-built descriptor of sitewhere.proto, CPython json + float() (correctly rounded, as
 * Double.parseDouble), exact rational arithmetic (fractions.Fraction) for the orientation sign
 * and ray crossing, Kafka's published murmur2 test vectors, and the reference's own fixtures
 * (EventsHelper / MqttTests payload shapes, the 'workarea' zone of the construction dataset).
 */
#ifndef SW_ORACLE_H
#define SW_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Same column set and bit layout as include/swgpu.h (kept textually separate on purpose: the
 * oracle does not include product headers). */
typedef struct orc_cols {
    uint64_t *meta, *token_hash, *name_hash;
    int64_t *event_date;
    double *v0, *v1, *v2;
    uint64_t *token_span, *name_span, *originator_span, *message_span, *metadata_span,
        *alternate_id_span;
} orc_cols;

/* ids carried by an enriched event (one element per decoded row): GDeviceEvent.id, .deviceAssignmentId, .deviceId */
typedef struct orc_event_ids {
    uint64_t *event_hi, *event_lo, *assignment_hi, *assignment_lo, *device_hi, *device_lo;
} orc_event_ids;

uint32_t orc_murmur2(const uint8_t *data, uint32_t len);
uint32_t orc_partition(const uint8_t *key, uint32_t len, uint32_t n_partitions);
uint64_t orc_hash64(const uint8_t *data, uint32_t len);

/* returns number of rows written (rows beyond max_rows are counted, not written) */
uint64_t orc_decode_pb(const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint64_t max_rows,
                       const orc_cols *out, uint8_t *status, uint32_t *first_row);
uint64_t orc_decode_json(const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint64_t max_rows,
                         const orc_cols *out, uint8_t *status, uint32_t *first_row);
/* GEnrichedEventPayload records (sw_oracle_enriched.c); `ids` may be NULL */
uint64_t orc_decode_enriched(const uint8_t *bytes, const uint64_t *rec_off, uint32_t n, uint64_t max_rows,
                             const orc_cols *out, const orc_event_ids *ids, uint8_t *status, uint32_t *first_row);
/* one JSON number token -> double bits as Jackson + Double.parseDouble give; returns 0 on ok */
int orc_json_number_to_double(const char *tok, uint32_t len, double *out);
/* ISO-8601 / epoch handling of eventDate strings as Jackson StdDateFormat; returns 0 on ok */
int orc_parse_date(const char *s, uint32_t len, int64_t *out_ms);

int orc_sign_det2x2(double x1, double y1, double x2, double y2);
/* JTS 1.13 Polygon.contains(Point) for one closed ring (npts includes the repeated first point) */
int orc_polygon_contains(const double *xs, const double *ys, uint32_t npts, double px, double py);
int orc_polygon_is_rectangle(const double *xs, const double *ys, uint32_t npts);
/* the index-free loop of ZoneTestRuleProcessor.onLocation; returns total pair count */
uint64_t orc_zone_test(uint32_t n_tests, const uint32_t *vtx_off, const double *lon, const double *lat,
                       const uint8_t *cond, const double *plon, const double *plat,
                       const uint64_t *meta, uint32_t n, uint64_t max_pairs, uint64_t *pairs);

/* device-state fold */
typedef struct orc_state orc_state;
orc_state *orc_state_new(void);
void orc_state_free(orc_state *);
void orc_state_merge(orc_state *, const uint8_t *bytes, const uint64_t *rec_off, const orc_cols *rows,
                     uint32_t n_rows, uint64_t seq_base, const uint64_t *id_hi, const uint64_t *id_lo);
void orc_state_merge_seq(orc_state *, const uint8_t *bytes, const uint64_t *rec_off, const orc_cols *rows,
                         uint32_t n_rows, uint64_t seq_base, const uint64_t *seq, const uint64_t *id_hi,
                         const uint64_t *id_lo);
/* fold of enriched rows: the key's token is the 16 bytes of the assignment id (msb, lsb little-endian), the
 * event id is the payload's own; seq as in orc_state_merge_seq */
void orc_state_merge_enriched(orc_state *, const uint8_t *bytes, const uint64_t *rec_off, const orc_cols *rows,
                              const orc_event_ids *ids, uint32_t n_rows, uint64_t seq_base, const uint64_t *seq);
/* presence: while now_ms_plus1 != 0 every folded event also touches its device's row (see sw_oracle.c) */
void orc_state_set_presence_now(orc_state *, uint64_t now_ms_plus1);
uint64_t orc_presence_scan(orc_state *, uint64_t now_ms, uint64_t interval_ms, uint64_t max_out, uint64_t *last_interaction,
                           uint64_t *key_off, uint8_t *key_bytes);
uint64_t orc_state_size(orc_state *, uint64_t *key_bytes);
/* rows sorted by key bytes; key_off has n+1 entries */
void orc_state_snapshot(orc_state *, uint64_t *seq, uint64_t *id_hi, uint64_t *id_lo,
                        int64_t *event_date, double *value, uint64_t *key_off, uint8_t *key_bytes);

/* the GInboundEventPayload a decoded row is forwarded as (sw_oracle_encode.c) */
uint64_t orc_encode_inbound(int format, const uint8_t *bytes, const uint64_t *rec_off, const orc_cols *c, uint32_t n_rows,
                            const uint8_t *source_id, uint32_t source_len, int64_t now_ms, uint8_t *out, uint64_t out_cap,
                            uint64_t *out_off, uint8_t *enc_status);

#ifdef __cplusplus
}
#endif
#endif
