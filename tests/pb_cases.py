"""Adversarial / edge-case protobuf payloads with the status the reference decoder gives
(ProtobufDeviceEventDecoder.java:75-304 on protobuf-java 3.3.0 lite).  Used by the oracle tests
(known answers) and by the GPU parity tests (GPU vs oracle)."""
import struct

from sitewhere_b200 import synth as S

OK, DECODE, RUNTIME, PASS = 0, 1, 2, 3
TOK = "12345-SYN-0000042"


def _loc_body(**kw):
    rec = S.pb_location_record(TOK, 34.5, -84.25, **kw)
    return rec


def cases():
    c = []

    def add(name, rec, status, rows=None):
        c.append((name, bytes(rec), status, rows))

    hdr3, hdr4, hdr5 = S.pb_header(3), S.pb_header(4), S.pb_header(5)
    # ---- well formed ---------------------------------------------------------------------------------
    add("loc_min", S.pb_location_record(TOK, 1.5, 2.5), OK, 1)
    add("loc_full", S.pb_location_record(TOK, 34.10469794977326, -84.23966646194458, elev=0.0,
                                         date=1527854400123, metadata={"fromMQTT": "true"}, update_state=True,
                                         originator="orig-1"), OK, 1)
    add("loc_update_false", S.pb_location_record(TOK, 1.0, 2.0, update_state=False), OK, 1)
    add("loc_negzero_nan", S.pb_location_record(TOK, -0.0, float("nan"), elev=float("inf")), OK, 1)
    add("alert", S.pb_alert_record(TOK, "engine.overheat", "eeeeyyyyyy!", date=1, metadata={"a": "b", "c": "d"}), OK, 1)
    add("mx_one", S.pb_measurements_record(TOK, [("fuel.level", 123.4)], date=5), OK, 1)
    add("mx_three", S.pb_measurements_record(TOK, [("a", 1.0), ("bb", -2.5), ("ccc", 1e300)], date=5,
                                             metadata={"k": "v"}, update_state=True, originator="o"), OK, 3)
    add("mx_zero", S.pb_measurements_record(TOK, [], date=5), OK, 0)
    add("empty_token", S.pb_location_record("", 1.0, 2.0), OK, 1)
    add("utf8_token", S.pb_location_record("dév-☃", 1.0, 2.0), OK, 1)
    add("long_token", S.pb_location_record("T" * 300, 1.0, 2.0), OK, 1)
    add("trailing_garbage", S.pb_location_record(TOK, 1.0, 2.0) + b"\xff\xff\xff", OK, 1)
    # ---- last-wins / unknown fields ---------------------------------------------------------------------
    m = (S.pb_field_len(1, b"first") + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0) + S.pb_field_len(1, b"second")
         + S.pb_field_f64(2, 9.0))
    add("repeated_scalars_last_wins", hdr3 + S.pb_delimited(m), OK, 1)
    m = (S.pb_field_varint(99, 7) + S.pb_field_len(1, TOK.encode()) + S.pb_field_len(77, b"xyz") + S.pb_field_f64(2, 1.0)
         + S.varint((55 << 3) | 5) + b"\1\2\3\4" + S.pb_field_f64(3, 2.0) + S.pb_field_fixed64(66, 1234))
    add("unknown_fields_skipped", hdr3 + S.pb_delimited(m), OK, 1)
    # field 2 with the wrong wire type is an unknown field -> latitude missing
    m = S.pb_field_len(1, TOK.encode()) + S.pb_field_varint(2, 5) + S.pb_field_f64(3, 2.0)
    add("wrong_wiretype_is_unknown", hdr3 + S.pb_delimited(m), DECODE)
    # unknown group, properly closed / mismatched / unterminated
    grp = S.varint((9 << 3) | 3) + S.pb_field_varint(1, 1) + S.varint((9 << 3) | 4)
    m = S.pb_field_len(1, TOK.encode()) + grp + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0)
    add("unknown_group_ok", hdr3 + S.pb_delimited(m), OK, 1)
    nested = (S.varint((9 << 3) | 3) + S.varint((8 << 3) | 3) + S.pb_field_varint(1, 1) + S.varint((8 << 3) | 4)
              + S.varint((9 << 3) | 4))
    m = S.pb_field_len(1, TOK.encode()) + nested + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0)
    add("unknown_group_nested_ok", hdr3 + S.pb_delimited(m), OK, 1)
    bad = S.varint((9 << 3) | 3) + S.pb_field_varint(1, 1) + S.varint((8 << 3) | 4)
    m = S.pb_field_len(1, TOK.encode()) + bad + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0)
    add("unknown_group_mismatched_end", hdr3 + S.pb_delimited(m), DECODE)
    m = S.pb_field_len(1, TOK.encode()) + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0) + S.varint((9 << 3) | 3)
    add("unknown_group_unterminated", hdr3 + S.pb_delimited(m), DECODE)
    m = S.pb_field_len(1, TOK.encode()) + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0) + S.varint((9 << 3) | 4)
    add("stray_end_group", hdr3 + S.pb_delimited(m), DECODE)
    for wt in (6, 7):
        m = S.pb_field_len(1, TOK.encode()) + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0) + S.varint((9 << 3) | wt)
        add(f"invalid_wiretype_{wt}", hdr3 + S.pb_delimited(m), DECODE)
    m = S.pb_field_len(1, TOK.encode()) + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0) + b"\x00"
    add("tag_zero_invalid", hdr3 + S.pb_delimited(m), DECODE)
    # ---- missing required ---------------------------------------------------------------------------------
    add("missing_lat", hdr3 + S.pb_delimited(S.pb_field_len(1, TOK.encode()) + S.pb_field_f64(3, 2.0)), DECODE)
    add("missing_hw", hdr3 + S.pb_delimited(S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0)), DECODE)
    add("alert_missing_msg", hdr4 + S.pb_delimited(S.pb_field_len(1, TOK.encode()) + S.pb_field_len(2, b"t")), DECODE)
    add("mx_missing_value", hdr5 + S.pb_delimited(S.pb_field_len(1, TOK.encode())
                                                  + S.pb_field_len(2, S.pb_field_len(1, b"name"))), DECODE)
    add("metadata_missing_value", hdr3 + S.pb_delimited(
        S.pb_field_len(1, TOK.encode()) + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0)
        + S.pb_field_len(6, S.pb_field_len(1, b"k"))), DECODE)
    add("header_missing_command", S.pb_delimited(S.pb_field_len(2, b"orig")) + S.pb_delimited(b""), DECODE)
    add("header_unknown_enum", S.pb_delimited(S.pb_field_varint(1, 42)) + S.pb_delimited(b""), DECODE)
    add("header_enum_zero", S.pb_delimited(S.pb_field_varint(1, 0)) + S.pb_delimited(b""), DECODE)
    # ---- truncation ---------------------------------------------------------------------------------------
    full = S.pb_location_record(TOK, 1.0, 2.0, elev=3.0, date=4)
    add("truncated_mid_double", full[:-3], DECODE)
    add("truncated_at_field_boundary_missing_req", full[:30], DECODE)
    # declared length longer than the stream: parses to EOF, still complete -> OK
    m = S.pb_field_len(1, TOK.encode()) + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0)
    add("length_longer_than_stream", hdr3 + S.varint(len(m) + 50) + m, OK, 1)
    add("empty_payload", b"", RUNTIME)
    add("header_only", hdr3, RUNTIME)
    add("header_len_varint_truncated", b"\x80", DECODE)
    add("header_len_varint_malformed", b"\xff" * 11, DECODE)
    add("string_len_beyond_end", hdr3 + S.pb_delimited(b"\x0a\x7f" + b"abc"), DECODE)
    add("string_negative_len", hdr3 + S.pb_delimited(b"\x0a" + S.varint(0xFFFFFFFF) + b"abc"), DECODE)
    # 10-byte varint tag whose low 32 bits are a valid tag (upper bits dropped)
    t = S.varint((1 << 40) | 10)
    m = t + S.varint(len(TOK)) + TOK.encode() + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0)
    add("long_varint_tag_low32", hdr3 + S.pb_delimited(m), OK, 1)
    # bool as multi-byte varint; non-canonical zero
    m = (S.pb_field_len(1, TOK.encode()) + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0) + S.varint(56)
         + b"\x80\x80\x01")
    add("bool_multibyte_true", hdr3 + S.pb_delimited(m), OK, 1)
    m = (S.pb_field_len(1, TOK.encode()) + S.pb_field_f64(2, 1.0) + S.pb_field_f64(3, 2.0) + S.varint(56)
         + b"\x80\x80\x00")
    add("bool_multibyte_false", hdr3 + S.pb_delimited(m), OK, 1)
    # nested measurement declared longer than the enclosing message: ends at EOF cleanly
    mx = S.pb_field_len(1, b"nm") + S.pb_field_f64(2, 7.0)
    m = S.pb_field_len(1, TOK.encode()) + S.varint(18) + S.varint(len(mx) + 9) + mx
    add("nested_len_beyond_stream_ok", hdr5 + S.pb_delimited(m), OK, 1)
    # nested message with stray end-group
    mx = S.pb_field_len(1, b"nm") + S.pb_field_f64(2, 7.0) + S.varint((3 << 3) | 4)
    m = S.pb_field_len(1, TOK.encode()) + S.pb_field_len(2, mx)
    add("nested_stray_end_group", hdr5 + S.pb_delimited(m), DECODE)
    # ---- non-event commands ---------------------------------------------------------------------------------
    for cmd in (1, 2, 6, 7, 8):
        add(f"passthrough_cmd_{cmd}", S.pb_header(cmd, "o") + S.pb_delimited(S.pb_field_len(1, TOK.encode())), PASS)
    # zero-length header message: command missing
    add("header_empty_message", b"\x00" + S.pb_delimited(b""), DECODE)
    # header length prefix negative (5-byte varint with the sign bit): empty header -> command missing
    add("header_negative_len", S.varint(0xFFFFFFFF) + b"\x08\x03", DECODE)
    return c
