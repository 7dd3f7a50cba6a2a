"""JSON payloads with the outcome JsonDeviceRequestDecoder gives (JsonDeviceRequestDecoder.java:48-62,
JsonDeviceRequestMarshaler.java:57-160 on Jackson 2.9 defaults).  Each Jackson behaviour of SURVEY.md §8
a-J4 has dedicated vectors.  Used by the oracle tests (known answers) and by the GPU parity tests."""
import datetime

from sitewhere_b200 import synth as S

OK, DECODE, RUNTIME, PASS = 0, 1, 2, 3
TOK = "20544-OPENHAB-7313064"
D0 = 1527854400000


def utc_ms(y, mo, d, h=0, mi=0, sec=0, ms=0, days=0):
    t = datetime.datetime(y, mo, d, tzinfo=datetime.timezone.utc) + datetime.timedelta(days=days, hours=h, minutes=mi,
                                                                                      seconds=sec, milliseconds=ms)
    return int((t - datetime.datetime(1970, 1, 1, tzinfo=datetime.timezone.utc)) / datetime.timedelta(milliseconds=1))


def env(request, type_="DeviceMeasurement", token='"%s"' % TOK, extra="", originator="null"):
    return ('{"deviceToken":%s,"originator":%s,"type":"%s","request":%s%s}' % (token, originator, type_, request, extra)).encode()


def mx(value="123.4", name='"fuel.level"', extra=""):
    return '{"eventType":"Measurement","eventDate":"2018-06-01T12:00:00.000Z","updateState":false,"name":%s,"value":%s%s}' % (
        name, value, extra)


def cases():
    """(name, payload bytes, status, checks) — checks: dict of expected fields for OK rows"""
    c = []

    def add(case_name, payload, status, **checks):
        c.append((case_name, payload if isinstance(payload, bytes) else payload.encode(), status, checks))

    # ---- the reference's own fixture shapes --------------------------------------------------------------
    add("fixture_measurement", S.json_measurement(TOK, "fuel.level", "123.4", D0), OK, v0=123.4, name=b"fuel.level",
        date=D0, token=TOK.encode(), kind=0)
    add("fixture_location", S.json_location("25683-LAIPAC-S911-7539094", "34.10469794977326", "-84.23966646194458", "0.0",
                                            D0 + 5, update_state=True, metadata={"fromMQTT": "true"}), OK,
        v0=34.10469794977326, v1=-84.23966646194458, v2=0.0, date=D0 + 5, kind=1, update=True, has_md=True)
    add("fixture_alert", S.json_alert("4545-UNO-9430527", "engine.overheat", "eeeeyyyyyy!", D0 - 600000), OK,
        name=b"engine.overheat", message=b"eeeeyyyyyy!", kind=2, level=0, source=0)
    # ---- envelope rules (Marshaler :57-84) -----------------------------------------------------------------
    add("type_missing", '{"deviceToken":"t","request":%s}' % mx(), DECODE)
    add("type_invalid", env(mx(), type_="Bogus"), DECODE)
    add("type_wrong_case", env(mx(), type_="devicemeasurement"), DECODE)
    add("type_number", '{"deviceToken":"t","type":3,"request":%s}' % mx(), DECODE)
    add("type_null", '{"deviceToken":"t","type":null,"request":%s}' % mx(), DECODE)
    add("request_missing", '{"deviceToken":"t","type":"DeviceMeasurement"}', DECODE)
    add("token_missing", '{"type":"DeviceMeasurement","request":%s}' % mx(), DECODE)
    add("token_null_is_null_token", env(mx(), token="null"), OK, null_token=True)
    add("token_number_is_null_token", env(mx(), token="17"), OK, null_token=True)
    add("originator_string", env(mx(), originator='"cmd-1"'), OK, originator=b"cmd-1")
    add("originator_number_ignored", env(mx(), originator="5"), OK, originator=None)
    add("unknown_top_level_keys_ignored", env(mx(), extra=',"foo":{"a":[1,2,{"b":null}]},"bar":1e5'), OK, v0=123.4)
    add("key_order_free", '{"request":%s,"type":"DeviceMeasurement","deviceToken":"%s"}' % (mx(), TOK), OK, v0=123.4)
    add("duplicate_type_last_wins", '{"type":"Bogus","deviceToken":"t","type":"DeviceMeasurement","request":%s}' % mx(), OK)
    add("duplicate_request_last_wins", '{"type":"DeviceMeasurement","deviceToken":"t","request":{"bogus":1},"request":%s}' % mx("7"),
        OK, v0=7.0)
    add("duplicate_request_last_bad", '{"type":"DeviceMeasurement","deviceToken":"t","request":%s,"request":{"bogus":1}}' % mx(), DECODE)
    add("request_null", env("null"), OK, null_request=True)
    add("request_string", env('"x"'), DECODE)
    add("request_array", env("[]"), DECODE)
    add("request_number", env("5"), DECODE)
    add("request_empty_object", env("{}"), OK, v0=0.0, name=None, date=None)
    for t in ("RegisterDevice", "DeviceStream", "DeviceStreamData", "Acknowledge"):
        add("passthrough_" + t, env("{}", type_=t), PASS)
    add("passthrough_type_but_request_missing", '{"deviceToken":"t","type":"RegisterDevice"}', DECODE)
    # ---- root / grammar (JsonParseException) ------------------------------------------------------------------
    add("empty", b"", DECODE)
    add("whitespace_only", b"  \n\t ", DECODE)
    add("root_array", b"[1,2]", DECODE)
    add("root_string", b'"x"', DECODE)
    add("root_number", b"12", DECODE)
    add("truncated", env(mx())[:-1], DECODE)
    add("trailing_garbage_ignored", env(mx()) + b" xyz {", OK, v0=123.4)
    add("leading_ws_and_bom", b"\xef\xbb\xbf \r\n" + env(mx()), OK, v0=123.4)
    add("utf16_passthrough", env(mx()).decode().encode("utf-16"), PASS)
    add("utf16le_nobom_passthrough", env(mx()).decode().encode("utf-16-le"), PASS)
    add("trailing_comma_object", env('{"name":"a",}'), DECODE)
    add("trailing_comma_array", env(mx(), extra=',"x":[1,]'), DECODE)
    add("missing_colon", env('{"name" "a"}'), DECODE)
    add("missing_comma", env('{"name":"a" "value":1}'), DECODE)
    add("single_quotes", env("{'name':'a'}"), DECODE)
    add("unquoted_key", env('{name:"a"}'), DECODE)
    add("comment", env(mx() + "/*c*/"), DECODE)
    add("leading_zero", env(mx("0123.4")), DECODE)
    add("leading_plus", env(mx("+1")), DECODE)
    add("bare_dot", env(mx(".5")), DECODE)
    add("trailing_dot", env(mx("5.")), DECODE)
    add("exp_no_digits", env(mx("1e")), DECODE)
    add("nan_literal", env(mx("NaN")), DECODE)
    add("infinity_literal", env(mx("Infinity")), DECODE)
    add("hex_number", env(mx("0x10")), DECODE)
    add("minus_only", env(mx("-")), DECODE)
    add("literal_bad", env(mx("tru")), DECODE)
    add("literal_suffix", env(mx("nullx")), DECODE)
    add("control_char_in_string", env(mx(name='"a\x01b"')), DECODE)
    add("tab_in_string", env(mx(name='"a\tb"')), DECODE)
    add("bad_escape", env(mx(name='"a\\qb"')), DECODE)
    add("bad_unicode_escape", env(mx(name='"a\\u12G4"')), DECODE)
    add("unterminated_string", b'{"deviceToken":"abc', DECODE)
    add("invalid_utf8_lead", env(mx(name='"a%sb"')).replace(b"%s", b"\xff"), DECODE)
    add("invalid_utf8_cont", env(mx(name='"a%sb"')).replace(b"%s", b"\xc3\x28"), DECODE)
    add("valid_utf8_name", env(mx(name='"température"')), OK, name="température".encode())
    add("utf8_4byte", env(mx(name='"m😀"')), OK, name="m😀".encode())
    add("ctrl_between_tokens", env(mx()).replace(b',"type"', b',\x0b"type"'), DECODE)
    add("deep_nesting_ok", env(mx(), extra=',"x":' + "[" * 200 + "]" * 200), OK)
    add("deep_nesting_passthrough", env(mx(), extra=',"x":' + "[" * 300 + "]" * 300), PASS)
    add("deep_nesting_then_error_is_passthrough", env(mx(), extra=',"x":' + "[" * 300), PASS)
    add("error_before_deep_nesting", b'{"a":tru,"x":' + b"[" * 300 + b"]" * 300 + b"}", DECODE)
    # ---- escapes ------------------------------------------------------------------------------------------------
    add("escaped_key_depth1", env(mx()).replace(b'"type"', b'"t\\u0079pe"'), PASS)
    add("escaped_key_depth2", env(mx().replace('"name"', '"n\\u0061me"')), PASS)
    add("escaped_key_depth3_ok", env(mx(extra=',"metadata":{"a\\nb":"c"}')), OK, has_md=True)
    add("escaped_token", env(mx(), token='"a\\u0031"'), PASS)
    add("escaped_name", env(mx(name='"fuel\\u002elevel"')), PASS)
    add("escaped_message_flag", S.json_alert(TOK, "t", 'say \\"hi\\"', D0), OK, escaped=True, message=b'say \\"hi\\"')
    add("escaped_type_value", env(mx()).replace(b'"DeviceMeasurement"', b'"Device\\u004deasurement"'), PASS)
    add("surrogate_pair_escape_in_message", S.json_alert(TOK, "t", "\\ud83d\\ude00", D0), OK, escaped=True)
    # ---- bean: unknown properties, per-class property sets ----------------------------------------------------------
    add("unknown_property", env(mx(extra=',"bogus":1')), DECODE)
    add("location_prop_in_measurement", env(mx(extra=',"latitude":1.0')), DECODE)
    add("measurement_prop_in_location", env('{"latitude":1,"longitude":2,"value":3}', type_="DeviceLocation"), DECODE)
    add("alert_type_prop_in_measurement", env(mx(extra=',"type":"x"')), DECODE)
    add("case_sensitive_property", env('{"Name":"a","value":1}'), DECODE)
    add("duplicate_property_last_wins", env('{"name":"a","value":true,"value":2.5,"name":"b"}'), OK, v0=2.5, name=b"b")
    add("duplicate_property_last_bad", env('{"name":"a","value":2.5,"value":true}'), DECODE)
    add("all_base_props", env('{"alternateId":"alt-1","eventType":"Measurement","eventDate":1,"updateState":true,'
                              '"metadata":{"k":"v","n":null,"i":5},"name":"a","value":1}'), OK,
        alt=b"alt-1", date=1, update=True, has_md=True)
    add("metadata_null", env(mx(extra=',"metadata":null')), OK, has_md=False)
    add("metadata_string", env(mx(extra=',"metadata":"x"')), DECODE)
    add("metadata_array", env(mx(extra=',"metadata":[]')), DECODE)
    add("metadata_nested_object_value", env(mx(extra=',"metadata":{"k":{"a":1}}')), DECODE)
    add("metadata_nested_array_value", env(mx(extra=',"metadata":{"k":[1]}')), DECODE)
    add("metadata_empty", env(mx(extra=',"metadata":{}')), OK, has_md=True)
    # ---- doubles ------------------------------------------------------------------------------------------------------
    add("int_value", env(mx("12")), OK, v0=12.0)
    add("neg_zero_int_is_pos_zero", env(mx("-0")), OK, v0bits=0)
    add("neg_zero_float", env(mx("-0.0")), OK, v0bits=0x8000000000000000)
    add("nasty_small_double", env(mx("2.2250738585072012e-308")), OK, v0bits=1)
    add("nasty_small_double_upper_e", env(mx("2.2250738585072012E-308")), OK, v0=2.2250738585072012e-308)
    add("big_int_value", env(mx("123456789012345678901234567890")), OK, v0=1.2345678901234568e29)
    add("long_min", env(mx("-9223372036854775808")), OK, v0=-9.223372036854775808e18)
    add("huge_exponent_inf", env(mx("1e400")), OK, v0=float("inf"))
    add("tiny_exponent_zero", env(mx("-1e-400")), OK, v0bits=0x8000000000000000)
    add("value_null_is_zero", env(mx("null")), OK, v0=0.0)
    add("value_true", env(mx("true")), DECODE)
    add("value_array", env(mx("[1]")), DECODE)
    add("value_object", env(mx("{}")), DECODE)
    add("value_string_number", env(mx('"1.5"')), OK, v0=1.5)
    add("value_string_trimmed", env(mx('" 7e2 "')), OK, v0=700.0)
    add("value_string_plus_dot", env(mx('"+.5"')), OK, v0=0.5)
    add("value_string_trailing_dot", env(mx('"5."')), OK, v0=5.0)
    add("value_string_neg_zero", env(mx('"-0"')), OK, v0bits=0x8000000000000000)
    add("value_string_leading_zeros", env(mx('"007.50"')), OK, v0=7.5)
    add("value_string_empty", env(mx('""')), OK, v0=0.0)
    add("value_string_null", env(mx('"null"')), OK, v0=0.0)
    add("value_string_nan", env(mx('"NaN"')), OK, v0bits=0x7FF8000000000000)
    add("value_string_inf", env(mx('"-Infinity"')), OK, v0=float("-inf"))
    add("value_string_INF", env(mx('"INF"')), OK, v0=float("inf"))
    add("value_string_hex_passthrough", env(mx('"0x1p3"')), PASS)
    add("value_string_suffix_passthrough", env(mx('"1.5d"')), PASS)
    add("value_string_garbage_passthrough", env(mx('"abc"')), PASS)
    add("value_string_garbage_and_unknown_prop", env(mx('"abc"', extra=',"bogus":1')), DECODE)
    add("value_string_escaped_passthrough", env(mx('"1\\u002e5"')), PASS)
    add("location_nulls", env('{"latitude":null,"longitude":1.5}', type_="DeviceLocation"), OK, has_v0=False, has_v1=True,
        has_v2=False, v1=1.5)
    add("location_empty_string_is_null", env('{"latitude":"","longitude":"2","elevation":"null"}', type_="DeviceLocation"), OK,
        has_v0=False, has_v1=True, has_v2=False, v1=2.0)
    add("location_int_coords", env('{"latitude":34,"longitude":-84,"elevation":0}', type_="DeviceLocation"), OK, v0=34.0,
        v1=-84.0, v2=0.0)
    # ---- booleans --------------------------------------------------------------------------------------------------------
    for txt, val in (("true", True), ("false", False), ("null", False), ("1", True), ("0", False), ("-0", False), ("17", True),
                     ('"true"', True), ('"True"', True), ('"false"', False), ('"False"', False), ('" true "', True),
                     ('""', False), ('"null"', False)):
        add("updateState_" + txt, env('{"name":"a","value":1,"updateState":%s}' % txt), OK, update=val)
    for txt in ('"TRUE"', '"yes"', "1.0", "[true]", "{}"):
        add("updateState_bad_" + txt, env('{"name":"a","value":1,"updateState":%s}' % txt), DECODE)
    # ---- enums -----------------------------------------------------------------------------------------------------------
    add("eventType_other", env(mx().replace('"Measurement"', '"Location"')), OK, etype=1)
    add("eventType_null", env(mx().replace('"Measurement"', "null")), OK, etype=None)
    add("eventType_index", env(mx().replace('"Measurement"', "5")), OK, etype=5)
    add("eventType_index_string", env(mx().replace('"Measurement"', '" 2 "')), OK, etype=2)
    add("eventType_index_oob", env(mx().replace('"Measurement"', "6")), DECODE)
    add("eventType_index_negative", env(mx().replace('"Measurement"', "-1")), DECODE)
    add("eventType_index_wraps_int", env(mx().replace('"Measurement"', "4294967297")), OK, etype=1)
    add("eventType_bad_name", env(mx().replace('"Measurement"', '"measurement"')), DECODE)
    add("eventType_padded_name", env(mx().replace('"Measurement"', '" Measurement"')), DECODE)
    add("eventType_float", env(mx().replace('"Measurement"', "1.0")), DECODE)
    add("eventType_bool", env(mx().replace('"Measurement"', "true")), DECODE)
    add("eventType_empty", env(mx().replace('"Measurement"', '""')), DECODE)
    add("alert_levels", S.json_alert(TOK, "t", "m", D0, level="Critical", source="System"), OK, level=3, source=1)
    add("alert_level_index", S.json_alert(TOK, "t", "m", D0).replace(b'"Info"', b"2"), OK, level=2)
    add("alert_level_null", S.json_alert(TOK, "t", "m", D0).replace(b'"Info"', b"null"), OK, null_level=True)
    add("alert_source_null", S.json_alert(TOK, "t", "m", D0).replace(b'"Device"', b"null"), OK, null_source=True)
    add("alert_level_bad", S.json_alert(TOK, "t", "m", D0, level="Fatal"), DECODE)
    add("alert_defaults", env('{"type":"t","message":"m"}', type_="DeviceAlert"), OK, level=0, source=0, name=b"t")
    # ---- strings ---------------------------------------------------------------------------------------------------------
    add("name_null", env(mx(name="null")), OK, name=None)
    add("name_number_passthrough", env(mx(name="15")), PASS)
    add("name_bool_passthrough", env(mx(name="true")), PASS)
    add("name_object", env(mx(name="{}")), DECODE)
    add("name_array", env(mx(name='["a"]')), DECODE)
    add("message_number_coerced", S.json_alert(TOK, "t", "X", D0).replace(b'"X"', b"1.50"), OK, coerced=True, message=b"1.50")
    add("alternateId_bool_coerced", env(mx(extra=',"alternateId":false')), OK, coerced=True, alt=b"false")
    add("empty_name", env(mx(name='""')), OK, name=b"")
    # ---- dates ------------------------------------------------------------------------------------------------------------
    def dt(text):
        return env('{"name":"a","value":1,"eventDate":%s}' % text)
    add("date_epoch_int", dt("1527854400123"), OK, date=1527854400123)
    add("date_negative_int", dt("-5"), OK, date=-5)
    add("date_bigint_wraps", dt("18446744073709551617"), OK, date=1)
    add("date_float", dt("1.5e3"), DECODE)
    add("date_bool", dt("true"), DECODE)
    add("date_null", dt("null"), OK, date=None)
    add("date_empty_string", dt('""'), OK, date=None)
    add("date_iso_z", dt('"2018-06-01T12:00:00.123Z"'), OK, date=1527854400123)
    add("date_iso_no_ms", dt('"2018-06-01T12:00:00Z"'), OK, date=1527854400000)
    add("date_iso_no_seconds", dt('"2018-06-01T12:00Z"'), OK, date=1527854400000)
    add("date_iso_no_zone_is_utc", dt('"2018-06-01T12:00:00.000"'), OK, date=1527854400000)
    add("date_iso_offset_colon", dt('"2018-06-01T14:00:00.000+02:00"'), OK, date=1527854400000)
    add("date_iso_offset_nocolon", dt('"2018-06-01T07:30:00.000-0430"'), OK, date=1527854400000)
    add("date_iso_offset_hours", dt('"2018-06-01T13:00:00+01"'), OK, date=1527854400000)
    add("date_iso_frac1", dt('"2018-06-01T12:00:00.5Z"'), OK, date=1527854400500)
    add("date_iso_frac2", dt('"2018-06-01T12:00:00.12Z"'), OK, date=1527854400120)
    add("date_iso_frac6_truncates", dt('"2018-06-01T12:00:00.123999Z"'), OK, date=1527854400123)
    add("date_iso_frac10_bad", dt('"2018-06-01T12:00:00.1234567890Z"'), DECODE)
    add("date_plain", dt('"2018-06-01"'), OK, date=1527811200000)
    add("date_trimmed", dt('" 2018-06-01T12:00:00Z "'), OK, date=1527854400000)
    add("date_lenient_overflow", dt('"2018-13-32T25:61:61Z"'), OK, date=utc_ms(2019, 1, 1, 25, 61, 61, days=31))
    add("date_lenient_zero_month_day", dt('"2018-00-00T00:00:00Z"'), OK, date=utc_ms(2017, 12, 1, days=-1))
    add("date_leap", dt('"2020-02-29T23:59:59.999Z"'), OK, date=utc_ms(2020, 2, 29, 23, 59, 59, 999))
    add("date_iso_bad_sep", dt('"2018-06-01 12:00:00Z"'), DECODE)
    add("date_iso_short", dt('"2018-06-1"'), DECODE)
    add("date_iso_bad_zone", dt('"2018-06-01T12:00:00+5"'), DECODE)
    add("date_iso_lower_t", dt('"2018-06-01t12:00:00Z"'), DECODE)
    add("date_digits_string", dt('"1527854400123"'), OK, date=1527854400123)
    add("date_negative_digits_string", dt('"-17"'), OK, date=-17)
    add("date_minus_only_string", dt('"-"'), DECODE)
    add("date_digits_too_long_rfc", dt('"99999999999999999999"'), PASS)
    add("date_rfc1123_passthrough", dt('"Fri, 01 Jun 2018 12:00:00 GMT"'), PASS)
    add("date_garbage_passthrough", dt('"yesterday"'), PASS)
    add("date_old_year_passthrough", dt('"1500-01-01T00:00:00Z"'), PASS)
    add("date_nonascii_passthrough", dt('"2018-06-01T12:00:00Z' + chr(0xA0) + '"'), PASS)
    add("date_object", dt("{}"), DECODE)
    return c
