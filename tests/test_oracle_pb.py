"""Pins the protobuf half of the CPU oracle (oracle/sw_oracle.c: orc_decode_pb) against
(1) google.protobuf's own runtime parsing a hand-built descriptor of sitewhere.proto
    (sitewhere-communication/src/main/proto/sitewhere.proto:1-109) — an independent implementation
    of the wire format, and
(2) known answers for malformed payloads per ProtobufDeviceEventDecoder.java:75-304.
CPU only."""
import struct

import numpy as np
import pytest

import oracle
from sitewhere_b200 import synth as S
from sitewhere_b200 import _lib as L
from tests import pb_cases


def _descriptor_pool():
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

    fdp = descriptor_pb2.FileDescriptorProto()
    fdp.name = "sitewhere.proto"
    fdp.syntax = "proto2"
    F = descriptor_pb2.FieldDescriptorProto

    def msg(parent, name, fields):
        m = parent.message_type.add() if hasattr(parent, "message_type") else parent.nested_type.add()
        m.name = name
        for (fname, num, label, ftype, tname) in fields:
            f = m.field.add()
            f.name, f.number, f.label, f.type = fname, num, label, ftype
            if tname:
                f.type_name = tname
        return m

    REQ, OPT, REP = F.LABEL_REQUIRED, F.LABEL_OPTIONAL, F.LABEL_REPEATED
    model = msg(fdp, "Model", [])
    msg(model, "Metadata", [("name", 1, REQ, F.TYPE_STRING, ""), ("value", 2, REQ, F.TYPE_STRING, "")])
    msg(model, "DeviceLocation", [("hardwareId", 1, REQ, F.TYPE_STRING, ""), ("latitude", 2, REQ, F.TYPE_DOUBLE, ""),
                                  ("longitude", 3, REQ, F.TYPE_DOUBLE, ""), ("elevation", 4, OPT, F.TYPE_DOUBLE, ""),
                                  ("eventDate", 5, OPT, F.TYPE_FIXED64, ""),
                                  ("metadata", 6, REP, F.TYPE_MESSAGE, ".Model.Metadata"),
                                  ("updateState", 7, OPT, F.TYPE_BOOL, "")])
    msg(model, "DeviceAlert", [("hardwareId", 1, REQ, F.TYPE_STRING, ""), ("alertType", 2, REQ, F.TYPE_STRING, ""),
                               ("alertMessage", 3, REQ, F.TYPE_STRING, ""), ("eventDate", 4, OPT, F.TYPE_FIXED64, ""),
                               ("metadata", 5, REP, F.TYPE_MESSAGE, ".Model.Metadata"),
                               ("updateState", 6, OPT, F.TYPE_BOOL, "")])
    msg(model, "Measurement", [("measurementId", 1, REQ, F.TYPE_STRING, ""),
                               ("measurementValue", 2, REQ, F.TYPE_DOUBLE, "")])
    msg(model, "DeviceMeasurements", [("hardwareId", 1, REQ, F.TYPE_STRING, ""),
                                      ("measurement", 2, REP, F.TYPE_MESSAGE, ".Model.Measurement"),
                                      ("eventDate", 3, OPT, F.TYPE_FIXED64, ""),
                                      ("metadata", 4, REP, F.TYPE_MESSAGE, ".Model.Metadata"),
                                      ("updateState", 5, OPT, F.TYPE_BOOL, "")])
    sw = msg(fdp, "SiteWhere", [])
    en = sw.enum_type.add()
    en.name = "Command"
    for i, nm in enumerate(["SEND_REGISTRATION", "SEND_ACKNOWLEDGEMENT", "SEND_DEVICE_LOCATION", "SEND_DEVICE_ALERT",
                            "SEND_DEVICE_MEASUREMENTS", "SEND_DEVICE_STREAM", "SEND_DEVICE_STREAM_DATA",
                            "REQUEST_DEVICE_STREAM_DATA"], start=1):
        v = en.value.add()
        v.name, v.number = nm, i
    msg(sw, "Header", [("command", 1, REQ, F.TYPE_ENUM, ".SiteWhere.Command"), ("originator", 2, OPT, F.TYPE_STRING, "")])
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fdp)
    get = lambda n: message_factory.GetMessageClass(pool.FindMessageTypeByName(n))
    return dict(Header=get("SiteWhere.Header"), DeviceLocation=get("Model.DeviceLocation"),
                DeviceAlert=get("Model.DeviceAlert"), DeviceMeasurements=get("Model.DeviceMeasurements"))


def _read_delimited(buf, pos):
    n = 0
    shift = 0
    while True:
        b = buf[pos]
        pos += 1
        n |= (b & 0x7F) << shift
        shift += 7
        if not b & 0x80:
            break
    return buf[pos: pos + n], pos + n


def _bits(x):
    return struct.unpack("<Q", struct.pack("<d", x))[0]


def test_wellformed_against_protobuf_runtime():
    cls = _descriptor_pool()
    data, rec_off, truth = S.pb_mixed(3000, 200, seed=S.SEED_BASE + 2)
    cols, status, first_row, total = oracle.decode_pb(data, rec_off)
    assert (status == 0).all() and total == 3000
    raw = data.tobytes()
    for i in range(3000):
        rec = raw[int(rec_off[i]): int(rec_off[i + 1])]
        hb, p = _read_delimited(rec, 0)
        hdr = cls["Header"]()
        hdr.ParseFromString(hb)
        body, _ = _read_delimited(rec, p)
        r = int(first_row[i])
        m = int(cols.meta[r])
        tok = rec[int(cols.token_span[r]) & 0xFFFFFFFF:][: int(cols.token_span[r]) >> 32].decode()
        if hdr.command == 3:
            msg = cls["DeviceLocation"]()
            msg.ParseFromString(body)
            assert m & 0xFF == L.KIND_LOCATION
            assert _bits(cols.v0[r]) == _bits(msg.latitude) and _bits(cols.v1[r]) == _bits(msg.longitude)
            assert _bits(cols.v2[r]) == _bits(msg.elevation)
        elif hdr.command == 4:
            msg = cls["DeviceAlert"]()
            msg.ParseFromString(body)
            assert m & 0xFF == L.KIND_ALERT
            ns = int(cols.name_span[r])
            assert rec[ns & 0xFFFFFFFF:][: ns >> 32].decode() == msg.alertType
            ms = int(cols.message_span[r])
            assert rec[ms & 0xFFFFFFFF:][: ms >> 32].decode() == msg.alertMessage
        else:
            msg = cls["DeviceMeasurements"]()
            msg.ParseFromString(body)
            assert m & 0xFF == L.KIND_MEASUREMENT and len(msg.measurement) == 1
            ns = int(cols.name_span[r])
            assert rec[ns & 0xFFFFFFFF:][: ns >> 32].decode() == msg.measurement[0].measurementId
            assert _bits(cols.v0[r]) == _bits(msg.measurement[0].measurementValue)
        assert tok == msg.hardwareId
        assert int(cols.event_date[r]) == msg.eventDate
        assert int(cols.token_hash[r]) & 0xFFFFFFFF == oracle.murmur2(msg.hardwareId.encode())


def test_known_answers():
    cs = pb_cases.cases()
    data, rec_off = S.pack_records([c[1] for c in cs])
    cols, status, first_row, total = oracle.decode_pb(data, rec_off)
    for i, (name, rec, st, rows) in enumerate(cs):
        assert int(status[i]) == st, f"{name}: status {status[i]} != {st}"
        got = int(first_row[i + 1]) - int(first_row[i])
        assert got == (rows or 0), f"{name}: rows {got} != {rows}"


def test_field_semantics():
    cls = _descriptor_pool()
    cs = {c[0]: c for c in pb_cases.cases()}
    names = ["loc_full", "mx_three", "repeated_scalars_last_wins", "loc_negzero_nan", "bool_multibyte_true",
             "bool_multibyte_false", "loc_update_false", "alert", "unknown_fields_skipped"]
    data, rec_off = S.pack_records([cs[n][1] for n in names])
    cols, status, first_row, total = oracle.decode_pb(data, rec_off)
    raw = data.tobytes()

    def row(name, k=0):
        i = names.index(name)
        return int(first_row[i]) + k, raw[int(rec_off[i]): int(rec_off[i + 1])]

    r, rec = row("loc_full")
    fl = (int(cols.meta[r]) >> 8) & 0xFFFF
    assert fl & L.F_UPDATE_STATE and fl & L.F_HAS_DATE and fl & L.F_HAS_ORIGINATOR and fl & L.F_HAS_METADATA
    assert cols.v0[r] == 34.10469794977326 and cols.v1[r] == -84.23966646194458 and cols.v2[r] == 0.0
    assert cols.event_date[r] == 1527854400123
    os_ = int(cols.originator_span[r])
    assert rec[os_ & 0xFFFFFFFF:][: os_ >> 32] == b"orig-1"
    assert int(cols.metadata_span[r]) >> 32 == 1
    # three measurements share token/date/flags; order preserved
    vals = []
    for k in range(3):
        r, rec = row("mx_three", k)
        ns = int(cols.name_span[r])
        vals.append((rec[ns & 0xFFFFFFFF:][: ns >> 32], float(cols.v0[r])))
        assert (int(cols.meta[r]) >> 8) & L.F_UPDATE_STATE
        assert int(cols.meta[r]) >> 32 == names.index("mx_three")
    assert vals == [(b"a", 1.0), (b"bb", -2.5), (b"ccc", 1e300)]
    r, rec = row("repeated_scalars_last_wins")
    ts = int(cols.token_span[r])
    assert rec[ts & 0xFFFFFFFF:][: ts >> 32] == b"second" and cols.v0[r] == 9.0
    r, rec = row("loc_negzero_nan")
    assert _bits(cols.v0[r]) == _bits(-0.0) and np.isnan(cols.v1[r]) and np.isinf(cols.v2[r])
    assert not (int(cols.meta[r]) >> 8) & L.F_HAS_DATE   # reference substitutes new Date()
    r, _ = row("bool_multibyte_true")
    assert (int(cols.meta[r]) >> 8) & L.F_UPDATE_STATE
    r, _ = row("bool_multibyte_false")
    assert not (int(cols.meta[r]) >> 8) & L.F_UPDATE_STATE
    r, _ = row("loc_update_false")
    assert not (int(cols.meta[r]) >> 8) & L.F_UPDATE_STATE
    r, rec = row("alert")
    assert int(cols.metadata_span[r]) >> 32 == 2 and (int(cols.meta[r]) >> 24) & 3 == 0  # level Info
    r, rec = row("unknown_fields_skipped")
    assert cols.v0[r] == 1.0 and cols.v1[r] == 2.0


def test_row_capacity_and_empty():
    data, rec_off, _ = S.pb_mixed(50, 5, seed=1)
    cols, status, first_row, total = oracle.decode_pb(data, rec_off, max_rows=10)
    assert total == 50 and cols.n == 10
    cols, status, first_row, total = oracle.decode_pb(np.zeros(0, np.uint8), np.zeros(1, np.uint64))
    assert total == 0 and first_row[0] == 0
