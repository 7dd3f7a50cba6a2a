"""Pins the enriched-event half of the CPU oracle (oracle/sw_oracle_enriched.c: orc_decode_enriched) against
google.protobuf's runtime on a hand-built descriptor of device-event-model.proto / sitewhere-common.proto
(tests/enriched_proto.py): well-formed payloads of every event type, repeated fields (message merge), damaged
payloads, and hand-made known answers.  CPU only."""
import random
import struct

import numpy as np

import oracle
from sitewhere_b200 import synth as S
from sitewhere_b200 import _lib as L
from tests import enriched_proto as E


import re
_OVERLONG_VARINT32 = re.compile(rb"[\x80-\xff]{4}[\x10-\xff]")
lenient = [0]


def check_against_protobuf(recs, strict_meta=True):
    data, rec_off = S.pack_records(recs)
    cols, ids, status, first_row, total = oracle.decode_enriched(data, rec_off)
    n_ok = 0
    for i, rec in enumerate(recs):
        st, d = E.expected_row(rec)
        if not strict_meta and st == 1 and int(status[i]) != 1 and _OVERLONG_VARINT32.search(rec):
            # a tag or length varint with bits beyond 32: protobuf-java's readRawVarint32 drops them silently
            # (the oracle follows it), python's parser rejects the message
            n_ok += int(status[i]) == 0
            lenient[0] += 1
            continue
        assert int(status[i]) == st, (i, rec.hex(), int(status[i]), st)
        if st != 0:
            assert first_row[i + 1] == first_row[i]
            continue
        n_ok += 1
        r = int(first_row[i])
        m = int(cols.meta[r])
        fl = (m >> 8) & 0xFFFF
        assert m & 0xFF == d["kind"] and (m >> 28) & 7 == d["etype"] and (m >> 27) & 1 == 1 and m >> 32 == i
        assert bool(fl & L.F_HAS_DATE) == (d["date"] != 0) and int(cols.event_date[r]) == d["date"] - (1 << 64 if d["date"] >> 63 else 0)
        assert (int(ids.event_hi[r]), int(ids.event_lo[r])) == d["event_id"]
        assert (int(ids.assignment_hi[r]), int(ids.assignment_lo[r])) == d["assignment"]
        assert (int(ids.device_hi[r]), int(ids.device_lo[r])) == d["device"]
        key = struct.pack("<QQ", *d["assignment"])
        assert int(cols.token_hash[r]) == oracle.hash64(key) and int(cols.token_span[r]) == 0
        assert bool(fl & L.F_HAS_ALT_ID) == d["has_alt"]
        if d["has_alt"]:
            sp = int(cols.alternate_id_span[r])
            assert rec[sp & 0xFFFFFFFF:][: sp >> 32].decode() == d["alt"]
        # (python's parser keeps a map entry that holds an unknown field as an unknown field of the parent instead
        #  of an entry; protobuf-java skips the unknown field and puts the entry: not compared on damaged input)
        assert not strict_meta or bool(fl & L.F_HAS_METADATA) == (d["n_meta"] > 0)
        span = lambda s: rec[int(s) & 0xFFFFFFFF:][: int(s) >> 32]
        bits = lambda a: int(np.float64(a).view(np.uint64))
        if d["kind"] == L.KIND_MEASUREMENT:
            assert span(cols.name_span[r]).decode() == d["name"] and bits(cols.v0[r]) == d["v0"]
            assert fl & L.F_HAS_NAME and fl & L.F_HAS_V0
            assert int(cols.name_hash[r]) == oracle.hash64(d["name"].encode())
        elif d["kind"] == L.KIND_ALERT:
            assert span(cols.name_span[r]).decode() == d["name"] and span(cols.message_span[r]).decode() == d["message"]
            assert (m >> 24) & 3 == d["level"] and (m >> 26) & 1 == d["source"]
        else:
            assert bool(fl & L.F_HAS_V0) == d["has_lat"] and bool(fl & L.F_HAS_V1) == d["has_lon"] and bool(fl & L.F_HAS_V2) == d["has_elev"]
            assert bits(cols.v0[r]) == d["v0"] and bits(cols.v1[r]) == d["v1"] and bits(cols.v2[r]) == d["v2"]
            assert int(cols.name_span[r]) == 0 and int(cols.name_hash[r]) == 0
    assert total == n_ok
    return n_ok


def test_wellformed_payloads_of_every_type():
    rnd = random.Random(5)
    recs = [E.random_payload(rnd, kinds=(1, 2, 3, 6)).SerializeToString() for _ in range(4000)]
    assert check_against_protobuf(recs) > 2500


def test_repeated_fields_merge_as_protobuf_merges():
    """two serialised payloads back to back = every top-level field twice: message fields merge, scalars of the
    later one replace, a different oneof member replaces, the same member merges.  (Serialised proto3 messages
    never carry explicit default values, the one case in which protobuf-java's builder merge and a parse into the
    same object differ; -0.0, which python serialises and Java does not, is left out here for that reason.)"""
    rnd = random.Random(6)
    recs = []
    for _ in range(3000):
        a, b, c = (E.random_payload(rnd, sparse=0.5, neg_zero=False) for _ in range(3))
        recs.append(a.SerializeToString() + b.SerializeToString())
        ev = c.event.SerializeToString()
        if len(ev) < 128:       # a third occurrence of the event field alone
            recs.append(a.SerializeToString() + b.SerializeToString() + b"\x12" + bytes([len(ev)]) + ev)
    check_against_protobuf(recs)


def test_damaged_payloads():
    rnd = random.Random(7)
    rng = np.random.default_rng(7)
    base = [E.random_payload(rnd).SerializeToString() for _ in range(400)]
    recs = []
    specials = [0x00, 0x01, 0x08, 0x09, 0x0A, 0x0B, 0x0C, 0x12, 0x1A, 0x22, 0x7F, 0x80, 0xC0, 0xE0, 0xED, 0xF4, 0xFF, 0x03, 0x04]
    for _ in range(20000):
        r = bytearray(rnd.choice(base))
        for _ in range(rnd.randrange(1, 4)):
            if not r:
                break
            op = rnd.randrange(6)
            p = rnd.randrange(len(r))
            if op == 0:
                r[p] ^= 1 << rnd.randrange(8)
            elif op == 1:
                r[p] = rnd.choice(specials)
            elif op == 2:
                del r[p: p + rnd.randrange(1, 5)]
            elif op == 3:
                r[p:p] = bytes(rnd.choice(specials) for _ in range(rnd.randrange(1, 4)))
            elif op == 4:
                r = r[: p]
            else:
                o = rnd.choice(base)
                r = r[:p] + bytearray(o[rnd.randrange(len(o) + 1):])
        recs.append(bytes(r))
    recs += [b"", b"\x00", b"\x0a", b"\x0a\x00", b"\x12\x00", b"\x0c", b"\x0b\x0c", b"\x0b\x14", b"\x0f"]
    lenient[0] = 0
    n_ok = check_against_protobuf(recs, strict_meta=False)
    assert 1000 < n_ok < len(recs) - 1000 and lenient[0] < len(recs) // 100


def test_known_answers():
    C = E.classes()
    p = C["GEnrichedEventPayload"]()
    p.event.location.latitude.value = 34.10469794977326      # the MQTT test location of the reference (MqttTests.java:114-115)
    p.event.location.longitude.value = -84.23966646194458
    p.event.location.event.eventType = 1
    p.event.location.event.deviceAssignmentId.msb = 1
    p.event.location.event.deviceAssignmentId.lsb = 2
    good = p.SerializeToString()
    ok = lambda rec: int(oracle.decode_enriched(*S.pack_records([rec]))[2][0])
    assert ok(good) == 0
    assert ok(good + b"\x0c") == 1                    # END_GROUP at top level: checkLastTagWas(0)
    assert ok(good + b"\x00") == 1                    # tag 0
    assert ok(good + b"\x0b\x0c") == 0                # an unknown group is skipped
    assert ok(good + b"\x0b\x14") == 1                # ... unless its end tag names another field
    assert ok(good + b"\x1f") == 1                    # wire type 7
    assert ok(good[:-1]) == 1                          # truncated
    p2 = C["GEnrichedEventPayload"]()
    p2.CopyFrom(p)
    p2.event.location.event.eventType = 6              # UNRECOGNIZED -> SiteWhereException in asApiDeviceEventType
    assert ok(p2.SerializeToString()) == 1
    p3 = C["GEnrichedEventPayload"]()
    p3.context.assignmentStatus = 1
    assert ok(p3.SerializeToString()) == 1             # EVENT_NOT_SET -> "Unable to convert event to API"
    m = C["GEnrichedEventPayload"]()
    m.event.measurement.name = "ok"
    raw = m.SerializeToString().replace(b"ok", b"\xc0\x80")   # overlong encoding: readStringRequireUtf8 throws
    assert ok(raw) == 1
    m.event.measurement.name = "fuel.level"             # the reference's fixture measurement (fuel.level = 123.4)
    m.event.measurement.value = 123.4
    cols, ids, status, fr, total = oracle.decode_enriched(*S.pack_records([m.SerializeToString()]))
    assert total == 1 and float(cols.v0[0]) == 123.4 and int(cols.meta[0]) & 0xFF == 0
