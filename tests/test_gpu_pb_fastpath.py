"""The protobuf decoder recognises the exact layouts ProtobufDeviceEventEncoder emits and decodes them
on a short path; everything else takes the generic wire walker.  Sweep the boundary between the
two (length bytes reaching 128, optional fields appearing, fields reordered) and require parity
with the oracle either way."""
import numpy as np
import pytest

from sitewhere_b200 import synth as S
from tests.helpers import check_decode_parity

pytestmark = pytest.mark.gpu


def test_length_boundaries_and_near_canonical_layouts(ctx):
    recs = []
    for tl in list(range(0, 8)) + list(range(84, 96)) + [126, 127, 128, 129, 200]:
        tok = "T" * tl
        recs.append(S.pb_location_record(tok, 1.5, -2.5, elev=3.0, date=77))            # canonical
        recs.append(S.pb_location_record(tok, 1.5, -2.5, date=77))                      # no elevation
        recs.append(S.pb_location_record(tok, 1.5, -2.5, elev=3.0))                     # no date
        recs.append(S.pb_location_record(tok, 1.5, -2.5, elev=3.0, date=77, update_state=True))
        recs.append(S.pb_location_record(tok, 1.5, -2.5, elev=3.0, date=77, originator="o"))
        for nl in (0, 1, 5, 100, 115, 116, 117, 118, 127, 128):
            recs.append(S.pb_measurements_record(tok, [("n" * nl, 4.25)], date=9))
            recs.append(S.pb_alert_record(tok, "a" * nl, "m" * (nl // 2), date=9))
        recs.append(S.pb_measurements_record(tok, [("a", 1.0), ("b", 2.0)], date=9))
        recs.append(S.pb_measurements_record(tok, [("a", 1.0)]))                         # no date
    # reordered fields: same content, not the encoder's order
    m = S.pb_field_f64(3, -2.5) + S.pb_field_len(1, b"tok") + S.pb_field_fixed64(5, 7) + S.pb_field_f64(2, 1.5) + S.pb_field_f64(4, 0.0)
    recs.append(S.pb_header(3) + S.pb_delimited(m))
    # canonical size but a wrong tag inside
    good = bytearray(S.pb_location_record("x" * 21, 1.0, 2.0, elev=3.0, date=4))
    for pos in (0, 1, 2, 3, 4, 5, 27, 36, 45, 54):
        for delta in (1, 0x80):
            b = bytearray(good)
            b[pos] = (b[pos] + delta) & 0xFF
            recs.append(bytes(b))
    data, rec_off = S.pack_records(recs)
    res = ctx.decode_pb(data, rec_off)
    check_decode_parity(res, data, rec_off)
