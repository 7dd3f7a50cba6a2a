"""GPU parity: swgpu_decode_pb through the C ABI vs the CPU oracle, bit for bit."""
import numpy as np
import pytest

import oracle
from sitewhere_b200 import synth as S
from sitewhere_b200 import _lib as L
import sitewhere_b200 as sw
from tests import pb_cases
from tests.helpers import check_decode_parity

pytestmark = pytest.mark.gpu


def test_locations_c2_shape(ctx):
    vo, zlon, zlat, cond = S.star_zones(8)
    lon, lat = S.locations_for_zones(100_000, vo, zlon, zlat, seed=S.SEED_BASE + 1)
    data, rec_off, truth = S.pb_locations(100_000, 5000, lon, lat, seed=S.SEED_BASE + 1)
    res = ctx.decode_pb(data, rec_off)
    check_decode_parity(res, data, rec_off)
    assert res.n_rows == 100_000
    assert np.array_equal(res.cols.v0[:res.n_rows].view(np.uint64), lat.view(np.uint64))
    assert np.array_equal(res.cols.v1[:res.n_rows].view(np.uint64), lon.view(np.uint64))


def test_mixed_c3_shape(ctx):
    data, rec_off, truth = S.pb_mixed(200_000, 10_000, seed=S.SEED_BASE + 2)
    res = ctx.decode_pb(data, rec_off)
    check_decode_parity(res, data, rec_off)


def test_adversarial_cases(ctx):
    cs = pb_cases.cases()
    recs = [c[1] for c in cs]
    data, rec_off = S.pack_records(recs)
    res = ctx.decode_pb(data, rec_off)
    check_decode_parity(res, data, rec_off)
    for i, (name, rec, st, rows) in enumerate(cs):
        assert int(res.status[i]) == st, name


def test_adversarial_mixed_into_bulk(ctx):
    """bad records in the middle of tiles must not disturb their neighbours' rows"""
    rng = np.random.default_rng(5)
    good, goff, _ = S.pb_mixed(3000, 100, seed=9)
    raw = good.tobytes()
    recs = [raw[int(goff[i]): int(goff[i + 1])] for i in range(3000)]
    cs = pb_cases.cases()
    for k in range(600):
        recs.insert(int(rng.integers(0, len(recs))), cs[k % len(cs)][1])
    data, rec_off = S.pack_records(recs)
    res = ctx.decode_pb(data, rec_off)
    check_decode_parity(res, data, rec_off)


def test_random_corruption_fuzz(ctx):
    """flip / truncate bytes of valid payloads: every record's status and rows must match the oracle"""
    rng = np.random.default_rng(77)
    good, goff, _ = S.pb_mixed(4000, 100, seed=10)
    raw = good.tobytes()
    recs = []
    for i in range(4000):
        r = bytearray(raw[int(goff[i]): int(goff[i + 1])])
        mode = i % 4
        if mode == 0:
            r[int(rng.integers(0, len(r)))] = int(rng.integers(0, 256))
        elif mode == 1:
            r = r[: int(rng.integers(0, len(r) + 1))]
        elif mode == 2:
            for _ in range(3):
                r[int(rng.integers(0, len(r)))] ^= 1 << int(rng.integers(0, 8))
        else:
            p = int(rng.integers(0, len(r)))
            r[p:p] = bytes(rng.integers(0, 256, size=int(rng.integers(1, 6)), dtype=np.uint8))
        recs.append(bytes(r))
    data, rec_off = S.pack_records(recs)
    res = ctx.decode_pb(data, rec_off)
    ocols, ostatus = check_decode_parity(res, data, rec_off)
    assert len(set(ostatus.tolist())) >= 2     # the fuzz does produce failures


def test_fanout_capacity_and_large_records(ctx):
    recs = [S.pb_measurements_record("tok-%d" % i, [("m%d" % k, float(k)) for k in range(i % 7)], date=i) for i in range(500)]
    recs.append(S.pb_measurements_record("big", [("name-%04d" % k, k * 0.5) for k in range(3000)], date=1))  # > staged tile
    recs.append(S.pb_location_record("after-big", 1.0, 2.0))
    data, rec_off = S.pack_records(recs)
    res = ctx.decode_pb(data, rec_off, max_rows=8)      # forces the capacity retry
    check_decode_parity(res, data, rec_off)
    assert res.n_rows == sum(i % 7 for i in range(500)) + 3000 + 1


def test_empty_batch_and_mirror_api(ctx):
    res = ctx.decode_pb(np.zeros(0, np.uint8), np.zeros(1, np.uint64))
    assert res.n_rows == 0 and res.first_row.tolist() == [0]
    dec = sw.ProtobufDeviceEventDecoder(ctx)
    out = dec.decode(S.pb_location_record("25683-LAIPAC-S911-7539094", 34.10469794977326, -84.23966646194458,
                                          elev=0.0, date=1527854400000, update_state=True))
    assert len(out) == 1 and out[0].deviceToken == "25683-LAIPAC-S911-7539094"
    assert out[0].request["latitude"] == 34.10469794977326 and out[0].request["updateState"] is True
    with pytest.raises(sw.EventDecodeException):
        dec.decode(b"\x02\x08\x03\x05abc")
    with pytest.raises(sw.DecoderRuntimeException):
        dec.decode(b"")
    out = dec.decode(S.pb_measurements_record("t", [("a", 1.0), ("b", 2.0)], date=3))
    assert [(o.request["name"], o.request["value"]) for o in out] == [("a", 1.0), ("b", 2.0)]
