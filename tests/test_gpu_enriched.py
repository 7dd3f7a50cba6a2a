"""GPU parity: the enriched-event payload (GEnrichedEventPayload, what the rule-processing and device-state
consumers read off Kafka) through the C ABI vs the CPU oracle, bit for bit: decode (rows, ids, status, order),
zone test on the decoded rows, and the fused decode + device-state merge keyed by assignment id."""
import random

import numpy as np
import pytest

import oracle
from sitewhere_b200 import synth as S
from sitewhere_b200 import _lib as L
from tests import enriched_proto as E
from tests.helpers import assert_cols_equal
from tests.test_gpu_merge import snap_equal

pytestmark = pytest.mark.gpu


def _damage(rnd, base, n_out):
    specials = [0x00, 0x01, 0x08, 0x09, 0x0A, 0x0B, 0x0C, 0x12, 0x1A, 0x22, 0x7F, 0x80, 0xC0, 0xE0, 0xED, 0xF4, 0xFF, 0x03, 0x04]
    out = []
    for _ in range(n_out):
        r = bytearray(rnd.choice(base))
        for _ in range(rnd.randrange(1, 4)):
            if not r:
                break
            op, p = rnd.randrange(6), rnd.randrange(len(r))
            if op == 0:
                r[p] ^= 1 << rnd.randrange(8)
            elif op == 1:
                r[p] = rnd.choice(specials)
            elif op == 2:
                del r[p: p + rnd.randrange(1, 5)]
            elif op == 3:
                r[p:p] = bytes(rnd.choice(specials) for _ in range(rnd.randrange(1, 4)))
            elif op == 4:
                r = r[:p]
            else:
                o = rnd.choice(base)
                r = r[:p] + bytearray(o[rnd.randrange(len(o) + 1):])
        out.append(bytes(r))
    return out


def _check(ctx, recs):
    data, rec_off = S.pack_records(recs)
    res, ids = ctx.decode_enriched(data, rec_off)
    ocols, oids, ostatus, ofirst, total = oracle.decode_enriched(data, rec_off)
    assert res.n_rows == total
    assert np.array_equal(res.status, ostatus), np.nonzero(res.status != ostatus)[0][:10]
    assert np.array_equal(res.first_row, ofirst)
    assert_cols_equal(res.cols, ocols, total)
    for k in L.ID_NAMES:
        assert np.array_equal(ids[k], getattr(oids, k)[:total]), k
    return res, ids, (data, rec_off, ocols, oids, total)


def test_wellformed_merged_and_damaged_payloads(ctx):
    rnd = random.Random(15)
    base = [E.random_payload(rnd, kinds=(1, 2, 3, 6)).SerializeToString() for _ in range(6000)]
    res, _, _ = _check(ctx, base)
    assert 4000 < res.n_rows < 6000
    merged = [rnd.choice(base) + rnd.choice(base) for _ in range(4000)]      # every top-level field twice
    _check(ctx, merged)
    res, _, _ = _check(ctx, _damage(rnd, base[:500], 40000) + [b"", b"\x00", b"\x0a", b"\x12\x00", b"\x0c", b"\x0b\x0c", b"\x0b\x14"])
    assert 2000 < int((res.status == 0).sum()) < 38000
    big = [rnd.choice(base) for _ in range(200_000)]                           # many tiles, dense and ordered passes
    _check(ctx, big)
    _check(ctx, [b for b in big if E.expected_row(b)[0] == 0][:50_000])        # an all-rows batch: the dense pass alone


def test_long_strings_and_records_beyond_the_staging_tile(ctx):
    C = E.classes()
    recs = []
    for k in range(300):
        p = C["GEnrichedEventPayload"]()
        p.event.alert.type = "t" * (k * 7 % 500)
        p.event.alert.alertMessage = "é" * (k * 13 % 3000)          # non-ASCII, records of several KB
        p.event.alert.event.deviceAssignmentId.lsb = k % 5
        p.event.alert.event.eventDate = 1 + k
        for j in range(k % 4):
            p.event.alert.event.metadata["key%d" % j] = "v" * (50 * j)
        recs.append(p.SerializeToString())
    _check(ctx, recs)


def test_zone_test_and_state_merge_on_enriched_events(ctx):
    rnd = random.Random(16)
    zones = S.star_zones(8, nvert=32, cond_mode="alternate")
    ctx.set_zones(*zones)
    C = E.classes()
    recs = []
    lon, lat = S.locations_for_zones(20_000, *zones[:3], seed=5, boundary_rate=0.01)
    for i in range(20_000):
        if i % 3 == 0:
            p = C["GEnrichedEventPayload"]()
            p.event.location.latitude.value = float(lat[i])
            p.event.location.longitude.value = float(lon[i])
            p.event.location.event.eventType = 1
            p.event.location.event.deviceAssignmentId.msb = 7
            p.event.location.event.deviceAssignmentId.lsb = rnd.randrange(300)
            p.event.location.event.id.msb = i
            p.event.location.event.eventDate = 1_527_854_400_000 + i
            recs.append(p.SerializeToString())
        else:
            recs.append(E.random_payload(rnd, n_assign=300).SerializeToString())
    recs += _damage(rnd, recs[:300], 500)
    rnd.shuffle(recs)
    data, rec_off = S.pack_records(recs)
    # decode + zone test in one host-buffer call
    cols, status, first_row, n_rows, pairs = ctx.decode_zone_test("enriched", data, rec_off)
    ocols, oids, ostatus, ofirst, total = oracle.decode_enriched(data, rec_off)
    assert n_rows == total and np.array_equal(status, ostatus) and np.array_equal(first_row, ofirst)
    assert_cols_equal(cols, ocols, total)
    want, tp = oracle.zone_test(*zones, ocols.v1[:total], ocols.v0[:total], meta=ocols.meta[:total])
    assert pairs.shape[0] == tp and tp > 1000 and np.array_equal(pairs, want)
    # fused decode + merge keyed by assignment id, in three batches; then snapshot / restore
    ctx.state_clear()
    st = oracle.State()
    seq = 0
    n = len(recs)
    for a, b in ((0, n // 3), (n // 3, 2 * n // 3), (2 * n // 3, n)):
        d, o = S.pack_records(recs[a:b])
        stt = np.zeros(b - a, np.uint8)
        rows = ctx.decode_merge("enriched", d, o, seq, None, None, stt)
        oc, oi, ost, _, tot = oracle.decode_enriched(d, o)
        assert rows == tot and np.array_equal(stt, ost)
        st.merge_enriched(d, o, oc, oi, tot, seq)
        seq += (b - a) << 12
    snap = ctx.state_snapshot()
    snap_equal(snap, st.snapshot())
    assert snap["seq"].shape[0] > 300
    ctx.state_restore(snap)
    snap_equal(ctx.state_snapshot(), st.snapshot())
    ctx.state_clear()


def test_forward_pass_equals_general_pass(ctx):
    """the forward pass over the serialiser's own field order (option enriched.fast_path, default on) and the general
    proto3 parser must give identical results: serialised payloads (taken by the forward pass), the same concatenated
    (fields twice: merge), damaged, with multi-byte lengths and non-ASCII strings, and the bench generator's records"""
    rnd = random.Random(17)
    base = [E.random_payload(rnd, kinds=(1, 2, 3, 6)).SerializeToString() for _ in range(3000)]
    lon, lat = S.locations_for_zones(4000, *S.star_zones(4)[:3], seed=6)
    gd, goff, truth = S.enriched_locations(4000, 100, lon, lat, seed=6)
    graw = gd.tobytes()
    recs = base + [rnd.choice(base) + rnd.choice(base) for _ in range(1000)] + _damage(rnd, base[:300], 8000)
    recs += [graw[int(goff[i]): int(goff[i + 1])] for i in range(4000)]
    C = E.classes()
    for k in range(100):
        p = C["GEnrichedEventPayload"]()
        p.context.deviceStatus.value = "status" * (k % 30)
        for j in range(k % 5):
            p.context.deviceMetadata["k%d" % j] = "v" * (40 * j)
        p.context.assignmentStatus = k % 4
        p.event.measurement.name = ("n" * (k % 200)) if k % 7 else "é"
        p.event.measurement.value = float(k) - 3.5
        p.event.measurement.event.eventDate = k * 1_000_003
        p.event.measurement.event.deviceAssignmentId.lsb = k
        recs.append(p.SerializeToString())
    rnd.shuffle(recs)
    _, _, (data, rec_off, ocols, oids, total) = _check(ctx, recs)          # forward pass + general pass vs oracle
    fast, fids = ctx.decode_enriched(data, rec_off)
    ctx.set_option("enriched.fast_path", 0)
    try:
        slow, sids = ctx.decode_enriched(data, rec_off)
        _check(ctx, recs)                                                   # general pass alone vs oracle
    finally:
        ctx.set_option("enriched.fast_path", 1)
    assert fast.n_rows == slow.n_rows and np.array_equal(fast.status, slow.status) and np.array_equal(fast.first_row, slow.first_row)
    assert_cols_equal(fast.cols, slow.cols, fast.n_rows)
    for k in L.ID_NAMES:
        assert np.array_equal(fids[k], sids[k]), k
    # the generator's records decode to the generator's truth
    d2, o2 = S.pack_records([graw[int(goff[i]): int(goff[i + 1])] for i in range(4000)])
    res, ids = ctx.decode_enriched(d2, o2)
    assert res.n_rows == 4000 and np.array_equal(res.cols.v0[:4000], truth["lat"]) and np.array_equal(res.cols.v1[:4000], truth["lon"])
    assert np.array_equal(res.cols.event_date[:4000], truth["date"]) and np.array_equal(ids["event_hi"], truth["event_id"])
