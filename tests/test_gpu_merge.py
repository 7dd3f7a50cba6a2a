"""GPU parity: swgpu_merge (device-state last-writer fold) vs the oracle's sequential fold."""
import numpy as np
import pytest

import oracle
from sitewhere_b200 import synth as S
import sitewhere_b200 as sw

pytestmark = pytest.mark.gpu


def snap_equal(a, b):
    for k in ("seq", "id_hi", "id_lo", "event_date", "key_off", "key_bytes"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["value"].view(np.uint64), b["value"].view(np.uint64))


def test_merge_batches_match_sequential_fold(ctx):
    ctx.state_clear()
    st = oracle.State()
    logic = sw.DeviceStateProcessingLogic(ctx)
    seq = 0
    for b in range(4):
        data, rec_off, _ = S.pb_mixed(50_000, 2000, seed=200 + b)
        res = ctx.decode_pb(data, rec_off)
        ocols, _, _, total = oracle.decode_pb(data, rec_off)
        id_hi = np.full(total, b, np.uint64)
        id_lo = np.arange(total, dtype=np.uint64)
        logic.process(data, rec_off, res.cols, id_hi, id_lo)
        st.merge(data, rec_off, ocols, seq, id_hi, id_lo)
        seq += total
    snap_equal(ctx.state_snapshot(), st.snapshot())
    n, kb = ctx.state_size()
    assert n == st.snapshot()["seq"].shape[0]


def test_hot_key_contention_and_growth(ctx):
    """all events on 3 devices (heavy same-key contention) then a burst of new keys forcing a rehash"""
    ctx.state_clear()
    st = oracle.State()
    seq = 0
    for b, (n, ndev) in enumerate([(100_000, 3), (100_000, 40_000), (20_000, 3)]):
        data, rec_off, _ = S.pb_mixed(n, ndev, seed=300 + b)
        res = ctx.decode_pb(data, rec_off)
        ocols, _, _, total = oracle.decode_pb(data, rec_off)
        id_hi = np.full(total, 7 + b, np.uint64)
        id_lo = np.arange(total, dtype=np.uint64)[::-1].copy()
        ctx.merge(data, rec_off, res.cols, seq, id_hi, id_lo)
        st.merge(data, rec_off, ocols, seq, id_hi, id_lo)
        seq += total
    snap_equal(ctx.state_snapshot(), st.snapshot())


def test_snapshot_restore_roundtrip_and_replay_idempotence(ctx):
    ctx.state_clear()
    data, rec_off, _ = S.pb_mixed(30_000, 500, seed=400)
    res = ctx.decode_pb(data, rec_off)
    ids = np.arange(res.n_rows, dtype=np.uint64)
    ctx.merge(data, rec_off, res.cols, 0, ids, ids)
    snap = ctx.state_snapshot()
    ctx.merge(data, rec_off, res.cols, 0, ids, ids)            # replay of the same batch: idempotent
    snap_equal(ctx.state_snapshot(), snap)
    ctx.state_clear()
    assert ctx.state_size()[0] == 0
    ctx.state_restore(snap)
    snap_equal(ctx.state_snapshot(), snap)
    # merging a further batch after a restore behaves as if the state had never left the GPU
    d2, o2, _ = S.pb_mixed(10_000, 500, seed=401)
    r2 = ctx.decode_pb(d2, o2)
    ids2 = np.arange(r2.n_rows, dtype=np.uint64) + 10**6
    ctx.merge(d2, o2, r2.cols, 30_000, ids2, ids2)
    st = oracle.State()
    oc, _, _, t = oracle.decode_pb(data, rec_off)
    st.merge(data, rec_off, oc, 0, ids, ids)
    oc2, _, _, t2 = oracle.decode_pb(d2, o2)
    st.merge(d2, o2, oc2, 30_000, ids2, ids2)
    snap_equal(ctx.state_snapshot(), st.snapshot())
    ctx.state_clear()


def test_colliding_bucket_keys_are_told_apart_by_bytes(ctx):
    """names that differ only in bytes (same length, same device) and tokens that are prefixes of each
    other must stay separate keys"""
    ctx.state_clear()
    recs = []
    for i in range(200):
        recs.append(S.pb_measurements_record("dev", [("ab", 1.0 + i)], date=i))
        recs.append(S.pb_measurements_record("dev", [("ba", 2.0 + i)], date=i))
        recs.append(S.pb_measurements_record("deva", [("b", 3.0 + i)], date=i))
        recs.append(S.pb_measurements_record("de", [("vab", 4.0 + i)], date=i))
        recs.append(S.pb_alert_record("dev", "ab", "m", date=i))
        recs.append(S.pb_location_record("dev", 1.0, 2.0, date=i))
    data, rec_off = S.pack_records(recs)
    res = ctx.decode_pb(data, rec_off)
    ids = np.arange(res.n_rows, dtype=np.uint64)
    ctx.merge(data, rec_off, res.cols, 0, ids, ids)
    st = oracle.State()
    oc, _, _, t = oracle.decode_pb(data, rec_off)
    st.merge(data, rec_off, oc, 0, ids, ids)
    s = ctx.state_snapshot()
    assert s["seq"].shape[0] == 6
    snap_equal(s, st.snapshot())
    ctx.state_clear()


def test_fused_decode_merge_matches_decode_then_merge(ctx):
    """swgpu_decode_merge (one host-buffer call, rows never leave the GPU) leaves the same state as the oracle's
    decode + sequential fold: several batches, failed records, a fan-out record (ids are per RECORD: every row of
    a record carries its id), both formats"""
    from tests import pb_cases, json_cases
    ctx.state_clear()
    st = oracle.State()
    logic = sw.DeviceStateProcessingLogic(ctx)
    seq = 0
    rng = np.random.default_rng(9)
    for b in range(3):
        good, goff, _ = S.pb_mixed(20_000, 700, seed=500 + b)
        raw = good.tobytes()
        recs = [raw[int(goff[i]): int(goff[i + 1])] for i in range(20_000)]
        for c in pb_cases.cases():
            recs.insert(int(rng.integers(0, len(recs))), c[1])
        recs.insert(100, S.pb_measurements_record("fan%d" % b, [("m%03d" % k, float(k + b)) for k in range(300)], date=7 + b))
        data, rec_off = S.pack_records(recs)
        n = len(recs)
        id_hi = np.full(n, 40 + b, np.uint64)
        id_lo = rng.integers(0, 1 << 62, size=n).astype(np.uint64)
        status = np.zeros(n, np.uint8)
        rows = ctx.decode_merge("pb", data, rec_off, seq, id_hi, id_lo, status)
        ocols, ostatus, ofirst, total = oracle.decode_pb(data, rec_off)
        assert rows == total and np.array_equal(status, ostatus[:n])
        rec_of_row = (ocols.meta[:total].view(np.uint64) >> np.uint64(32)).astype(np.int64)
        st.merge(data, rec_off, ocols, seq, id_hi[rec_of_row], id_lo[rec_of_row])
        seq += total
    jrecs = [c[1] for c in json_cases.cases()]
    for i in range(3000):
        jrecs.append(S.json_measurement(S.token_str(i % 97), "fuel.level", repr(float(i) / 7), S.EVENT_DATE_BASE + i))
    data, rec_off = S.pack_records(jrecs)
    n = len(jrecs)
    id_hi = np.full(n, 99, np.uint64)
    id_lo = np.arange(n, dtype=np.uint64)
    logic.seq = seq
    rows = logic.process_records("json", data, rec_off, id_hi, id_lo)
    ocols, ostatus, ofirst, total = oracle.decode_json(data, rec_off)
    assert rows == total
    rec_of_row = (ocols.meta[:total].view(np.uint64) >> np.uint64(32)).astype(np.int64)
    st.merge(data, rec_off, ocols, seq, id_hi[rec_of_row], id_lo[rec_of_row])
    snap_equal(ctx.state_snapshot(), st.snapshot())
    assert ctx.decode_merge("pb", np.zeros(0, np.uint8), np.zeros(1, np.uint64), 0, np.zeros(0, np.uint64), np.zeros(0, np.uint64)) == 0
    ctx.state_clear()
