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
        st.merge_by_record(data, rec_off, ocols, total, seq, id_hi, id_lo)
        seq += n << 12                      # SWGPU_SEQ_STRIDE(n)
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
    st.merge_by_record(data, rec_off, ocols, total, seq, id_hi, id_lo)
    snap_equal(ctx.state_snapshot(), st.snapshot())
    assert ctx.decode_merge("pb", np.zeros(0, np.uint8), np.zeros(1, np.uint64), 0, np.zeros(0, np.uint64), np.zeros(0, np.uint64)) == 0
    ctx.state_clear()


def test_long_keys_inline_arena_and_growth(ctx):
    """keys of every size class: in the slot's first two sectors, in its fourth, and in the long-key arena
    (which starts small and must grow and replay); 200-byte tokens and names as in ADVICE r1"""
    ctx.state_clear()
    st = oracle.State()
    rng = np.random.default_rng(21)
    seq = 0
    for b in range(3):
        recs = []
        for i in range(3000):
            tl = int(rng.choice([1, 7, 8, 21, 36, 40, 64, 65, 200, 300]))
            nl = int(rng.choice([0, 1, 8, 12, 16, 33, 200]))
            tok = "t%d_" % (i % 400) + "x" * max(0, tl - 5)
            name = "n%d_" % (i % 7) + "y" * max(0, nl - 3)
            recs.append(S.pb_measurements_record(tok, [(name, float(i))], date=1000 * b + i))
        recs.append(S.pb_measurements_record("\x00" * 16, [("\x00" * 8, 1.5)], date=5))       # all-zero key words
        recs.append(S.pb_measurements_record("\x00" * 16, [("\x00" * 9, 2.5)], date=6))
        data, rec_off = S.pack_records(recs)
        n = len(recs)
        ids = np.arange(n, dtype=np.uint64) + np.uint64(b * 10_000)
        if b == 1:                       # unfused path on the same key population
            res = ctx.decode_pb(data, rec_off)
            ocols, _, _, total = oracle.decode_pb(data, rec_off)
            ctx.merge(data, rec_off, res.cols, seq, ids[:total], ids[:total])
            st.merge(data, rec_off, ocols, seq, ids[:total], ids[:total])
            seq += total
        else:
            rows = ctx.decode_merge("pb", data, rec_off, seq, ids, ids)
            ocols, _, _, total = oracle.decode_pb(data, rec_off)
            assert rows == total
            st.merge_by_record(data, rec_off, ocols, total, seq, ids, ids)
            seq += n << 12
    a, b_ = ctx.state_snapshot(), st.snapshot()
    snap_equal(a, b_)
    ctx.state_restore(a)
    snap_equal(ctx.state_snapshot(), b_)
    ctx.state_clear()


def test_null_request_rows_do_not_reach_state(ctx):
    """JSON "request": null decodes to an OK row flagged NULL_REQUEST; the reference drops it before the state"""
    ctx.state_clear()
    recs = [b'{"deviceToken":"d1","type":"DeviceMeasurement","request":null}',
            S.json_measurement("d1", "fuel.level", "1.5", S.EVENT_DATE_BASE),
            b'{"deviceToken":"d2","type":"DeviceLocation","request":null}']
    data, rec_off = S.pack_records(recs)
    ids = np.arange(3, dtype=np.uint64)
    rows = ctx.decode_merge("json", data, rec_off, 0, ids, ids)
    assert rows == 3
    snap = ctx.state_snapshot()
    assert snap["seq"].shape[0] == 1 and int(snap["id_lo"][0]) == 1
    st = oracle.State()
    ocols, _, _, total = oracle.decode_json(data, rec_off)
    st.merge_by_record(data, rec_off, ocols, total, 0, ids, ids)
    snap_equal(snap, st.snapshot())
    res = ctx.decode_json(data, rec_off)              # and through the unfused path
    ctx.state_clear()
    ctx.merge(data, rec_off, res.cols, 0, ids, ids)
    assert ctx.state_snapshot()["seq"].shape[0] == 1
    ctx.state_clear()


def test_merge_dev_reports_dropped_rows():
    """a device-resident merge cannot grow the long-key arena: the drop is reported by swgpu_state_size (once),
    and after a reserve the replayed batch lands"""
    import torch
    from sitewhere_b200._lib import SwgpuError
    ctx = sw.Context(0)          # a fresh context: its long-key arena has its initial, small size
    dev = torch.device("cuda", ctx.device)
    recs = [S.pb_measurements_record("L%04d" % i + "z" * 300, [("m", float(i))], date=i) for i in range(2000)]
    data, rec_off = S.pack_records(recs)
    n = len(recs)
    d_bytes = torch.zeros(data.shape[0] + 64, dtype=torch.uint8, device=dev)
    d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
    d_off = torch.from_numpy(rec_off.view(np.int64)).to(dev)
    ids = torch.arange(n, dtype=torch.int64, device=dev)
    d_status = torch.zeros(n, dtype=torch.uint8, device=dev)
    d_rows = torch.zeros(1, dtype=torch.int32, device=dev)
    ctx.decode_merge_dev("pb", d_bytes, d_off, n, 0, ids, ids, d_status, d_rows)
    ctx.sync()
    assert int(d_rows.item()) == n
    with pytest.raises(SwgpuError, match="replay"):
        ctx.state_size()
    ctx.state_reserve(4 * n, n * 400)
    ctx.decode_merge_dev("pb", d_bytes, d_off, n, 0, ids, ids, d_status, d_rows)
    ctx.sync()
    keys, _ = ctx.state_size()
    assert keys == n
    st = oracle.State()
    ocols, _, _, total = oracle.decode_pb(data, rec_off)
    idn = np.arange(n, dtype=np.uint64)
    st.merge_by_record(data, rec_off, ocols, total, 0, idn, idn)
    snap_equal(ctx.state_snapshot(), st.snapshot())
    assert ctx.state_size()[0] == n      # the drop was reported once; nothing is pending now
    ctx.close()


def test_presence_rows_and_scan(ctx):
    """lastInteractionDate / presenceMissingDate per device and the presence scan, against the oracle: three batches
    stamped at different times (unfused, fused protobuf, fused JSON), scans that report some devices, a device that
    comes back, a scan with too small a result buffer"""
    ctx.state_clear()
    st = oracle.State()
    t0 = 1_600_000_000_000
    seq = 0
    try:
        # batch 1 (t0): devices 0..299, unfused merge
        d1, o1, _ = S.pb_mixed(20_000, 300, seed=700)
        ctx.set_option("state.presence_now_ms", t0)
        st.set_presence_now(t0)
        res = ctx.decode_pb(d1, o1)
        oc, _, _, tot = oracle.decode_pb(d1, o1)
        ids = np.arange(tot, dtype=np.uint64)
        ctx.merge(d1, o1, res.cols, seq, ids, ids)
        st.merge(d1, o1, oc, seq, ids, ids)
        seq += tot
        # batch 2 (t0 + 60 s): devices 0..99 only, fused
        d2, o2, _ = S.pb_mixed(5_000, 100, seed=701)
        ctx.set_option("state.presence_now_ms", t0 + 60_000)
        st.set_presence_now(t0 + 60_000)
        n2 = len(o2) - 1
        ids2 = np.arange(n2, dtype=np.uint64)
        rows = ctx.decode_merge("pb", d2, o2, seq, ids2, ids2)
        oc2, _, _, tot2 = oracle.decode_pb(d2, o2)
        assert rows == tot2
        st.merge_by_record(d2, o2, oc2, tot2, seq, ids2, ids2)
        seq += n2 << 12
        snap_equal(ctx.state_snapshot(), st.snapshot())
        # scan at t0 + 90 s with a 45 s interval: the devices last seen at t0 are missing
        last, keys, found = ctx.presence_scan(t0 + 90_000, 45_000)
        olast, okeys = st.presence_scan(t0 + 90_000, 45_000)
        assert found == len(okeys) and 150 < found < 300
        assert np.array_equal(last.astype(np.uint64), olast) and keys == okeys
        assert all(k[2] == 0x7F for k in keys)
        # the same scan again reports nothing (SendOnce); the snapshot shows the presenceMissingDate
        last, keys, found = ctx.presence_scan(t0 + 91_000, 45_000)
        assert found == 0 and st.presence_scan(t0 + 91_000, 45_000)[1] == []
        snap_equal(ctx.state_snapshot(), st.snapshot())
        # batch 3 (t0 + 120 s), JSON: some missing devices come back -> presenceMissingDate cleared
        jrecs = [S.json_measurement(S.token_str(d), "fuel.level", "1.5", S.EVENT_DATE_BASE + d) for d in range(50, 250, 3)]
        dj, oj = S.pack_records(jrecs)
        ctx.set_option("state.presence_now_ms", t0 + 120_000)
        st.set_presence_now(t0 + 120_000)
        idj = np.arange(len(jrecs), dtype=np.uint64)
        rows = ctx.decode_merge("json", dj, oj, seq, idj, idj)
        ocj, _, _, totj = oracle.decode_json(dj, oj)
        st.merge_by_record(dj, oj, ocj, totj, seq, idj, idj)
        seq += len(jrecs) << 12
        snap_equal(ctx.state_snapshot(), st.snapshot())
        # a later scan with room for 10 devices only: 10 reported and marked, E_CAPACITY semantics, the rest next time
        last, keys, found = ctx.presence_scan(t0 + 400_000, 100_000, max_devices=10)
        assert len(keys) == 10 and found > 10
        olast, okeys = st.presence_scan(t0 + 400_000, 100_000)
        assert found == len(okeys)
        last2, keys2, found2 = ctx.presence_scan(t0 + 400_000, 100_000)
        assert found2 == found - 10 and sorted(keys + keys2) == sorted(okeys)
        snap_equal(ctx.state_snapshot(), st.snapshot())
    finally:
        ctx.set_option("state.presence_now_ms", 0)
        ctx.state_clear()
