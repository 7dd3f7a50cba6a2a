"""GPU parity: the fused host-buffer call swgpu_decode_zone_test (chunked, overlapped copies) must give
exactly what the oracle's decode followed by its index-free zone loop gives — across chunk
boundaries, with failed records, measurement fan-out beyond a chunk's row capacity, and more alerts
per chunk than its pair buffer holds."""
import numpy as np
import pytest

import oracle
from sitewhere_b200 import synth as S
from sitewhere_b200.api import pinned_array, pinned_cols
from tests import pb_cases, json_cases
from tests.helpers import assert_cols_equal

pytestmark = pytest.mark.gpu


def check(ctx, fmt, data, rec_off, zones, chunk):
    ctx.set_zones(*zones)
    cols, status, first_row, n_rows, pairs = ctx.decode_zone_test(fmt, data, rec_off, chunk_records=chunk)
    ocols, ostatus, ofirst, ototal = getattr(oracle, "decode_pb" if fmt == "pb" else "decode_json")(data, rec_off)
    assert n_rows == ototal
    assert np.array_equal(status, ostatus)
    assert np.array_equal(first_row, ofirst)
    assert_cols_equal(cols, ocols, ototal)
    want, total = oracle.zone_test(*zones, ocols.v1[:ototal], ocols.v0[:ototal], meta=ocols.meta[:ototal])
    assert pairs.shape[0] == total
    assert np.array_equal(pairs, want)
    return n_rows, total


def test_pb_locations_many_chunks(ctx):
    zones = S.star_zones(16)
    lon, lat = S.locations_for_zones(50_000, *zones[:3], seed=11, boundary_rate=0.01)
    data, rec_off, _ = S.pb_locations(50_000, 3000, lon, lat, seed=11)
    rows, alerts = check(ctx, "pb", data, rec_off, zones, chunk=4096)      # 13 chunks, last one ragged
    assert rows == 50_000 and alerts > 1000
    check(ctx, "pb", data, rec_off, zones, chunk=0)                        # default chunk: single chunk


def test_pb_mixed_with_failures_and_fanout(ctx):
    zones = S.star_zones(8, cond_mode="alternate")      # 32 Outside-condition tests: ~32 alerts per location
    good, goff, _ = S.pb_mixed(6000, 200, seed=12, zones=zones)
    raw = good.tobytes()
    recs = [raw[int(goff[i]): int(goff[i + 1])] for i in range(6000)]
    cs = pb_cases.cases()
    rng = np.random.default_rng(2)
    for k in range(400):
        recs.insert(int(rng.integers(0, len(recs))), cs[k % len(cs)][1])
    # one payload with 900 measurements: more rows than a 512-record chunk has capacity for
    recs.insert(1000, S.pb_measurements_record("fan", [("m%03d" % k, float(k)) for k in range(900)], date=5))
    data, rec_off = S.pack_records(recs)
    check(ctx, "pb", data, rec_off, zones, chunk=512)


def test_json_pipeline(ctx):
    zones = S.star_zones(8)
    recs = []
    rng = np.random.default_rng(3)
    lon, lat = S.locations_for_zones(4000, *zones[:3], seed=13, boundary_rate=0.02)
    for i in range(4000):
        recs.append(S.json_location(S.token_str(int(rng.integers(0, 500))), repr(float(lat[i])), repr(float(lon[i])), "0.0",
                                    S.EVENT_DATE_BASE + i))
        if i % 7 == 0:
            recs.append(S.json_measurement("tok", "fuel.level", "1.5", S.EVENT_DATE_BASE + i))
    for c in json_cases.cases():
        recs.insert(int(rng.integers(0, len(recs))), c[1])
    data, rec_off = S.pack_records(recs)
    check(ctx, "json", data, rec_off, zones, chunk=700)


def test_pinned_buffers_and_column_subset(ctx):
    zones = S.star_zones(8)
    lon, lat = S.locations_for_zones(20_000, *zones[:3], seed=14)
    data, rec_off, _ = S.pb_locations(20_000, 1000, lon, lat, seed=14)
    ctx.set_zones(*zones)
    p_data = pinned_array((data.shape[0],), np.uint8)
    p_data[:] = data
    p_off = pinned_array((rec_off.shape[0],), np.uint64)
    p_off[:] = rec_off
    names = ["meta", "v0", "v1", "token_span"]
    p_cols = pinned_cols(20_000, names)
    p_pairs = pinned_array((20_000,), np.uint64)
    for _ in range(2):       # twice: the ring and its events are reused
        cols, status, first_row, n_rows, pairs = ctx.decode_zone_test("pb", p_data, p_off, p_cols, None, None, p_pairs,
                                                                      chunk_records=3000)
        ocols, ostatus, ofirst, ototal = oracle.decode_pb(data, rec_off)
        assert n_rows == ototal
        for nm in names:
            assert np.array_equal(getattr(cols, nm)[:ototal].view(np.uint64), getattr(ocols, nm)[:ototal].view(np.uint64)), nm
        want, total = oracle.zone_test(*zones, ocols.v1[:ototal], ocols.v0[:ototal], meta=ocols.meta[:ototal])
        assert np.array_equal(pairs, want)


def test_empty_batch(ctx):
    ctx.set_zones(*S.rect_zones(2))
    cols, status, first_row, n_rows, pairs = ctx.decode_zone_test("pb", np.zeros(0, np.uint8), np.zeros(1, np.uint64))
    assert n_rows == 0 and pairs.shape[0] == 0 and first_row.tolist() == [0]


def test_small_batch_path(ctx):
    """batches of up to 16,384 records with the chunk size left to the library go through one staging block each
    way: same results as the ring, for both formats, with failed records, a subset of columns, a batch whose
    fan-out does not fit the block (falls back to the ring) and one whose alerts exceed the caller's buffer"""
    zones = S.star_zones(8, cond_mode="alternate")
    good, goff, _ = S.pb_mixed(3000, 200, seed=31, zones=zones)
    raw = good.tobytes()
    recs = [raw[int(goff[i]): int(goff[i + 1])] for i in range(3000)]
    rng = np.random.default_rng(7)
    for k, c in enumerate(pb_cases.cases()):
        recs.insert(int(rng.integers(0, len(recs))), c[1])
    data, rec_off = S.pack_records(recs)
    check(ctx, "pb", data, rec_off, zones, chunk=0)                     # small block (alerts retried with a larger buffer)
    check(ctx, "pb", data, rec_off, zones, chunk=1000)                  # ring, for comparison
    recs.insert(500, S.pb_measurements_record("fan", [("m%03d" % k, float(k)) for k in range(5000)], date=5))
    data, rec_off = S.pack_records(recs)
    check(ctx, "pb", data, rec_off, zones, chunk=0)                     # 5,000 rows from one record: does not fit the block
    jrecs = [c[1] for c in json_cases.cases()]
    lon, lat = S.locations_for_zones(800, *zones[:3], seed=15, boundary_rate=0.05)
    for i in range(800):
        jrecs.insert(int(rng.integers(0, len(jrecs))),
                     S.json_location(S.token_str(i % 50), repr(float(lat[i])), repr(float(lon[i])), "1.0", S.EVENT_DATE_BASE + i))
    data, rec_off = S.pack_records(jrecs)
    check(ctx, "json", data, rec_off, zones, chunk=0)
    # column subset + no status / first_row
    ctx.set_zones(*zones)
    names = ["meta", "v1"]
    p_cols = pinned_cols(len(jrecs), names)
    cols, status, first_row, n_rows, pairs = ctx.decode_zone_test("json", data, rec_off, p_cols, None, None, None)
    ocols, _, _, ototal = oracle.decode_json(data, rec_off)
    assert n_rows == ototal
    for nm in names:
        assert np.array_equal(getattr(cols, nm)[:ototal].view(np.uint64), getattr(ocols, nm)[:ototal].view(np.uint64)), nm


def test_reserve_batch_then_async_submit_and_wait():
    """swgpu_reserve_batch sizes the buffers up front; swgpu_submit_decode_zone_test / swgpu_wait run the same
    call on the context's worker thread: three batches queued at once, waited for in order, each equal to the oracle"""
    import ctypes as C
    import sitewhere_b200 as sw
    from sitewhere_b200 import _lib as L
    ctx = sw.Context(0)
    try:
        zones = S.star_zones(8)
        ctx.set_zones(*zones)
        assert ctx.lib.swgpu_reserve_batch(ctx.h, 300_000, 300_000 * 64) == 0
        jobs = []
        for k, n in enumerate((40_000, 300_000, 2_000)):                 # ring path, ring path, small-batch path
            lon, lat = S.locations_for_zones(n, *zones[:3], seed=40 + k, boundary_rate=0.01)
            data, rec_off, _ = S.pb_locations(n, 3000, lon, lat, seed=40 + k)
            p_data = pinned_array((data.shape[0],), np.uint8)
            p_data[:] = data
            p_off = pinned_array((n + 1,), np.uint64)
            p_off[:] = rec_off
            cols = pinned_cols(n)
            status = pinned_array((n,), np.uint8)
            first = pinned_array((n + 1,), np.uint32)
            pairs = pinned_array((n,), np.uint64)
            out_rows, n_pairs, ticket = C.c_uint32(), C.c_uint64(), C.c_uint64()
            s = cols.struct()
            rc = ctx.lib.swgpu_submit_decode_zone_test(ctx.h, 0, p_data.ctypes.data, p_off.ctypes.data, n, n, C.byref(s),
                                                       status.ctypes.data, first.ctypes.data, C.byref(out_rows), n,
                                                       pairs.ctypes.data, C.byref(n_pairs), 0, C.byref(ticket))
            assert rc == 0 and ticket.value == k + 1
            jobs.append((ticket.value, data, rec_off, cols, status, first, pairs, out_rows, n_pairs, p_data, p_off, s))
        for (t, data, rec_off, cols, status, first, pairs, out_rows, n_pairs, *_keep) in jobs:
            assert ctx.lib.swgpu_wait(ctx.h, t) == 0
            ocols, ostatus, ofirst, ototal = oracle.decode_pb(data, rec_off)
            assert out_rows.value == ototal and np.array_equal(status, ostatus) and np.array_equal(first, ofirst)
            assert_cols_equal(cols, ocols, ototal)
            want, total = oracle.zone_test(*zones, ocols.v1[:ototal], ocols.v0[:ototal], meta=ocols.meta[:ototal])
            assert n_pairs.value == total and np.array_equal(pairs[:total], want)
        assert ctx.lib.swgpu_wait(ctx.h, 99) == L.E_INVALID
    finally:
        ctx.close()


def _dev_decode_zone(ctx, fmt, data, rec_off, cap_rows=None, cap_pairs=1 << 20):
    """swgpu_decode_zone_test_dev on device buffers -> (per-column dict, status, first_row, n_rows, pairs)"""
    import torch
    import sitewhere_b200 as sw
    from sitewhere_b200 import _lib as L
    dev = torch.device("cuda", 0)
    n = rec_off.shape[0] - 1
    cap = cap_rows or max(n, 1)
    st = torch.cuda.current_stream(dev)
    c2 = sw.Context(0, stream=st.cuda_stream)
    try:
        d_bytes = torch.zeros(data.shape[0] + 64, dtype=torch.uint8, device=dev)
        d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
        d_off = torch.from_numpy(rec_off.view(np.int64)).to(dev)
        cols = sw.DeviceCols(cap, dev, optional=True)
        d_status = torch.zeros(max(n, 1), dtype=torch.uint8, device=dev)
        d_first = torch.zeros(n + 1, dtype=torch.int32, device=dev)
        d_rows = torch.zeros(1, dtype=torch.int32, device=dev)
        d_pairs = torch.zeros(cap_pairs, dtype=torch.int64, device=dev)
        d_np = torch.zeros(1, dtype=torch.int64, device=dev)
        return c2, d_bytes, d_off, n, cols, d_status, d_first, d_rows, d_pairs, d_np
    except Exception:
        c2.close()
        raise


@pytest.mark.parametrize("shape", ["dense_locations", "mixed_with_failures", "outside_conditions", "json"])
def test_decode_zone_test_dev_equals_oracle(ctx, shape):
    """the device-buffer form against the oracle: a dense batch, a batch with failed records and measurement
    fan-out (the ordered pass takes over), Outside conditions, and JSON"""
    import torch
    fmt = "json" if shape == "json" else "pb"
    zones = S.star_zones(8, cond_mode="alternate") if shape == "outside_conditions" else S.star_zones(16)
    if shape == "dense_locations" or shape == "outside_conditions":
        lon, lat = S.locations_for_zones(30_001, *zones[:3], seed=21, boundary_rate=0.02)
        data, rec_off, _ = S.pb_locations(30_001, 2000, lon, lat, seed=21)
    elif shape == "mixed_with_failures":
        good, goff, _ = S.pb_mixed(9000, 300, seed=22, zones=zones)
        raw = good.tobytes()
        recs = [raw[int(goff[i]): int(goff[i + 1])] for i in range(9000)]
        cs = pb_cases.cases()
        rng = np.random.default_rng(5)
        for k in range(300):
            recs.insert(int(rng.integers(0, len(recs))), cs[k % len(cs)][1])
        recs.insert(500, S.pb_measurements_record("fan", [("m%03d" % k, float(k)) for k in range(40)], date=5))
        data, rec_off = S.pack_records(recs)
    else:
        lon, lat = S.locations_for_zones(5000, *zones[:3], seed=23, boundary_rate=0.02)
        recs = [S.json_location(S.token_str(i % 100), repr(float(lat[i])), repr(float(lon[i])), "1.5", S.EVENT_DATE_BASE + i)
                for i in range(5000)]
        data, rec_off = S.pack_records(recs)
    ocols, ostatus, ofirst, ototal = getattr(oracle, "decode_pb" if fmt == "pb" else "decode_json")(data, rec_off)
    want, total = oracle.zone_test(*zones, ocols.v1[:ototal], ocols.v0[:ototal], meta=ocols.meta[:ototal])
    n = rec_off.shape[0] - 1
    c2, d_bytes, d_off, n, cols, d_status, d_first, d_rows, d_pairs, d_np = _dev_decode_zone(
        ctx, fmt, data, rec_off, cap_rows=max(ototal, n) + 7, cap_pairs=max(total, 1) + 5)
    try:
        c2.set_zones(*zones)
        for _ in range(2):   # twice: the work area and flags are reused
            c2.decode_zone_test_dev(fmt, d_bytes, d_off, n, cols, d_status, d_first, d_rows, d_pairs, d_np)
        torch.cuda.synchronize()
        assert int(d_rows.item()) == ototal
        assert np.array_equal(d_status.cpu().numpy()[:n], ostatus)
        assert np.array_equal(d_first.cpu().numpy().view(np.uint32), ofirst)
        for name in ("meta", "token_hash", "name_hash", "event_date", "token_span", "name_span"):
            assert np.array_equal(getattr(cols, name).cpu().numpy()[:ototal].view(np.uint64),
                                  np.asarray(getattr(ocols, name)[:ototal]).view(np.uint64)), name
        for name in ("v0", "v1", "v2"):
            assert np.array_equal(getattr(cols, name).cpu().numpy()[:ototal].view(np.uint64),
                                  np.asarray(getattr(ocols, name)[:ototal]).view(np.uint64)), name
        assert int(d_np.item()) == total
        assert np.array_equal(d_pairs.cpu().numpy()[:total].view(np.uint64), want)
    finally:
        c2.close()
