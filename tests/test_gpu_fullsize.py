"""BASELINE.json's configurations at (or as near as a test may run to) their full sizes, checked through
properties that do not need the oracle to process every event:

  C2  10 M protobuf DeviceLocation events + 1,024 concave 64-vertex zones   (full size)
  C3  mixed measurement / location / alert events, 1 M device tokens        (10 M events here; the config's 100 M
      is ten such batches — the merge is associative in the batch boundaries, which is one of the properties)
  C5  65,536 zones, 64 KB micro-batches                                      (full zone count; 64 micro-batches)

Each case also checks a bounded sample against the oracle bit for bit."""
import numpy as np
import pytest
import torch

import oracle
import sitewhere_b200 as sw
from sitewhere_b200 import synth as S

pytestmark = pytest.mark.gpu


def _to_dev(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def test_c2_full_size_decode_and_zone_test_properties(ctx):
    n = 10_000_000
    dev = torch.device("cuda", 0)
    zones = S.star_zones(32)
    lon, lat = S.locations_for_zones(n, *zones[:3], seed=S.SEED_BASE + 77, boundary_rate=1e-4)
    data, rec_off, truth = S.pb_locations(n, 1_000_000, lon, lat, seed=S.SEED_BASE + 78)
    ctx.set_zones(*zones)
    st = torch.cuda.ExternalStream(ctx.stream)
    with torch.cuda.stream(st):
        d_bytes = torch.empty(data.shape[0] + 64, dtype=torch.uint8, device=dev)
        d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
        d_off = _to_dev(rec_off.view(np.int64), dev)
        cols = sw.DeviceCols(n, dev, optional=False)
        d_status = torch.empty(n, dtype=torch.uint8, device=dev)
        d_first = torch.empty(n + 1, dtype=torch.int32, device=dev)
        d_nrows = torch.zeros(1, dtype=torch.int32, device=dev)
        d_pairs = torch.empty(n, dtype=torch.int64, device=dev)
        d_np = torch.zeros(1, dtype=torch.int64, device=dev)
        ctx.decode_pb_dev(d_bytes, d_off, n, cols, d_status, d_first, d_nrows)
        ctx.zone_test_dev(cols.v1, cols.v0, cols.meta, n, d_nrows, d_pairs, d_np)
    ctx.sync()
    torch.cuda.synchronize()
    # ---- decode: one row per record, in order, every field bit-identical to what the generator encoded
    assert int(d_nrows.item()) == n
    assert int(d_status.max().item()) == 0
    assert torch.equal(d_first, torch.arange(n + 1, dtype=torch.int32, device=dev))
    assert np.array_equal(cols.v0.cpu().numpy().view(np.uint64), truth["lat"].view(np.uint64))
    assert np.array_equal(cols.v1.cpu().numpy().view(np.uint64), truth["lon"].view(np.uint64))
    assert np.array_equal(cols.v2.cpu().numpy().view(np.uint64), truth["elev"].view(np.uint64))
    assert np.array_equal(cols.event_date.cpu().numpy(), truth["date"])
    meta = cols.meta.cpu().numpy().view(np.uint64)
    assert np.array_equal(meta >> np.uint64(32), np.arange(n, dtype=np.uint64))          # source record
    assert np.all((meta & np.uint64(0xff)) == 1)                                         # kind = location
    th = cols.token_hash.cpu().numpy().view(np.uint64)
    sample = np.random.default_rng(5).choice(n, 100_000, replace=False)
    want_m2 = S.murmur2_rows(S.token_bytes(truth["dev"][sample]))
    assert np.array_equal((th[sample] & np.uint64(0xffffffff)).astype(np.uint32), want_m2.astype(np.uint32))
    # ---- zone test: pairs ordered by (row, test); the same pairs whichever way the batch is cut
    n_pairs = int(d_np.item())
    pairs = d_pairs[:n_pairs].cpu().numpy().view(np.uint64)
    assert n_pairs > 2_000_000
    assert np.all(pairs[1:] > pairs[:-1])
    parts = []
    with torch.cuda.stream(st):
        for k in range(5):
            a, b = k * 2_000_000, (k + 1) * 2_000_000
            pk = torch.empty(b - a, dtype=torch.int64, device=dev)
            npk = torch.zeros(1, dtype=torch.int64, device=dev)
            ctx.zone_test_dev(cols.v1[a:b], cols.v0[a:b], cols.meta[a:b], b - a, None, pk, npk)
            ctx.sync()
            part = pk[: int(npk.item())].cpu().numpy().view(np.uint64)
            parts.append(part + (np.uint64(a) << np.uint64(32)))
    assert np.array_equal(np.concatenate(parts), pairs)
    # ---- a sample of rows against the oracle's index-free JTS loop
    rows = np.sort(sample[:50_000])
    want, total = oracle.zone_test(*zones, lon[rows], lat[rows])
    want = (rows[(want >> np.uint64(32)).astype(np.int64)].astype(np.uint64) << np.uint64(32)) | (want & np.uint64(0xffffffff))
    mask = np.isin((pairs >> np.uint64(32)).astype(np.int64), rows)
    assert np.array_equal(pairs[mask], want)


def _snap_equal(a, b):
    for k in ("seq", "id_hi", "id_lo", "event_date", "key_off", "key_bytes"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["value"].view(np.uint64), b["value"].view(np.uint64))


def _canon(snap):
    """snapshot rows ordered by their (unique) sequence number; keys reduced to length + weighted byte sum"""
    order = np.argsort(snap["seq"], kind="stable")
    ko = snap["key_off"].astype(np.int64)
    kb = snap["key_bytes"].astype(np.uint64)
    w = kb * ((np.arange(kb.shape[0], dtype=np.uint64) - np.repeat(ko[:-1], np.diff(ko)).astype(np.uint64)) % np.uint64(251) + np.uint64(1))
    out = {k: snap[k][order] for k in ("seq", "id_hi", "id_lo", "event_date")}
    out["value"] = snap["value"].view(np.uint64)[order]
    out["key_len"] = np.diff(ko)[order]
    out["key_sum"] = np.add.reduceat(w, ko[:-1])[order] if kb.shape[0] else np.zeros(0, np.uint64)
    return out


def _canon_equal(a, b):
    for k in a:
        assert np.array_equal(a[k], b[k]), k


def test_c3_mixed_events_merge_properties(ctx):
    """10 M mixed events over 1 M device tokens: the state after the fold has one entry per distinct key, holds
    the newest event of each, and does not depend on how the stream is cut into batches or on replays"""
    n = 10_000_000
    data, rec_off, _ = S.pb_mixed(n, 1_000_000, seed=S.SEED_BASE + 91)
    dev = torch.device("cuda", 0)
    st = torch.cuda.ExternalStream(ctx.stream)
    with torch.cuda.stream(st):
        d_bytes = torch.empty(data.shape[0] + 64, dtype=torch.uint8, device=dev)
        d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
        d_off = _to_dev(rec_off.view(np.int64), dev)
        cols = sw.DeviceCols(n, dev, optional=False)
        d_status = torch.empty(n, dtype=torch.uint8, device=dev)
        d_first = torch.empty(n + 1, dtype=torch.int32, device=dev)
        d_nrows = torch.zeros(1, dtype=torch.int32, device=dev)
        ctx.decode_pb_dev(d_bytes, d_off, n, cols, d_status, d_first, d_nrows)
        ctx.sync()
        assert int(d_nrows.item()) == n and int(d_status.max().item()) == 0
        ids = torch.arange(n, dtype=torch.int64, device=dev)
        ctx.state_clear()
        ctx.merge_dev(d_bytes, d_off, n, cols, n, d_nrows, 0, ids, ids)
        ctx.sync()
    whole = _canon(ctx.state_snapshot(sort=False))
    # one entry per distinct (token, kind, name): count the distinct keys from the decoded columns
    # (the two 64-bit hashes folded into one word: a property check, not a proof)
    th = cols.token_hash.cpu().numpy().view(np.uint64)
    nh = cols.name_hash.cpu().numpy().view(np.uint64)
    kind = (cols.meta.cpu().numpy().view(np.uint64) & np.uint64(0xff))
    with np.errstate(over="ignore"):
        key = th ^ (np.where(kind == 1, np.uint64(0), nh) * np.uint64(0x9E3779B97F4A7C15)) ^ (kind << np.uint64(61))
    uniq, last = np.unique(key[::-1], return_index=True)              # first hit in reversed order = newest row
    assert whole["seq"].shape[0] == uniq.shape[0] == ctx.state_size()[0]
    # the stored winners are exactly the newest row of each key
    assert np.array_equal(whole["seq"], np.sort((n - 1 - last).astype(np.uint64)))
    assert np.array_equal(whole["id_lo"], whole["seq"])               # ids were the row numbers
    # replay of the same batch: nothing changes
    with torch.cuda.stream(st):
        ctx.merge_dev(d_bytes, d_off, n, cols, n, d_nrows, 0, ids, ids)
        ctx.sync()
    _canon_equal(_canon(ctx.state_snapshot(sort=False)), whole)
    # the same stream cut into three uneven batches (rows are sliced, the payload buffer is shared)
    with torch.cuda.stream(st):
        ctx.state_clear()
        for a, b in ((0, 1_234_567), (1_234_567, 7_000_001), (7_000_001, n)):
            sub = sw.DeviceCols.__new__(sw.DeviceCols)
            sub.__dict__.update({k: (v[a:b] if isinstance(v, torch.Tensor) else v) for k, v in cols.__dict__.items()})
            ctx.merge_dev(d_bytes, d_off, n, sub, b - a, None, a, ids[a:b], ids[a:b])
        ctx.sync()
    _canon_equal(_canon(ctx.state_snapshot(sort=False)), whole)
    # a bounded prefix against the oracle's sequential fold, bit for bit
    m = 300_000
    ctx.state_clear()
    res = ctx.decode_pb(data[: int(rec_off[m])], rec_off[: m + 1])
    ocols, _, _, total = oracle.decode_pb(data[: int(rec_off[m])], rec_off[: m + 1])
    idm = np.arange(total, dtype=np.uint64)
    ctx.merge(data[: int(rec_off[m])], rec_off[: m + 1], res.cols, 0, idm, idm)
    ost = oracle.State()
    ost.merge(data[: int(rec_off[m])], rec_off[: m + 1], ocols, 0, idm, idm)
    _snap_equal(ctx.state_snapshot(), ost.snapshot())
    ctx.state_clear()


def test_c5_many_zones_micro_batches(ctx):
    """65,536 zones (256 x 256 concave 64-vertex polygons), Kafka-shaped 64 KB micro-batches through the
    host-buffer call: every batch gives what one big batch gives, and sampled batches match the oracle"""
    zones = S.star_zones(256, nvert=64, seed=S.SEED_BASE + 5)
    assert zones[3].shape[0] == 65_536
    ctx.set_zones(*zones)
    per = 65_536 // 63                       # DeviceLocation payloads in 64 KB
    nb = 64
    n = per * nb
    lon, lat = S.locations_for_zones(n, *zones[:3], seed=S.SEED_BASE + 6, boundary_rate=1e-3)
    data, rec_off, _ = S.pb_locations(n, 10_000_000, lon, lat, seed=S.SEED_BASE + 7)
    whole = ctx.decode_zone_test("pb", data, rec_off)
    w_pairs = whole[4]
    assert whole[3] == n and w_pairs.shape[0] > n // 8
    got = []
    for k in range(nb):
        a, b = k * per, (k + 1) * per
        r = ctx.decode_zone_test("pb", data[a * 63: b * 63], rec_off[a: b + 1] - rec_off[a])
        assert r[3] == per
        got.append(r[4] + (np.uint64(a) << np.uint64(32)))
    assert np.array_equal(np.concatenate(got), w_pairs)
    for k in (0, 37):                        # index-free JTS loop over all 65,536 zones: two batches
        a, b = k * per, (k + 1) * per
        want, _ = oracle.zone_test(*zones, lon[a:b], lat[a:b])
        assert np.array_equal(got[k] - (np.uint64(a) << np.uint64(32)), want)
