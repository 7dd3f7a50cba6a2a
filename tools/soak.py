"""Differential soak: the fuzzers of tests/test_gpu_fuzz.py with fresh seeds for a given number of seconds.
    python tools/soak.py 240        (prints one line per round, stops at the first mismatch)"""
import sys, time
sys.path.insert(0, '.')
import numpy as np
import oracle
import sitewhere_b200 as sw
from sitewhere_b200 import synth as S
from tests import test_gpu_fuzz as F, pb_cases, json_cases
from tests.helpers import check_decode_parity

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
ctx = sw.Context()
t_end = time.time() + budget
seed = int(time.time()) & 0xffffff
rnd = 0
d0, o0, _ = S.pb_mixed(3000, 200, seed=77)
pb_base = F._records(d0, o0) + [bytes(c[1]) for c in pb_cases.cases()]
d1, o1, _ = S.json_measurements(1500, 100, seed=78)
js_base = F._records(d1, o1) + [bytes(c[1]) for c in json_cases.cases()]
js_base += [S.json_location("00001-SYNTH-000000001", "34.5", "-84.25", "12.5", 1540000000000),
            S.json_alert("00001-SYNTH-000000002", "engine.overheat", "too hot", 1540000000001)]
tot = dict(pb=0, json=0, zone=0, pipe=0)
while time.time() < t_end:
    rng = np.random.default_rng(seed + rnd)
    recs = F._mutate(rng, pb_base, F.PB_SPECIALS, 50_000, max_mut=1 + rnd % 4)
    data, off = S.pack_records(recs)
    check_decode_parity(ctx.decode_pb(data, off), data, off, "decode_pb"); tot["pb"] += len(recs)
    recs = F._mutate(rng, js_base, F.JSON_SPECIALS, 30_000, max_mut=1 + rnd % 3)
    data, off = S.pack_records(recs)
    check_decode_parity(ctx.decode_json(data, off), data, off, "decode_json"); tot["json"] += len(recs)
    # lattice polygons
    T = 120
    vo, zx, zy = [0], [], []
    for t in range(T):
        nv = int(rng.integers(3, 16)); cx, cy = rng.integers(-20, 20, size=2); span = int(rng.integers(1, 9))
        sc = (1.0, 0.5, 0.125, 3.0)[rnd % 4]
        zx += list((cx + rng.integers(-span, span + 1, size=nv)).astype(np.float64) * sc)
        zy += list((cy + rng.integers(-span, span + 1, size=nv)).astype(np.float64) * sc)
        vo.append(len(zx))
    zones = (np.array(vo, np.uint32), np.array(zx), np.array(zy), rng.integers(0, 2, size=T).astype(np.uint8))
    n = 8000
    lon = rng.integers(-60, 61, size=n).astype(np.float64) * sc * 0.5
    lat = rng.integers(-60, 61, size=n).astype(np.float64) * sc * 0.5
    lon[: n // 4] += rng.choice([1e-9, -1e-9, 2.0 ** -40, 0.25], size=n // 4)
    ctx.set_zones(*zones)
    got = ctx.zone_test(lon, lat, max_pairs=n * (T + 1))
    want, total = oracle.zone_test(*zones, lon, lat)
    assert got.shape[0] == total and np.array_equal(got, want), "zone mismatch round %d seed %d" % (rnd, seed)
    tot["zone"] += n
    # the fused pipeline on a mutated protobuf batch with random chunking
    recs = F._mutate(rng, pb_base, F.PB_SPECIALS, int(rng.choice([700, 5000, 16_384, 20_000])), max_mut=1)   # (small-batch path too)
    data, off = S.pack_records(recs)
    chunk = int(rng.choice([0, 0, 333, 1000, 4096, 7777]))
    cols, status, first_row, n_rows, pairs = ctx.decode_zone_test("pb", data, off, chunk_records=chunk)
    ocols, ostatus, ofirst, ototal = oracle.decode_pb(data, off)
    assert n_rows == ototal and np.array_equal(status, ostatus) and np.array_equal(first_row, ofirst)
    want, total = oracle.zone_test(*zones, ocols.v1[:ototal], ocols.v0[:ototal], meta=ocols.meta[:ototal])
    assert np.array_equal(pairs, want), "pipeline mismatch round %d seed %d chunk %d" % (rnd, seed, chunk)
    tot["pipe"] += len(recs)
    # the same through the small-batch path for JSON (ordered pass alone + single-launch zone kernel)
    recs = F._mutate(rng, js_base, F.JSON_SPECIALS, int(rng.choice([300, 3000, 12_000])), max_mut=1) + list(js_base[-2:]) * 50
    data, off = S.pack_records(recs)
    cols, status, first_row, n_rows, pairs = ctx.decode_zone_test("json", data, off)
    ocols, ostatus, ofirst, ototal = oracle.decode_json(data, off)
    assert n_rows == ototal and np.array_equal(status, ostatus) and np.array_equal(first_row, ofirst)
    want, total = oracle.zone_test(*zones, ocols.v1[:ototal], ocols.v0[:ototal], meta=ocols.meta[:ototal])
    assert np.array_equal(pairs, want), "small json pipeline mismatch round %d seed %d" % (rnd, seed)
    tot["pipe"] += len(recs)
    # device-state fold: a fresh state every fourth round, hot keys (few devices) or many keys, damaged records
    if rnd % 4 == 0:
        ctx.state_clear()
        ost = oracle.State()
        seq = 0
    ndev = int(rng.choice([3, 40, 5000]))
    d2, o2, _ = S.pb_mixed(15_000, ndev, seed=int(rng.integers(1 << 30)))
    recs = F._records(d2, o2)
    recs = recs + F._mutate(rng, recs[:2000], F.PB_SPECIALS, 2000, max_mut=1)
    order = rng.permutation(len(recs))
    recs = [recs[i] for i in order]
    data, off = S.pack_records(recs)
    ids_hi = rng.integers(0, 1 << 62, size=len(recs)).astype(np.uint64)
    ids_lo = rng.integers(0, 1 << 62, size=len(recs)).astype(np.uint64)
    rows = ctx.decode_merge("pb", data, off, seq, ids_hi, ids_lo)
    ocols, ostatus, ofirst, ototal = oracle.decode_pb(data, off)
    assert rows == ototal
    ost.merge_by_record(data, off, ocols, ototal, seq, ids_hi, ids_lo)
    seq += len(recs) << 12            # SWGPU_SEQ_STRIDE(n)
    a, b = ctx.state_snapshot(), ost.snapshot()
    for k in ("seq", "id_hi", "id_lo", "event_date", "key_off", "key_bytes"):
        assert np.array_equal(a[k], b[k]), "state mismatch %s round %d seed %d" % (k, rnd, seed)
    assert np.array_equal(a["value"].view(np.uint64), b["value"].view(np.uint64))
    tot["merge"] = tot.get("merge", 0) + ototal
    # enriched payloads: serialised (forward pass), concatenated (merge semantics) and damaged (general parser)
    import random as _random
    from tests import enriched_proto as EP
    from tests.test_gpu_enriched import _damage
    from sitewhere_b200 import _lib as L
    prnd = _random.Random(seed + rnd)
    ebase = [EP.random_payload(prnd, kinds=(1, 2, 3, 6)).SerializeToString() for _ in range(1500)]
    erecs = ebase + [prnd.choice(ebase) + prnd.choice(ebase) for _ in range(500)] + _damage(prnd, ebase[:300], 6000)
    prnd.shuffle(erecs)
    data, off = S.pack_records(erecs)
    res, ids = ctx.decode_enriched(data, off)
    ocols, oids, ostatus, ofirst, ototal = oracle.decode_enriched(data, off)
    assert res.n_rows == ototal and np.array_equal(res.status, ostatus) and np.array_equal(res.first_row, ofirst), "enriched %d %d" % (rnd, seed)
    from tests.helpers import assert_cols_equal
    assert_cols_equal(res.cols, ocols, ototal)
    for k in L.ID_NAMES:
        assert np.array_equal(ids[k], getattr(oids, k)[:ototal]), k
    tot["enriched"] = tot.get("enriched", 0) + len(erecs)
    # outbound encode of the rows of a mutated batch (device buffers)
    import torch
    for fmt, base, spec in (("pb", pb_base, F.PB_SPECIALS), ("json", js_base, F.JSON_SPECIALS)):
        recs = F._mutate(rng, base, spec, 8000, max_mut=1) + list(base[:500])
        data, off = S.pack_records(recs)
        ocols, ostatus, ofirst, ototal = getattr(oracle, "decode_" + fmt)(data, off)
        want, woff, wst = oracle.encode_inbound(fmt, data, off, ocols, ototal, b"soak", 1_700_000_000_000)
        dev = torch.device("cuda", 0)
        n = off.shape[0] - 1
        d_bytes = torch.zeros(data.shape[0] + 64, dtype=torch.uint8, device=dev)
        d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
        d_off = torch.from_numpy(off.view(np.int64)).to(dev)
        cap = max(ototal, n, 1)
        cols = sw.DeviceCols(cap, dev, optional=True)
        d_st = torch.zeros(max(n, 1), dtype=torch.uint8, device=dev)
        d_first = torch.zeros(n + 1, dtype=torch.int32, device=dev)
        d_rows = torch.zeros(1, dtype=torch.int32, device=dev)
        torch.cuda.synchronize()
        getattr(ctx, "decode_%s_dev" % fmt)(d_bytes, d_off, n, cols, d_st, d_first, d_rows)
        d_out = torch.zeros(want.shape[0] + 16, dtype=torch.uint8, device=dev)
        d_ooff = torch.zeros(cap + 1, dtype=torch.int64, device=dev)
        d_est = torch.zeros(cap, dtype=torch.uint8, device=dev)
        d_tot = torch.zeros(1, dtype=torch.int64, device=dev)
        ctx.encode_inbound_dev(fmt, d_bytes, d_off, cols, cap, d_rows, b"soak", 1_700_000_000_000, d_out, d_ooff, d_est, d_tot)
        ctx.sync() if hasattr(ctx, "sync") else None
        torch.cuda.synchronize()
        assert int(d_tot.item()) == want.shape[0], "encode size %s %d %d" % (fmt, rnd, seed)
        assert np.array_equal(d_est.cpu().numpy()[:ototal], wst) and np.array_equal(d_out.cpu().numpy()[: want.shape[0]], want), "encode %s %d %d" % (fmt, rnd, seed)
        assert np.array_equal(d_ooff.cpu().numpy().view(np.uint64)[: ototal + 1], woff)
        tot["encode"] = tot.get("encode", 0) + ototal
    rnd += 1
    print("round", rnd, tot, flush=True)
print("soak ok: seed", seed, "rounds", rnd, tot)
