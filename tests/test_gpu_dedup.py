"""GPU parity: swgpu_dedup_probe_dev (the alternate-id set, AlternateIdDeduplicator.isDuplicate as a set probe)
against the oracle's row-order restatement: ids repeated inside one batch and across batches, short, 72-byte and
longer ids (the last key words of a slot; beyond 72 bytes: ask the store), requests without an id, ids with escapes (ask the store), the
set growing from empty through several reallocations, clear."""
import numpy as np
import pytest

import oracle
from sitewhere_b200 import synth as S

pytestmark = pytest.mark.gpu


def _batch(rng, n, pool):
    recs = []
    for i in range(n):
        k = int(rng.integers(0, 10))
        tok = S.token_str(int(rng.integers(0, 50)))
        if k == 0:
            recs.append(S.json_measurement(tok, "fuel.level", "1.5", S.EVENT_DATE_BASE + i))            # no alternate id
            continue
        alt = pool[int(rng.integers(0, len(pool)))]
        recs.append(('{"deviceToken":"%s","type":"DeviceMeasurement","request":{"name":"n","value":%d,"alternateId":"%s"}}'
                     % (tok, i, alt)).encode())
    recs.append(b'{"deviceToken":"t","type":"DeviceLocation","request":{"latitude":1,"alternateId":"esc\\u0041ped"}}')
    recs.append(b'{"deviceToken":"t","type":"DeviceLocation","request":{"latitude":1,"alternateId":12345}}')
    recs.append(b'{"deviceToken":"t","type":"DeviceLocation","request":{"latitude":1,"alternateId":""}}')
    recs.append(b'{"deviceToken":"t","type":"DeviceLocation","request":{"latitude":1,"alternateId":""}}')
    return recs


def test_alternate_id_set_equals_row_order_restatement():
    import torch
    import sitewhere_b200 as sw
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(71)
    pool = ["id-%d" % k for k in range(3000)] + ["x" * 64 + "%08d" % k for k in range(300)] + ["L" * 200 + "%d" % k for k in range(300)]
    pool += ["a", "ab", "abcdefg", "abcdefgh", "abcdefghi", "café"]
    ctx = sw.Context(0, stream=torch.cuda.current_stream(dev).cuda_stream)
    ref = oracle.AltIdSet()
    try:
        for rnd in range(6):
            recs = _batch(rng, [50, 3000, 20_000, 1, 7000, 40_000][rnd], pool)
            data, rec_off = S.pack_records(recs)
            n = rec_off.shape[0] - 1
            ocols, ostatus, ofirst, total = oracle.decode_json(data, rec_off)
            want = ref.probe(data, rec_off, ocols, total)
            d_bytes = torch.zeros(data.shape[0] + 64, dtype=torch.uint8, device=dev)
            d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
            d_off = torch.from_numpy(rec_off.view(np.int64)).to(dev)
            cols = sw.DeviceCols(n + 5, dev, optional=True)
            d_st = torch.zeros(n, dtype=torch.uint8, device=dev)
            d_first = torch.zeros(n + 1, dtype=torch.int32, device=dev)
            d_rows = torch.zeros(1, dtype=torch.int32, device=dev)
            ctx.decode_json_dev(d_bytes, d_off, n, cols, d_st, d_first, d_rows)
            d_dup = torch.full((n + 5,), 9, dtype=torch.uint8, device=dev)
            ctx.dedup_probe_dev(d_bytes, d_off, cols, n + 5, d_rows, d_dup)
            torch.cuda.synchronize()
            got = d_dup.cpu().numpy()
            assert int(d_rows.item()) == total
            assert np.array_equal(got[:total], want), (rnd, np.nonzero(got[:total] != want)[0][:10])
            assert np.all(got[total:] == 9)                      # rows past the live ones are not touched
            assert ctx.dedup_size() == len(ref.seen)
        assert {0, 1, 2}.issubset(set(np.unique(want).tolist()))
        ctx.dedup_reserve(200_000)          # growing an occupied set keeps its ids
        assert ctx.dedup_size() == len(ref.seen)
        ctx.dedup_clear()
        assert ctx.dedup_size() == 0
        # after a clear every id is new again; the device-state table was never touched by any of this
        d_dup.fill_(9)
        ctx.dedup_probe_dev(d_bytes, d_off, cols, n + 5, d_rows, d_dup)
        torch.cuda.synchronize()
        fresh = oracle.AltIdSet().probe(data, rec_off, ocols, total)
        assert np.array_equal(d_dup.cpu().numpy()[:total], fresh)
        assert ctx.state_size()[0] == 0
    finally:
        ctx.close()


def test_host_form(ctx):
    ctx.dedup_clear()
    got = ctx.dedup_probe([b"a", b"b", b"a", b"", b"", b"x" * 72, b"x" * 73, b"b"])
    assert got.tolist() == [0, 0, 1, 0, 1, 0, 2, 1]
    assert ctx.dedup_probe([b"b", b"c", b"x" * 72]).tolist() == [1, 0, 1]
    assert ctx.dedup_size() == 5 and ctx.dedup_probe([]).shape[0] == 0
    ctx.dedup_clear()
