"""GPU parity: swgpu_encode_inbound_dev (the GInboundEventPayload of every decoded row, built on the device)
against the oracle's restatement of EventModelConverter.asGrpcInboundEventPayload + writeTo, byte for byte, on
the rows the GPU decoders themselves produce: mixed protobuf events with damaged records and measurement
fan-out, JSON requests of the three kinds with nulls, mismatched event types, escapes and non-ASCII strings;
ragged sizes, every output alignment, an output buffer that is too small, and the empty batch."""
import numpy as np
import pytest

import oracle
from sitewhere_b200 import synth as S
from tests import json_cases, pb_cases

pytestmark = pytest.mark.gpu
NOW = 1_700_000_000_123


def _run(fmt, recs, source=b"mqtt-source-1", out_cap=None, out_shift=0, optional=True):
    import torch
    import sitewhere_b200 as sw
    dev = torch.device("cuda", 0)
    data, rec_off = S.pack_records(recs)
    n = rec_off.shape[0] - 1
    ocols, ostatus, ofirst, ototal = getattr(oracle, "decode_" + fmt)(data, rec_off)
    if not optional:      # the caller did not ask for the optional span columns: the oracle sees them as absent too
        for name in ("originator_span", "message_span", "metadata_span", "alternate_id_span"):
            getattr(ocols, name)[:] = 0
    want, woff, wst = oracle.encode_inbound(fmt, data, rec_off, ocols, ototal, source, NOW)
    st = torch.cuda.current_stream(dev)
    ctx = sw.Context(0, stream=st.cuda_stream)
    try:
        d_bytes = torch.zeros(data.shape[0] + 64, dtype=torch.uint8, device=dev)
        d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
        d_off = torch.from_numpy(rec_off.view(np.int64)).to(dev)
        cap = max(ototal, n, 1) + 3
        cols = sw.DeviceCols(cap, dev, optional=optional)
        d_status = torch.zeros(max(n, 1), dtype=torch.uint8, device=dev)
        d_first = torch.zeros(n + 1, dtype=torch.int32, device=dev)
        d_rows = torch.zeros(1, dtype=torch.int32, device=dev)
        getattr(ctx, "decode_%s_dev" % fmt)(d_bytes, d_off, n, cols, d_status, d_first, d_rows)
        ocap = want.shape[0] + 16 if out_cap is None else out_cap
        d_out_all = torch.full((ocap + 64,), 0xEE, dtype=torch.uint8, device=dev)
        d_out = d_out_all[out_shift: out_shift + ocap]
        d_ooff = torch.zeros(cap + 1, dtype=torch.int64, device=dev)
        d_est = torch.full((cap,), 0x77, dtype=torch.uint8, device=dev)
        d_total = torch.zeros(1, dtype=torch.int64, device=dev)
        for _ in range(2):
            ctx.encode_inbound_dev(fmt, d_bytes, d_off, cols, cap, d_rows, source, NOW, d_out, d_ooff, d_est, d_total)
        torch.cuda.synchronize()
        assert int(d_rows.item()) == ototal
        return (want, woff, wst, ototal, d_out.cpu().numpy(), d_ooff.cpu().numpy().view(np.uint64), d_est.cpu().numpy(),
                int(d_total.item()), d_out_all.cpu().numpy(), cap)
    finally:
        ctx.close()


def _pb_recs(n, seed):
    good, goff, _ = S.pb_mixed(n, 300, seed=seed)
    raw = good.tobytes()
    recs = [raw[int(goff[i]): int(goff[i + 1])] for i in range(n)]
    cs = pb_cases.cases()
    rng = np.random.default_rng(seed)
    for k in range(len(cs)):
        recs.insert(int(rng.integers(0, len(recs))), cs[k][1])
    recs.insert(7, S.pb_measurements_record("fan", [("m%02d" % k, float(k)) for k in range(33)], date=5))
    recs.append(S.pb_measurements_record("z", [("zero", 0.0), ("", 1.5), ("neg", -0.0)], date=0))
    return recs


def _json_recs(n):
    recs = [c[1] for c in json_cases.cases()]
    for i in range(n):
        tok = S.token_str(i % 40)
        d = S.EVENT_DATE_BASE + i
        recs.append(S.json_measurement(tok, S.MEASUREMENT_NAMES[i % len(S.MEASUREMENT_NAMES)], repr(round(i * 0.37 - 200, 3)), d))
        recs.append(S.json_location(tok, repr(33.0 + i * 1e-3), repr(-84.0 - i * 1e-3), "0", d, update_state=bool(i & 1)))
        recs.append(S.json_alert(tok, S.ALERT_TYPES[i % len(S.ALERT_TYPES)], "alert raised " * (1 + i % 5), d,
                                 level=["Info", "Warning", "Error", "Critical"][i % 4], source=["Device", "System"][i % 2]))
    recs += [
        b'{"deviceToken":"t","originator":"orig","type":"DeviceMeasurement","request":{"name":"n","value":2,"alternateId":"alt-1"}}',
        b'{"deviceToken":"t","type":"DeviceMeasurement","request":{"name":"n"}}',
        b'{"deviceToken":"t","type":"DeviceMeasurement","request":{"value":1}}',
        b'{"deviceToken":"t","type":"DeviceAlert","request":{"type":"a","message":"m","level":null}}',
        b'{"deviceToken":"t","type":"DeviceLocation","request":{"latitude":1.5}}',
        b'{"deviceToken":"t","type":"DeviceLocation","request":{"eventType":"Alert","latitude":1.5}}',
        b'{"deviceToken":"t","type":"DeviceLocation","request":null}',
        b'{"deviceToken":"caf\\u00e9","type":"DeviceLocation","request":{"latitude":1}}',
        b'{"deviceToken":"t","originator":"caf\xc3\xa9","type":"DeviceLocation","request":{"latitude":1,"longitude":-0.0}}',
        b'{"deviceToken":"t","type":"DeviceAlert","request":{"type":"a","message":"say \\"hi\\"","metadata":{"k":"v"}}}',
        b'{"deviceToken":"' + b"x" * 300 + b'","type":"DeviceMeasurement","request":{"name":"' + b"n" * 200 + b'","value":1e300,"eventDate":"1969-12-31T23:59:59.000Z"}}',
    ]
    return recs


def _compare(res, exact_fit=True):
    want, woff, wst, total, out, ooff, est, got_total, _, cap = res
    assert got_total == want.shape[0]
    assert np.array_equal(est[:total], wst)
    assert np.all(est[total:] == 0x77)                       # rows past the live ones are not touched
    assert np.array_equal(ooff[: total + 1], woff)
    assert np.all(ooff[total: cap + 1] == woff[total])
    if exact_fit:
        assert np.array_equal(out[: want.shape[0]], want)


@pytest.mark.parametrize("n", [1, 255, 256, 257, 5000])
def test_pb_rows(n):
    res = _run("pb", _pb_recs(n, seed=40 + n) if n > 300 else _pb_recs(n, seed=40 + n)[:n])
    _compare(res)
    assert (res[2] == 0).sum() >= min(n, 200) // 2


def test_json_rows_and_statuses():
    res = _run("json", _json_recs(1200))
    _compare(res)
    st = res[2]
    assert {0, 1, 3, 4}.issubset(set(np.unique(st).tolist()))


def test_without_optional_columns():
    _compare(_run("json", _json_recs(100), optional=False))
    _compare(_run("pb", _pb_recs(600, seed=3), optional=False))


@pytest.mark.parametrize("shift", list(range(0, 16, 3)) + [15])
def test_every_output_alignment_and_no_write_outside(shift):
    res = _run("pb", _pb_recs(700, seed=50), out_shift=shift, source=b"s" * (shift + 1))
    _compare(res)
    want, out_all = res[0], res[8]
    assert np.all(out_all[:shift] == 0xEE) and np.all(out_all[shift + want.shape[0]:] == 0xEE)


def test_output_too_small_reports_exact_size_and_writes_whole_tiles_only():
    recs = _pb_recs(3000, seed=60)
    full = _run("pb", recs)
    want, woff = full[0], full[1]
    cap = int(woff[300 * 1]) + 11            # room for the first tile and a bit
    res = _run("pb", recs, out_cap=cap)
    _compare(res, exact_fit=False)           # offsets, statuses and the total are exact all the same
    out = res[4]
    tile_end = int(woff[256])
    assert np.array_equal(out[:tile_end], want[:tile_end])
    assert np.all(out[tile_end:] == 0xEE)    # the second tile did not fit: none of it was written


def test_empty_batch_and_empty_source():
    res = _run("pb", [])
    assert res[7] == 0 and res[5][0] == 0
    _compare(_run("json", _json_recs(50), source=b""))
