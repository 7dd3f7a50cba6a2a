"""Mutation fuzzing of both decoders against the oracle: well-formed payloads are damaged at random (bit flips,
byte overwrites with structural characters, deletions, insertions, truncations, splices of two payloads) and
every output of the GPU decode — status, first_row, row count, every column — must still equal the oracle's,
bit for bit.  The hand-written cases of pb_cases.py / json_cases.py pin each rule once; this pins their
interactions (a rule reached through a damaged neighbour field, reads near the end of a record, ...)."""
import numpy as np
import pytest

from sitewhere_b200 import synth as S
from tests import json_cases, pb_cases
from tests.helpers import check_decode_parity

pytestmark = pytest.mark.gpu

JSON_SPECIALS = np.frombuffer(b'"\\{}[]:,-+.eE0123456789tfn u\x00\x1f\x7f\x80\xc3\xe2\xf0 \t\n', dtype=np.uint8)
PB_SPECIALS = np.array([0x00, 0x01, 0x02, 0x03, 0x04, 0x05, 0x08, 0x0A, 0x10, 0x11, 0x12, 0x19, 0x1A, 0x21, 0x29, 0x7F,
                        0x80, 0xFF, 0x0B, 0x0C, 0x0D, 0x15, 0x3A], dtype=np.uint8)


def _records(data, rec_off):
    return [bytes(data[int(rec_off[i]): int(rec_off[i + 1])]) for i in range(len(rec_off) - 1)]


def _mutate(rng, base, specials, n_out, max_mut=3):
    out = []
    nb = len(base)
    for _ in range(n_out):
        r = bytearray(base[int(rng.integers(nb))])
        for _ in range(int(rng.integers(1, max_mut + 1))):
            if not r:
                break
            op = int(rng.integers(7))
            p = int(rng.integers(len(r)))
            if op == 0:
                r[p] ^= 1 << int(rng.integers(8))
            elif op == 1:
                r[p] = int(specials[int(rng.integers(len(specials)))])
            elif op == 2:
                del r[p: p + int(rng.integers(1, 6))]
            elif op == 3:
                ins = specials[rng.integers(len(specials), size=int(rng.integers(1, 5)))]
                r[p:p] = bytes(ins.tobytes())
            elif op == 4:
                del r[p:]                                        # truncate
            elif op == 5:
                o = base[int(rng.integers(nb))]                  # splice: head of this one, tail of another
                q = int(rng.integers(len(o) + 1))
                r = bytearray(bytes(r[:p]) + o[q:])
            else:
                q = int(rng.integers(len(r)))                    # duplicate a slice in place
                a, b = min(p, q), max(p, q)
                r[b:b] = r[a:b]
        out.append(bytes(r))
    return out


def _check(ctx, recs, fn):
    data, rec_off = S.pack_records(recs)
    res = getattr(ctx, fn)(data, rec_off)
    check_decode_parity(res, data, rec_off, fn)
    return res


def test_protobuf_mutations(ctx):
    rng = np.random.default_rng(20261107)
    data, rec_off, _ = S.pb_mixed(3000, 200, seed=77)
    base = _records(data, rec_off) + [bytes(c[1]) for c in pb_cases.cases()]
    for rnd in range(5):
        recs = _mutate(rng, base, PB_SPECIALS, 40_000)
        res = _check(ctx, recs, "decode_pb")
        assert 0 < int((res.status[: len(recs)] == 0).sum()) < len(recs)          # both outcomes are exercised


def test_json_mutations(ctx):
    rng = np.random.default_rng(20261108)
    d1, o1, _ = S.json_measurements(1500, 100, seed=78)
    base = _records(d1, o1)
    base += [S.json_location("00001-SYNTH-000000001", "34.5", "-84.25", "12.5", 1540000000000),
             S.json_alert("00001-SYNTH-000000002", "engine.overheat", "too hot", 1540000000001)]
    base += [bytes(c[1]) for c in json_cases.cases()]
    for rnd in range(5):
        recs = _mutate(rng, base, JSON_SPECIALS, 30_000, max_mut=1 + rnd % 3)
        res = _check(ctx, recs, "decode_json")
        assert 0 < int((res.status[: len(recs)] == 0).sum()) < len(recs)


def test_zone_lattice_polygons(ctx):
    """random rings on a small integer lattice (self-intersections, repeated and collinear vertices, zero-area
    spikes, rectangles by chance) tested at lattice and half-lattice points: most queries hit a vertex, an edge
    or an edge's supporting line exactly, on every one of the four ray directions"""
    import oracle
    rng = np.random.default_rng(20261109)
    for rnd in range(6):
        T = 150
        vo, zx, zy = [0], [], []
        for t in range(T):
            nv = int(rng.integers(3, 14))   # (1 or 2 bounds: JTS cannot build the ring; refused by set_zones)
            cx, cy = rng.integers(-20, 20, size=2)
            span = int(rng.integers(1, 9))
            zx += list((cx + rng.integers(-span, span + 1, size=nv)).astype(np.float64) * (0.5 if rnd % 2 else 1.0))
            zy += list((cy + rng.integers(-span, span + 1, size=nv)).astype(np.float64) * (0.5 if rnd % 2 else 1.0))
            vo.append(len(zx))
        zones = (np.array(vo, np.uint32), np.array(zx), np.array(zy), (rng.integers(0, 2, size=T)).astype(np.uint8))
        n = 6000
        lon = rng.integers(-60, 61, size=n).astype(np.float64) * 0.5
        lat = rng.integers(-60, 61, size=n).astype(np.float64) * 0.5
        k = n // 4
        lon[:k] += rng.choice([1e-9, -1e-9, 2.0 ** -40, 0.25], size=k)      # just off the lattice, on one axis
        ctx.set_zones(*zones)
        got = ctx.zone_test(lon, lat, max_pairs=n * (T + 1))
        want, total = oracle.zone_test(*zones, lon, lat)
        assert got.shape[0] == total and np.array_equal(got, want)


@pytest.mark.parametrize("fmt", ["pb", "json"])
def test_every_base_alignment_and_offset_origin(ctx, fmt):
    """device-resident decode with the payload pointer at every alignment mod 16, record offsets that do not
    start at zero, and a batch size that leaves a ragged last warp tile: the staging copies work on
    16-byte-aligned windows around the records and must neither miss bytes nor leave their buffer"""
    import torch
    import oracle
    import sitewhere_b200 as sw
    dev = torch.device("cuda", 0)
    if fmt == "pb":
        data, rec_off, _ = S.pb_mixed(4999, 300, seed=55)
        recs = _records(data, rec_off) + [bytes(c[1]) for c in pb_cases.cases()]
    else:
        d1, o1, _ = S.json_measurements(2999, 100, seed=56)
        recs = _records(d1, o1) + [bytes(c[1]) for c in json_cases.cases()]
    data, rec_off = S.pack_records(recs)
    n = len(recs)
    ocols, ostatus, ofirst, ototal = getattr(oracle, "decode_" + fmt)(data, rec_off)
    st = torch.cuda.ExternalStream(ctx.stream)
    for shift in range(0, 16):
        origin = 1000 + 7 * shift                       # rec_off[0]: the batch is a window of a larger buffer
        with torch.cuda.stream(st):
            buf = torch.zeros(origin + shift + data.shape[0] + 256, dtype=torch.uint8, device=dev)
            buf[origin + shift: origin + shift + data.shape[0]].copy_(torch.from_numpy(data))
            d_off = torch.from_numpy((rec_off + np.uint64(origin)).view(np.int64)).to(dev)
            cols = sw.DeviceCols(ototal + 8, dev, optional=True)
            d_status = torch.empty(n, dtype=torch.uint8, device=dev)
            d_first = torch.empty(n + 1, dtype=torch.int32, device=dev)
            d_nrows = torch.zeros(1, dtype=torch.int32, device=dev)
            getattr(ctx, "decode_%s_dev" % fmt)(buf[shift:], d_off, n, cols, d_status, d_first, d_nrows)
        ctx.sync()
        assert int(d_nrows.item()) == ototal
        assert np.array_equal(d_status.cpu().numpy(), ostatus[:n])
        assert np.array_equal(d_first.cpu().numpy().view(np.uint32), ofirst)
        for name in ("meta", "token_hash", "name_hash", "event_date", "v0", "v1", "v2"):
            g = getattr(cols, name)[:ototal].cpu().numpy().view(np.uint64)
            assert np.array_equal(g, getattr(ocols, name)[:ototal].view(np.uint64)), (shift, name)
        # spans are offsets into the record: the same whatever the window
        for name in ("token_span", "name_span"):
            g = getattr(cols, name)[:ototal].cpu().numpy().view(np.uint64)
            assert np.array_equal(g, getattr(ocols, name)[:ototal].view(np.uint64)), (shift, name)
