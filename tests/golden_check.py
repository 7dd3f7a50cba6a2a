"""Checks a decoder / zone-test backend (the CPU oracle, or the CUDA path through the C ABI) against the
committed golden vectors of tests/golden/ (made by tests/golden/make_golden.py from google.protobuf,
CPython json/float and exact rational arithmetic)."""
import json
import os
import struct

import numpy as np

from sitewhere_b200 import _lib as L
from sitewhere_b200 import synth as S

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return json.load(open(os.path.join(HERE, name)))


def bits(x):
    return struct.unpack("<Q", struct.pack("<d", float(x)))[0]


def span(rec, s):
    return rec[int(s) & 0xFFFFFFFF:][: int(s) >> 32]


def check_pb(decode):
    """decode(data, rec_off) -> (cols, status, first_row, total)"""
    g = load("pb_decode.json")
    recs = [bytes.fromhex(e["hex"]) for e in g]
    data, rec_off = S.pack_records(recs)
    cols, status, first_row, total = decode(data, rec_off)
    for i, e in enumerate(g):
        assert int(status[i]) == e["status"], (e["source"], int(status[i]))
        if e["status"] != 0:
            assert first_row[i] == first_row[i + 1]
            continue
        rows = e["rows"]
        assert int(first_row[i + 1]) - int(first_row[i]) == len(rows), e["source"]
        rec = recs[i]
        for k, want in enumerate(rows):
            r = int(first_row[i]) + k
            m = int(cols.meta[r])
            fl = (m >> 8) & 0xFFFF
            assert m & 0xFF == want["kind"] and m >> 32 == i
            assert span(rec, cols.token_span[r]).decode() == want["token"]
            if "name" in want:
                assert span(rec, cols.name_span[r]).decode() == want["name"]
            if "message" in want:
                assert span(rec, cols.message_span[r]).decode() == want["message"]
            for c in ("v0", "v1", "v2"):
                if c in want:
                    assert bits(getattr(cols, c)[r]) == want[c], (e["source"], c)
            if want["date"] is None:
                assert not fl & L.F_HAS_DATE
            else:
                assert fl & L.F_HAS_DATE and int(cols.event_date[r]) & 0xFFFFFFFFFFFFFFFF == want["date"]
            assert bool(fl & L.F_UPDATE_STATE) == want["update"]
            assert (int(cols.metadata_span[r]) >> 32) == want["n_md"]
            if e["originator"] is None:
                assert not fl & L.F_HAS_ORIGINATOR
            else:
                assert span(rec, cols.originator_span[r]).decode() == e["originator"]
    return len(g)


def check_json(decode):
    g = load("json_decode.json")
    recs = [e["text"].encode() if "text" in e else bytes.fromhex(e["hex"]) for e in g]
    data, rec_off = S.pack_records(recs)
    cols, status, first_row, total = decode(data, rec_off)
    from tests.test_oracle_json import check_row
    for i, e in enumerate(g):
        assert int(status[i]) == e["status"], (e["source"], int(status[i]))
        if e["status"] != 0:
            continue
        r = int(first_row[i])
        rec = recs[i]
        if "row" in e:
            want = e["row"]
            m = int(cols.meta[r])
            fl = (m >> 8) & 0xFFFF
            assert m & 0xFF == want["kind"]
            assert span(rec, cols.token_span[r]).decode() == want["token"]
            assert int(cols.event_date[r]) == want["date"] and fl & L.F_HAS_DATE
            assert bool(fl & L.F_UPDATE_STATE) == want["update"]
            for c in ("v0", "v1", "v2"):
                if c in want:
                    assert bits(getattr(cols, c)[r]) == want[c], (e["text"], c)
            if "name" in want:
                assert span(rec, cols.name_span[r]).decode() == want["name"]
            if "message" in want:
                assert span(rec, cols.message_span[r]).decode() == want["message"]
            if "level" in want:
                assert (m >> 24) & 3 == want["level"]
        else:
            checks = {}
            for k, v in e["checks"].items():
                if isinstance(v, dict) and "hex" in v:
                    v = bytes.fromhex(v["hex"])
                elif isinstance(v, dict) and "f64bits" in v:
                    v = struct.unpack("<d", struct.pack("<Q", v["f64bits"]))[0]
                checks[k] = v
            check_row(e["source"], rec, cols, r, checks)
    return len(g)


def check_double(parse):
    """parse(text) -> float"""
    g = load("double_parse.json")
    for e in g:
        got = bits(parse(e["text"]))
        assert "%016x" % got == e["bits"], (e["text"], "%016x" % got, e["bits"])
    return len(g)


def check_zone(contains_batch):
    """contains_batch(xs_closed, ys_closed, px[], py[]) -> bool[]"""
    g = load("zone_contains.json")
    n = 0
    for z in g:
        f = lambda b: struct.unpack("<d", struct.pack("<Q", b))[0]
        xs = [f(b) for b in z["x"]]
        ys = [f(b) for b in z["y"]]
        px = np.array([f(p[0]) for p in z["points"]])
        py = np.array([f(p[1]) for p in z["points"]])
        want = np.array([p[2] for p in z["points"]], dtype=bool)
        got = np.asarray(contains_batch(xs, ys, px, py), dtype=bool)
        assert np.array_equal(got, want), (z["name"], np.nonzero(got != want)[0][:5])
        n += len(want)
    return n
