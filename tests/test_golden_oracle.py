"""The CPU oracle (and the library's host helpers) against the committed golden vectors.  CPU only."""
import ctypes as C

import numpy as np

import oracle
import sitewhere_b200 as sw
from sitewhere_b200 import _lib as L
from tests import golden_check as G


def test_pb_decode_golden():
    assert G.check_pb(lambda d, o: oracle.decode_pb(d, o)) > 300


def test_json_decode_golden():
    assert G.check_json(lambda d, o: oracle.decode_json(d, o)) > 300


def test_double_parse_golden():
    g = G.load("double_parse.json")
    for e in g:
        t = e["text"]
        want = e["bits"]
        if not any(c in t for c in ".eE") and (int(want, 16) << 1) & 0xFFFFFFFFFFFFFFFF == 0:
            want = "0" * 16          # Jackson: an integer token of value zero is the int 0 -> +0.0, even "-0"
        assert "%016x" % G.bits(oracle.json_number_to_double(t)[1]) == want, t
    assert len(g) > 1000


def test_library_host_parse_double_golden():
    lib = L.load()

    def parse(t):
        out = C.c_double()
        b = t.encode()
        assert lib.swgpu_parse_double(b, len(b), C.byref(out)) == 0
        return out.value
    g = G.load("double_parse.json")
    for e in g:
        assert "%016x" % G.bits(parse(e["text"])) == e["bits"], e["text"]


def test_zone_contains_golden():
    def contains(xs, ys, px, py):
        return [oracle.polygon_contains(xs, ys, float(a), float(b)) for a, b in zip(px, py)]
    assert G.check_zone(contains) > 500


def test_murmur2_golden():
    for key, want in G.load("murmur2.json"):
        assert oracle.murmur2(key.encode()) == want & 0xFFFFFFFF
        assert sw.murmur2(key.encode()) == want & 0xFFFFFFFF
