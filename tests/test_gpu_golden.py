"""The CUDA path, through the C ABI, against the committed golden vectors (independent of the oracle)."""
import numpy as np
import pytest

from sitewhere_b200 import synth as S
from tests import golden_check as G

pytestmark = pytest.mark.gpu


def _dec(fn):
    def f(d, o):
        res = fn(d, o)
        return res.cols, res.status, res.first_row, res.n_rows
    return f


def test_pb_decode_golden(ctx):
    G.check_pb(_dec(ctx.decode_pb))


def test_json_decode_golden(ctx):
    G.check_json(_dec(ctx.decode_json))


def test_double_parse_golden_through_json_kernel(ctx):
    g = G.load("double_parse.json")
    recs = []
    for e in g:
        t = e["text"]
        if t.lstrip("-").startswith("0") and len(t.lstrip("-")) > 1 and t.lstrip("-")[1].isdigit():
            t = None      # JSON forbids leading zeros: covered by the string-coerced form below
        recs.append(S.json_measurement("t", "m", t if t is not None else '"%s"' % e["text"], 1))
        recs.append(S.json_measurement("t", "m", '"%s"' % e["text"], 1))          # string -> double coercion
    data, rec_off = S.pack_records(recs)
    res = ctx.decode_json(data, rec_off)
    assert (res.status == 0).all()
    for k, e in enumerate(g):
        for j in (0, 1):
            got = "%016x" % G.bits(res.cols.v0[2 * k + j])
            want = e["bits"]
            is_int_token = j == 0 and not any(c in e["text"] for c in ".eE") and not recs[2 * k].count(b'"value":"')
            if is_int_token and int(want, 16) << 1 & 0xFFFFFFFFFFFFFFFF == 0:
                want = "0" * 16                                                        # integer zero is +0.0
            assert got == want, (e["text"], j, got, want)


def test_zone_contains_golden(ctx):
    def contains(xs, ys, px, py):
        vo = np.array([0, len(xs) - 1], np.uint32)
        ctx.set_zones(vo, np.array(xs[:-1]), np.array(ys[:-1]), np.array([0], np.uint8))
        pairs = ctx.zone_test(px, py)
        out = np.zeros(len(px), dtype=bool)
        out[(pairs >> np.uint64(32)).astype(np.int64)] = True
        return out
    G.check_zone(contains)
