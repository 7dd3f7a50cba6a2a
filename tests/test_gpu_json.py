"""GPU parity: swgpu_decode_json through the C ABI vs the CPU oracle, bit for bit (status, row order,
every column), on the reference's fixture shapes, every Jackson-quirk vector, number/date torture
sets and random corruptions."""
import random

import numpy as np
import pytest

import oracle
from sitewhere_b200 import synth as S
import sitewhere_b200 as sw
from tests import json_cases
from tests.helpers import check_decode_parity
from tests.test_oracle_json import TORTURE

pytestmark = pytest.mark.gpu


def test_canonical_measurements_c1_shape(ctx):
    data, rec_off, truth = S.json_measurements(60_000, 5000, seed=S.SEED_BASE)
    res = ctx.decode_json(data, rec_off)
    check_decode_parity(res, data, rec_off, fn="decode_json")
    assert res.n_rows == 60_000 and (res.status == 0).all()


def test_known_answer_vectors(ctx):
    cs = json_cases.cases()
    data, rec_off = S.pack_records([c[1] for c in cs])
    res = ctx.decode_json(data, rec_off)
    for i, (name, rec, st, checks) in enumerate(cs):
        assert int(res.status[i]) == st, f"{name}: gpu status {res.status[i]} != {st}"
    check_decode_parity(res, data, rec_off, fn="decode_json")


def test_mixed_event_types(ctx):
    rnd = random.Random(4)
    recs = []
    for i in range(20_000):
        tok = S.token_str(rnd.randrange(3000))
        k = rnd.random()
        d = S.EVENT_DATE_BASE + i
        if k < 0.6:
            recs.append(S.json_measurement(tok, rnd.choice(S.MEASUREMENT_NAMES), repr(round(rnd.uniform(-1e3, 1e3), 3)), d))
        elif k < 0.9:
            recs.append(S.json_location(tok, repr(rnd.uniform(34, 35)), repr(rnd.uniform(-85, -84)), "0.0", d,
                                        update_state=rnd.random() < 0.5,
                                        metadata={"k": "v"} if rnd.random() < 0.3 else None))
        else:
            recs.append(S.json_alert(tok, rnd.choice(S.ALERT_TYPES), "alert raised", d, level=rnd.choice(["Info", "Warning", "Error", "Critical"])))
    data, rec_off = S.pack_records(recs)
    res = ctx.decode_json(data, rec_off)
    check_decode_parity(res, data, rec_off, fn="decode_json")
    assert (res.status == 0).all()


def test_number_torture_as_values(ctx):
    rnd = random.Random(8)
    toks = [t for t in TORTURE] + ["-" + t for t in TORTURE if not t.startswith("-")]
    for _ in range(20_000):
        m = rnd.random()
        if m < 0.4:
            toks.append("%.17g" % (rnd.uniform(-1, 1) * 10.0 ** rnd.randint(-300, 300)))
        elif m < 0.7:
            toks.append("%de%d" % (rnd.randrange(10 ** rnd.randint(1, 19)), rnd.randint(-345, 310)))
        elif m < 0.85:
            toks.append("%d.%d" % (rnd.randrange(10 ** rnd.randint(1, 30)), rnd.randrange(10 ** rnd.randint(1, 30))))
        else:
            toks.append(str(rnd.randrange(10 ** rnd.randint(1, 45))))
    recs = []
    for i, t in enumerate(toks):
        if t.startswith("0") and len(t) > 1 and t[1].isdigit():
            t = "1" + t          # JSON forbids leading zeros
        if i % 3 == 0:
            recs.append(S.json_measurement("tok", "m", t, S.EVENT_DATE_BASE))
        elif i % 3 == 1:
            recs.append(S.json_location("tok", t, t, t, S.EVENT_DATE_BASE))
        else:
            recs.append(S.json_measurement("tok", "m", '"%s"' % t, S.EVENT_DATE_BASE))     # string-coerced
    data, rec_off = S.pack_records(recs)
    res = ctx.decode_json(data, rec_off)
    check_decode_parity(res, data, rec_off, fn="decode_json")


def test_mutation_fuzz(ctx):
    rnd = random.Random(6)
    bases = [S.json_location("25683-LAIPAC-S911-7539094", "34.1", "-84.2", "0.0", 1527854400000, metadata={"a": "b"}),
             S.json_measurement("20544-OPENHAB-7313064", "fuel.level", "123.4", 1527854400000),
             S.json_alert("4545-UNO-9430527", "engine.overheat", "eeeeyyyyyy!", 1527854400000)]
    alphabet = b'{}[]:,"\\ \t0123456789.eE+-tfnrualsxTZ\x00\x01\x7f\xc3\xa9'
    recs = []
    for i in range(30_000):
        r = bytearray(bases[i % 3])
        for _ in range(rnd.randint(1, 3)):
            p = rnd.randrange(len(r))
            op = rnd.random()
            if op < 0.5:
                r[p] = rnd.choice(alphabet)
            elif op < 0.75:
                del r[p]
            else:
                r.insert(p, rnd.choice(alphabet))
        recs.append(bytes(r))
    data, rec_off = S.pack_records(recs)
    res = ctx.decode_json(data, rec_off)
    ocols, ostatus = check_decode_parity(res, data, rec_off, fn="decode_json")
    assert len(set(ostatus.tolist())) >= 3


def test_large_record_and_empty_batch(ctx):
    big_md = ",".join('"k%d":"%s"' % (i, "v" * 50) for i in range(2000))      # ~120 KB record: global-memory path
    recs = [S.json_measurement("a", "m", "1.5", 1),
            ('{"deviceToken":"big","type":"DeviceMeasurement","request":{"name":"m","value":2.5,"metadata":{%s}}}' % big_md).encode(),
            S.json_measurement("b", "m", "3.5", 2)]
    data, rec_off = S.pack_records(recs)
    res = ctx.decode_json(data, rec_off)
    check_decode_parity(res, data, rec_off, fn="decode_json")
    assert res.cols.v0[:3].tolist() == [1.5, 2.5, 3.5]
    res = ctx.decode_json(np.zeros(0, np.uint8), np.zeros(1, np.uint64))
    assert res.n_rows == 0


def test_mirror_api(ctx):
    dec = sw.JsonDeviceRequestDecoder(ctx)
    out = dec.decode(S.json_measurement("20544-OPENHAB-7313064", "fuel.level", "123.4", 1527854400000))
    assert len(out) == 1 and out[0].deviceToken == "20544-OPENHAB-7313064" and out[0].originator is None
    assert out[0].request == {"eventType": "Measurement", "updateState": False, "eventDate": 1527854400000,
                              "name": "fuel.level", "value": 123.4}
    with pytest.raises(sw.EventDecodeException):
        dec.decode(b'{"deviceToken":"t","request":{}}')                # "Event type is required."
    with pytest.raises(sw.EventDecodeException):
        dec.decode(b'{"deviceToken":"t","type":"DeviceMeasurement","request":{"bogus":1}}')
    with pytest.raises(sw.PassThrough):
        dec.decode(b'{"deviceToken":"t","type":"RegisterDevice","request":{}}')


def test_fast_structural_pass_equals_general_pass(ctx):
    """the template recogniser for requests in the reference serialiser's own shape (option json.fast_path,
    default on) and the general pass must give identical results: canonical records of the three event types
    (taken by the template), the same in shuffled key order, with white space, repeated / unknown keys (not
    taken), every known-answer vector, number and date variants inside the canonical text, and mutated records"""
    from tests.test_gpu_fuzz import _mutate, JSON_SPECIALS
    rnd = random.Random(12)
    rng = np.random.default_rng(12)
    recs = [c[1] for c in json_cases.cases()]
    base = []
    for i in range(6000):
        tok = S.token_str(rnd.randrange(500))
        d = S.EVENT_DATE_BASE + i
        k = i % 3
        if k == 0:
            r = S.json_measurement(tok, rnd.choice(S.MEASUREMENT_NAMES), repr(round(rnd.uniform(-1e3, 1e3), 3)), d)
        elif k == 1:
            r = S.json_location(tok, repr(rnd.uniform(34, 35)), repr(rnd.uniform(-85, -84)), "12.5", d)
        else:
            r = S.json_alert(tok, rnd.choice(S.ALERT_TYPES), "alert raised", d, level=rnd.choice(["Info", "Error"]))
        base.append(r)
    recs += base
    import json as _json
    for r in base[:3000]:           # shuffled member order, optional white space, an extra or repeated key now and then
        o = _json.loads(r)
        req = list(o["request"].items())
        rnd.shuffle(req)
        top = list(o.items())
        rnd.shuffle(top)
        sep = rnd.choice([(",", ":"), (", ", ": "), (" , ", " : ")])

        def dump(items):
            out = []
            for k, v in items:
                out.append(_json.dumps(k) + sep[1] + (dump(req) if k == "request" else _json.dumps(v)))
            return "{" + sep[0].join(out) + "}"
        txt = dump(top)
        m = rnd.random()
        if m < 0.05:
            txt = txt[:-1] + ',"deviceToken":"again"}'
        elif m < 0.10:
            txt = txt[:-1] + ',"extra":[1,{"a":2}]}'
        elif m < 0.15:
            txt = txt.replace('"request"' + sep[1] + "{", '"request"' + sep[1] + '{"bogus":1,', 1)
        recs.append(txt.encode())
    recs += _mutate(rng, base[:2000], JSON_SPECIALS, 6000, max_mut=2)
    # canonical text around every torture number / odd dates / long and odd strings
    toks = [t for t in TORTURE] + ["-" + t for t in TORTURE if not t.startswith("-")] + ["0", "-0", "-0.0", "0.0", "1e5", "12345678901234567", "0.1234567890123456", "123456789012345.5", "1.0000000000000000000000001"]
    for i, t in enumerate(toks):
        recs.append(S.json_measurement(S.token_str(i % 50), "fuel.level", t, S.EVENT_DATE_BASE + i))
        recs.append(S.json_location(S.token_str(i % 50), t, "-84.25", t, S.EVENT_DATE_BASE + i, update_state=bool(i & 1)))
    for d in ("2018-13-45T25:61:61.999Z", "1582-01-01T00:00:00.000Z", "9999-12-31T23:59:59.999Z", "2018-06-01T12:00:00.000z", "2018-06-01 12:00:00.000Z"):
        recs.append(S.json_measurement("tok", "n", "1.5", S.EVENT_DATE_BASE).replace(b"2018-06-01T12:00:00.000Z", d.encode()))
    for tok in ("", "a", "x" * 31, "y" * 32, "z" * 40, "t\\u0041", "caf\u00e9", "tab\there"):
        recs.append(S.json_measurement(tok, tok[:8], "2.5", S.EVENT_DATE_BASE))
        recs.append(S.json_alert("dev", tok, "msg " + tok * 3, S.EVENT_DATE_BASE, level="Critical", source="System"))
    recs.append(S.json_measurement("tok", "n", "1.5", S.EVENT_DATE_BASE) + b" ")
    recs.append(S.json_measurement("tok", "n", "1.5", S.EVENT_DATE_BASE)[:-1])
    data, rec_off = S.pack_records(recs)
    ctx.set_option("json.stats", 1)
    fast = ctx.decode_json(data, rec_off)
    st_fast = ctx.json_stats()
    ctx.set_option("json.fast_path", 0)
    try:
        slow = ctx.decode_json(data, rec_off)
        st_slow = ctx.json_stats()
    finally:
        ctx.set_option("json.fast_path", 1)
        ctx.set_option("json.stats", 0)
    # the template really took the canonical records (and none with the option off)
    assert st_fast["template_records"] >= 6000 and st_slow["template_records"] == 0
    assert fast.n_rows == slow.n_rows and np.array_equal(fast.status, slow.status)
    assert np.array_equal(fast.first_row, slow.first_row)
    from tests.helpers import assert_cols_equal
    assert_cols_equal(fast.cols, slow.cols, fast.n_rows)
    check_decode_parity(fast, data, rec_off, fn="decode_json")
    assert (fast.status == 0).sum() > 7000


def test_template_reads_at_the_end_of_a_full_staging_tile(ctx):
    """a warp tile whose bytes fill the 8 KiB stage exactly, its last record cut inside the device token: the
    template's four-byte reads run past the record's end into the stage's slack, never out of the CTA's shared
    memory, and the record is left to the general pass (a decode exception, as in the oracle)"""
    canon = S.json_measurement(S.token_str(7), "fuel.level", "12.5", S.EVENT_DATE_BASE)
    for tile_bytes in (8192, 8191, 8176, 8160):
        recs = []
        for rep in range(3):                      # three such tiles one after another (16-byte aligned starts)
            body = [canon] * 31
            used = sum(len(r) for r in body)
            pad = tile_bytes - used
            assert pad > 20
            last = (b'{"deviceToken":"' + b"x" * pad)[:pad]
            recs += body + [last]
            if tile_bytes % 16:
                recs[-1] = recs[-1] + b"y" * (16 - tile_bytes % 16) if rep < 2 else recs[-1]
        data, rec_off = S.pack_records(recs)
        res = ctx.decode_json(data, rec_off)
        check_decode_parity(res, data, rec_off, fn="decode_json")
        assert int((res.status == 0).sum()) == 93
