"""Pins the JSON half of the CPU oracle (oracle/sw_oracle_json.c) against
(1) known answers for every Jackson behaviour listed in SURVEY.md §8 a-J4 (tests/json_cases.py),
(2) CPython's json module for the grammar (accept/reject) on mutated payloads,
(3) CPython float() for number tokens (correctly rounded, as Double.parseDouble) on a torture set,
(4) CPython datetime for ISO-8601 dates.
CPU only."""
import json
import math
import random
import struct

import numpy as np

import oracle
from sitewhere_b200 import synth as S
from sitewhere_b200 import _lib as L
from tests import json_cases


def bits(x):
    return struct.unpack("<Q", struct.pack("<d", x))[0]


def span(rec, s):
    return rec[int(s) & 0xFFFFFFFF:][: int(s) >> 32]


def check_row(name, rec, cols, r, checks):
    m = int(cols.meta[r])
    fl = (m >> 8) & 0xFFFF
    for k, v in checks.items():
        if k == "v0":
            assert bits(float(cols.v0[r])) == bits(float(v)), (name, cols.v0[r], v)
        elif k == "v1":
            assert bits(float(cols.v1[r])) == bits(float(v)), name
        elif k == "v2":
            assert bits(float(cols.v2[r])) == bits(float(v)), name
        elif k == "v0bits":
            assert bits(float(cols.v0[r])) == v, (name, hex(bits(float(cols.v0[r]))))
        elif k == "name":
            if v is None:
                assert not fl & L.F_HAS_NAME, name
            else:
                assert fl & L.F_HAS_NAME and span(rec, cols.name_span[r]) == v, (name, span(rec, cols.name_span[r]))
                assert int(cols.name_hash[r]) == oracle.hash64(v)
        elif k == "message":
            assert fl & L.F_HAS_MESSAGE and span(rec, cols.message_span[r]) == v, name
        elif k == "alt":
            assert fl & L.F_HAS_ALT_ID and span(rec, cols.alternate_id_span[r]) == v, name
        elif k == "token":
            assert span(rec, cols.token_span[r]) == v and int(cols.token_hash[r]) == oracle.hash64(v), name
        elif k == "originator":
            if v is None:
                assert not fl & L.F_HAS_ORIGINATOR, name
            else:
                assert fl & L.F_HAS_ORIGINATOR and span(rec, cols.originator_span[r]) == v, name
        elif k == "date":
            if v is None:
                assert not fl & L.F_HAS_DATE, name
            else:
                assert fl & L.F_HAS_DATE and int(cols.event_date[r]) == v, (name, int(cols.event_date[r]), v)
        elif k == "kind":
            assert m & 0xFF == v, name
        elif k == "update":
            assert bool(fl & L.F_UPDATE_STATE) == v, name
        elif k == "has_md":
            assert bool(fl & L.F_HAS_METADATA) == v, name
            if v:
                json.loads(span(rec, cols.metadata_span[r]))
        elif k == "null_token":
            assert bool(fl & L.F_NULL_TOKEN) == v, name
        elif k == "null_request":
            assert bool(fl & (1 << 13)) == v, name
        elif k == "null_level":
            assert bool(fl & (1 << 14)) == v, name
        elif k == "null_source":
            assert bool(fl & (1 << 15)) == v, name
        elif k == "escaped":
            assert bool(fl & L.F_STR_ESCAPED) == v, name
        elif k == "coerced":
            assert bool(fl & L.F_NAME_COERCED) == v, name
        elif k == "level":
            assert (m >> 24) & 3 == v, name
        elif k == "source":
            assert (m >> 26) & 1 == v, name
        elif k == "etype":
            if v is None:
                assert not (m >> 27) & 1, name
            else:
                assert (m >> 27) & 1 and (m >> 28) & 7 == v, name
        elif k in ("has_v0", "has_v1", "has_v2"):
            bit = {"has_v0": L.F_HAS_V0, "has_v1": L.F_HAS_V1, "has_v2": L.F_HAS_V2}[k]
            assert bool(fl & bit) == v, name
        else:
            raise KeyError(k)


def test_known_answers():
    cs = json_cases.cases()
    data, rec_off = S.pack_records([c[1] for c in cs])
    cols, status, first_row, total = oracle.decode_json(data, rec_off)
    for i, (name, rec, st, checks) in enumerate(cs):
        assert int(status[i]) == st, f"{name}: status {status[i]} != {st}"
        rows = int(first_row[i + 1]) - int(first_row[i])
        assert rows == (1 if st == 0 else 0), name
        if st == 0:
            check_row(name, rec, cols, int(first_row[i]), checks)
            assert int(cols.meta[int(first_row[i])]) >> 32 == i


def test_canonical_batches_match_python_json():
    data, rec_off, truth = S.json_measurements(5000, 300, seed=S.SEED_BASE)
    cols, status, first_row, total = oracle.decode_json(data, rec_off)
    assert (status == 0).all() and total == 5000
    raw = data.tobytes()
    for i in range(5000):
        rec = raw[int(rec_off[i]): int(rec_off[i + 1])]
        doc = json.loads(rec)
        assert bits(float(cols.v0[i])) == bits(float(doc["request"]["value"])), rec
        assert span(rec, cols.name_span[i]).decode() == doc["request"]["name"]
        assert span(rec, cols.token_span[i]).decode() == doc["deviceToken"]
        assert int(cols.event_date[i]) == int(truth["date"][i])
        assert int(cols.token_hash[i]) & 0xFFFFFFFF == oracle.murmur2(doc["deviceToken"].encode())


def test_grammar_agrees_with_python_json_on_mutations():
    """Random single-byte mutations of a valid payload: wherever CPython's strict parser rejects the
    document, so must the oracle (status DECODE or PASS); wherever the oracle says OK, CPython must
    accept the prefix that holds the first JSON value."""
    rnd = random.Random(5)
    base = S.json_location("25683-LAIPAC-S911-7539094", "34.1", "-84.2", "0.0", 1527854400000, metadata={"a": "b"})
    alphabet = b'{}[]:,"\\ \t0123456789.eE+-tfnrualsx\x00\x01\x7f\xc3\xa9'
    recs = []
    for _ in range(6000):
        r = bytearray(base)
        for _ in range(rnd.randint(1, 2)):
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
    cols, status, first_row, total = oracle.decode_json(data, rec_off)
    dec = json.JSONDecoder(strict=True)
    n_ok = 0
    for i, rec in enumerate(recs):
        try:
            txt = rec.decode("utf-8")
            if txt.startswith("﻿"):
                txt = txt[1:]
            # Jackson whitespace = space, tab, CR, LF
            doc, end = dec.raw_decode(txt.lstrip(" \t\r\n"))
            py_ok = isinstance(doc, dict)
        except (ValueError, UnicodeDecodeError, RecursionError):
            py_ok = False
        if int(status[i]) == 0:
            n_ok += 1
            assert py_ok, rec
        # CPython accepts a few things Jackson rejects (NaN/Infinity literals, lone surrogates in \u escapes are fine in both,
        # control chars are rejected by both in strict mode); the reverse direction is the strong check
        if not py_ok:
            assert int(status[i]) != 0, rec
    assert n_ok > 100


TORTURE = [
    "0", "-0", "0.0", "-0.0", "1", "-1", "0.1", "0.2", "0.3", "1e0", "1E0", "1e+0", "1e-0", "123.4", "1.7976931348623157e308",
    "1.7976931348623158e308", "1.7976931348623159e308", "1.797693134862315807e308", "1.8e308", "2e308", "1e309",
    "4.9e-324", "4.9406564584124654e-324", "2.4703282292062327e-324", "2.4703282292062328e-324", "2.47e-324", "1e-324", "1e-400",
    "2.2250738585072014e-308", "2.2250738585072011e-308", "2.2250738585072009e-308", "2.225073858507201e-308",
    "9007199254740992", "9007199254740993", "9007199254740994", "9007199254740995", "9007199254740993.0",
    "9007199254740992.5", "9007199254740993.5000000000000000000000000000000001", "1e23", "8.5e22", "1e22", "1e21", "9.5e21",
    "0.000001", "1e-5", "123456789012345678", "1234567890123456789", "12345678901234567890", "123456789012345678901",
    "0.1e1", "100e-2", "5e-324", "3e-324", "2e-324", "7.4109846876186981626485318930233205854758970392148714663837852375101326090531312779794975454245398856969484704316857659638998506553390969459816219401617281718945106978546710679176872575177347315553307795408549809608457500958111373034747658096871009590975442271004757307809711118935784838675653998783503015228055934046593739791790738723868299395818481660169122019456499931289798411362062484498678713572180352209017023903285791732520220528974020802906854021606612375549983402671300035812486479041385743401875520901590172592547146296175134159774938718574737870961645638908718119841271673056017045493004705269590165763776884908267986972573366521765567941072508764337560846003984904972149117463085539556354188641513168478436313080237596295773983001708984375e-324",
    "1.00000000000000011102230246251565404236316680908203125", "1.00000000000000011102230246251565404236316680908203124",
    "1.00000000000000011102230246251565404236316680908203126", "0.500000000000000166533453693773481063544750213623046875",
    "3.14159265358979323846264338327950288419716939937510", "179769313486231580793728971405303415079934132710037826936173778980444968292764750946649017977587207096330286416692887910946555547851940402630657488671505820681908902000708383676273854845817711531764475730270069855571366959622842914819860834936475292719074168444365510704342711559699508093042880177904174497791.999999999999999999999999999999999999999999999999999999999999999999999999",
    "1e-320", "1.5e-320", "0.99999999999999999", "0.999999999999999944488848768742172978818416595458984375",
    "0.9999999999999999444888487687421729788184165954589843749", "17976931348623157e292", "1e-2147483647", "1e2147483647",
    "1e99999999999999999999", "0e99999999999999999999", "0.0e-99999999999999999999", "1E+308", "1E-308",
]


def test_number_tokens_match_python_float():
    rnd = random.Random(9)
    toks = list(TORTURE) + ["-" + t for t in TORTURE if not t.startswith("-")]
    for _ in range(30000):
        mode = rnd.random()
        if mode < 0.3:
            toks.append(repr(rnd.uniform(-1000, 1000)))
        elif mode < 0.5:
            toks.append("%.17g" % rnd.uniform(-1e6, 1e6))
        elif mode < 0.7:
            toks.append("%de%d" % (rnd.randrange(10 ** rnd.randint(1, 19)), rnd.randint(-340, 308)))
        elif mode < 0.85:
            # halfway cases between adjacent doubles, in decimal with many digits
            x = rnd.uniform(1, 2) * 2.0 ** rnd.randint(-1070, 1020)
            y = math.nextafter(x, math.inf)
            from fractions import Fraction
            mid = (Fraction(x) + Fraction(y)) / 2
            digits = 800
            scaled = mid * 10 ** digits
            q = scaled.numerator // scaled.denominator
            s = str(q + rnd.choice((-1, 0, 0, 1)))
            toks.append(s + "e-%d" % digits if len(s) < 1200 else repr(x))
        else:
            toks.append(str(rnd.randrange(10 ** rnd.randint(1, 40))))
    for t in toks:
        t = t.replace("inf", "1e999").replace("nan", "0")
        if t.startswith("+") or t.startswith(".") or "e+-" in t:
            continue
        rc, got = oracle.json_number_to_double(t)
        assert rc == 0
        is_float = any(ch in t for ch in ".eE")
        want = float(t)
        if not is_float and want == 0:
            want = 0.0                      # integer zero is +0.0 even when written "-0"
        if t == "2.2250738585072012e-308":
            want = 5e-324                   # NumberInput.NASTY_SMALL_DOUBLE
        assert bits(got) == bits(want), (t, got, want)


def test_iso_dates_match_python_datetime():
    import datetime as dtm
    rnd = random.Random(3)
    epoch = dtm.datetime(1970, 1, 1, tzinfo=dtm.timezone.utc)
    for _ in range(5000):
        t = epoch + dtm.timedelta(milliseconds=rnd.randrange(-10_000_000_000_000, 40_000_000_000_000))
        if t.year < 1583:
            continue
        ms = int((t - epoch) / dtm.timedelta(milliseconds=1))
        off_min = rnd.choice([0, 0, 60, -60, 330, -270, 845])
        local = t + dtm.timedelta(minutes=off_min)
        form = rnd.randrange(4)
        base = local.strftime("%Y-%m-%dT%H:%M:%S") + ".%03d" % (ms % 1000)
        if off_min == 0 and form == 0:
            s = base + "Z"
        else:
            sign = "+" if off_min >= 0 else "-"
            a = abs(off_min)
            s = base + sign + ("%02d:%02d" if form % 2 else "%02d%02d") % (a // 60, a % 60)
        rc, got = oracle.parse_date(s)
        assert rc == 0 and got == ms, (s, got, ms)
    assert S.iso_date(1527854400123) == "2018-06-01T12:00:00.123Z"
    assert oracle.parse_date(S.iso_date(1527854400123)) == (0, 1527854400123)
