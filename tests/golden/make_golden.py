"""Generates the committed golden vectors of tests/golden/ from implementations that are INDEPENDENT of
both the oracle and the CUDA code:
  * google.protobuf's runtime on a hand-built descriptor of sitewhere.proto      -> pb_decode.json
  * CPython json / float() / datetime                                               -> json_decode.json, double_parse.json
  * exact rational arithmetic (fractions.Fraction) for JTS-1.13 Polygon.contains    -> zone_contains.json
  * kafka-clients' published murmur2 test vectors (UtilsTest.testMurmur2)           -> murmur2.json
The reference itself cannot run here (no JVM), so these are the strongest pins available; the expected
values of malformed inputs are the behaviours listed in SURVEY.md §8 (a-P2, a-J4), flagged "spec".
Run:  python tests/golden/make_golden.py        (rewrites the *.json files next to it)
"""
import json
import os
import random
import struct
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from sitewhere_b200 import synth as S  # noqa: E402
from tests import json_cases, pb_cases  # noqa: E402
from tests.test_oracle_pb import _descriptor_pool, _read_delimited  # noqa: E402
from tests.test_oracle_zone import WORKAREA, MQTT_LOCATION, jts_contains_exact  # noqa: E402
from tests.test_oracle_json import TORTURE  # noqa: E402


def bits(x):
    return struct.unpack("<Q", struct.pack("<d", x))[0]


def dump(name, obj):
    with open(os.path.join(HERE, name), "w") as f:
        json.dump(obj, f, indent=0, separators=(",", ":"))
        f.write("\n")
    print(name, os.path.getsize(os.path.join(HERE, name)), "bytes")


def gen_pb():
    cls = _descriptor_pool()
    out = []
    # well-formed payloads: expectations from google.protobuf
    data, rec_off, truth = S.pb_mixed(300, 40, seed=4242)
    raw = data.tobytes()
    recs = [raw[int(rec_off[i]): int(rec_off[i + 1])] for i in range(300)]
    recs += [c[1] for c in pb_cases.cases() if c[2] == 0 and c[0] not in ("length_longer_than_stream", "nested_len_beyond_stream_ok",
                                                                         "long_varint_tag_low32", "trailing_garbage")]
    for rec in recs:
        hb, p = _read_delimited(rec, 0)
        hdr = cls["Header"]()
        hdr.ParseFromString(hb)
        body, _ = _read_delimited(rec, p)
        e = {"hex": rec.hex(), "status": 0, "source": "google.protobuf"}
        if hdr.command == 3:
            m = cls["DeviceLocation"]()
            m.ParseFromString(body)
            e["rows"] = [{"kind": 1, "token": m.hardwareId, "v0": bits(m.latitude), "v1": bits(m.longitude),
                          "v2": bits(m.elevation), "date": m.eventDate if m.HasField("eventDate") else None,
                          "update": bool(m.updateState), "n_md": len(m.metadata)}]
        elif hdr.command == 4:
            m = cls["DeviceAlert"]()
            m.ParseFromString(body)
            e["rows"] = [{"kind": 2, "token": m.hardwareId, "name": m.alertType, "message": m.alertMessage,
                          "date": m.eventDate if m.HasField("eventDate") else None, "update": bool(m.updateState),
                          "n_md": len(m.metadata)}]
        else:
            m = cls["DeviceMeasurements"]()
            m.ParseFromString(body)
            e["rows"] = [{"kind": 0, "token": m.hardwareId, "name": x.measurementId, "v0": bits(x.measurementValue),
                          "date": m.eventDate if m.HasField("eventDate") else None, "update": bool(m.updateState),
                          "n_md": len(m.metadata)} for x in m.measurement]
        e["originator"] = hdr.originator if hdr.HasField("originator") else None
        out.append(e)
    # malformed / edge payloads: statuses per the reference decoder's behaviour (spec)
    for name, rec, st, rows in pb_cases.cases():
        if st != 0:
            out.append({"hex": rec.hex(), "status": st, "source": "spec:" + name})
    dump("pb_decode.json", out)


def gen_json():
    out = []
    rnd = random.Random(99)
    # canonical payloads: expectations from CPython json + float + the generator's own date
    for i in range(200):
        tok = S.token_str(rnd.randrange(1000))
        d = S.EVENT_DATE_BASE + rnd.randrange(10 ** 9)
        k = i % 3
        if k == 0:
            txt = rnd.choice(["%.17g" % rnd.uniform(-1e6, 1e6), repr(round(rnd.uniform(-1e3, 1e3), 3)),
                              "%de%d" % (rnd.randrange(10 ** 12), rnd.randint(-30, 30)), str(rnd.randrange(10 ** 25))])
            rec = S.json_measurement(tok, rnd.choice(S.MEASUREMENT_NAMES), txt, d)
        elif k == 1:
            rec = S.json_location(tok, repr(rnd.uniform(-90, 90)), repr(rnd.uniform(-180, 180)), repr(rnd.uniform(0, 1e4)), d,
                                  update_state=rnd.random() < 0.5)
        else:
            rec = S.json_alert(tok, rnd.choice(S.ALERT_TYPES), "msg %d" % i, d, level=rnd.choice(["Info", "Warning", "Error", "Critical"]))
        doc = json.loads(rec)
        rq = doc["request"]
        row = {"token": doc["deviceToken"], "date": d, "update": rq["updateState"]}
        if k == 0:
            v = float(rq["value"])
            if isinstance(rq["value"], int) and v == 0:
                v = 0.0
            row.update(kind=0, name=rq["name"], v0=bits(v))
        elif k == 1:
            row.update(kind=1, v0=bits(float(rq["latitude"])), v1=bits(float(rq["longitude"])), v2=bits(float(rq["elevation"])))
        else:
            row.update(kind=2, name=rq["type"], message=rq["message"], level=["Info", "Warning", "Error", "Critical"].index(rq["level"]))
        out.append({"text": rec.decode(), "status": 0, "row": row, "source": "cpython json/float"})
    for name, rec, st, checks in json_cases.cases():
        e = {"hex": rec.hex(), "status": st, "source": "spec:" + name}
        if st == 0:
            row = {}
            for k, v in checks.items():
                if isinstance(v, bytes):
                    v = {"hex": v.hex()}
                elif isinstance(v, float):
                    v = {"f64bits": bits(v)}
                row[k] = v
            e["checks"] = row
        out.append(e)
    dump("json_decode.json", out)


def gen_double():
    rnd = random.Random(5)
    toks = list(TORTURE) + ["-" + t for t in TORTURE if not t.startswith("-")]
    for _ in range(1500):
        m = rnd.random()
        if m < 0.4:
            toks.append("%.17g" % (rnd.uniform(-1, 1) * 10.0 ** rnd.randint(-300, 300)))
        elif m < 0.7:
            toks.append("%de%d" % (rnd.randrange(10 ** rnd.randint(1, 19)), rnd.randint(-345, 310)))
        else:
            toks.append("%d.%d" % (rnd.randrange(10 ** rnd.randint(1, 30)), rnd.randrange(10 ** rnd.randint(1, 30))))
    dump("double_parse.json", [{"text": t, "bits": "%016x" % bits(float(t))} for t in toks])


def gen_zone():
    rnd = random.Random(17)
    out = []
    polys = []
    xs = [p[1] for p in WORKAREA] + [WORKAREA[0][1]]
    ys = [p[0] for p in WORKAREA] + [WORKAREA[0][0]]
    polys.append(("workarea (deviceModel.groovy:106-112)", xs, ys))
    polys.append(("workarea envelope rectangle", [min(xs), max(xs), max(xs), min(xs), min(xs)], [min(ys), min(ys), max(ys), max(ys), min(ys)]))
    polys.append(("L shape", [0, 4, 4, 2, 2, 0, 0], [0, 0, 2, 2, 4, 4, 0]))
    vo, zl, za, _ = S.star_zones(2, nvert=24, seed=31)
    for t in range(4):
        polys.append(("star %d" % t, list(zl[vo[t]:vo[t + 1]]) + [zl[vo[t]]], list(za[vo[t]:vo[t + 1]]) + [za[vo[t]]]))
    for name, px, py in polys:
        pts = [(MQTT_LOCATION[1], MQTT_LOCATION[0])] if "workarea" in name else []
        for i in range(len(px) - 1):
            pts += [(px[i], py[i]), ((px[i] + px[i + 1]) / 2, (py[i] + py[i + 1]) / 2), (px[i] + 1e-4, py[i]), (px[i] - 1e-4, py[i])]
        pts += [(rnd.uniform(min(px), max(px)), rnd.uniform(min(py), max(py))) for _ in range(60)]
        pts += [(min(px), min(py)), (max(px), max(py)), (min(px) - 1, min(py))]
        out.append({"name": name, "x": [bits(v) for v in px], "y": [bits(v) for v in py],
                    "points": [[bits(a), bits(b), bool(jts_contains_exact(px, py, a, b))] for a, b in pts]})
    dump("zone_contains.json", out)


def gen_murmur():
    dump("murmur2.json", [["21", -973932308], ["foobar", -790332482], ["a-little-bit-long-string", -985981536],
                          ["a-little-bit-longer-string", -1486304829],
                          ["lkjh234lh9fiuh90y23oiuhsafujhadof229phr9h19h89h8", -58897971], ["abc", 479470107]])


if __name__ == "__main__":
    gen_pb()
    gen_json()
    gen_double()
    gen_zone()
    gen_murmur()
