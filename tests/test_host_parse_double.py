"""The decimal -> double routine of the JSON kernel (csrc/dparse.cuh: Clinger fast path, Eisel-Lemire,
exact big-integer fallback), exercised on the host through swgpu_parse_double and compared with
CPython float() (correctly rounded, as Double.parseDouble).  CPU only: the same source compiles for
the device."""
import ctypes as C
import math
import random
import struct
from fractions import Fraction

from sitewhere_b200 import _lib as L
from tests.test_oracle_json import TORTURE


def bits(x):
    return struct.unpack("<Q", struct.pack("<d", x))[0]


def parse(lib, t):
    out = C.c_double()
    b = t.encode()
    assert lib.swgpu_parse_double(b, len(b), C.byref(out)) == 0
    return out.value


def test_torture_and_random():
    lib = L.load()
    rnd = random.Random(21)
    toks = list(TORTURE) + ["-" + t for t in TORTURE if not t.startswith("-")]
    for _ in range(60000):
        mode = rnd.random()
        if mode < 0.25:
            toks.append(repr(rnd.uniform(-1000, 1000)))
        elif mode < 0.4:
            toks.append("%.17g" % (rnd.uniform(-1, 1) * 10.0 ** rnd.randint(-300, 300)))
        elif mode < 0.6:
            toks.append("%de%d" % (rnd.randrange(10 ** rnd.randint(1, 19)), rnd.randint(-345, 310)))
        elif mode < 0.7:
            toks.append("%d.%de%d" % (rnd.randrange(10 ** rnd.randint(1, 25)), rnd.randrange(10 ** rnd.randint(1, 25)),
                                      rnd.randint(-330, 290)))
        elif mode < 0.9:
            # decimal expansions at / next to the half-way point of two adjacent doubles
            x = rnd.uniform(1, 2) * 2.0 ** rnd.randint(-1074, 1022)
            y = math.nextafter(x, math.inf)
            if math.isinf(y):
                continue
            mid = (Fraction(x) + Fraction(y)) / 2
            digits = rnd.choice((20, 40, 400, 800, 1100))
            scaled = mid * 10 ** digits
            qd = scaled.numerator // scaled.denominator
            toks.append(str(qd + rnd.choice((-1, 0, 0, 1))) + "e-%d" % digits)
        else:
            toks.append(str(rnd.randrange(10 ** rnd.randint(1, 60))))
    bad = 0
    for t in toks:
        got, want = parse(lib, t), float(t)
        if bits(got) != bits(want):
            bad += 1
            if bad < 10:
                print("MISMATCH", t[:80], got.hex(), want.hex())
    assert bad == 0


def test_subnormal_and_overflow_boundaries():
    lib = L.load()
    for t in ["4.9406564584124654e-324", "2.4703282292062327e-324", "2.4703282292062328e-324", "2.4703282292062329e-324",
              "1.7976931348623157e308", "1.7976931348623158e308", "1.797693134862315807e308",
              "1.7976931348623158079372897140530341507993413271003782693617377898044496829276475094664901797758720709633028641669288791094655554785194040263065748867150582068190890200070838367627385484581771153176447573027006985557136695962284291481986083493647529271907416844436551070434271155969950809304288017790417449779e308",
              "1.7976931348623158079372897140530341507993413271003782693617377898044496829276475094664901797758720709633028641669288791094655554785194040263065748867150582068190890200070838367627385484581771153176447573027006985557136695962284291481986083493647529271907416844436551070434271155969950809304288017790417449778e308",
              "2.2250738585072011360574097967091319759348195463516456480234261097248222220210769455165295239081350879141491589130396211068700864386945946455276572074078206217433799881410632673292535522868813721490129811224514518898490572223072852551331557550159143974763979834118019993239625482890171070818506906306666559949382757725720157630626906633326475653000092458883164330377797918696120494973903778297049050510806099407302629371289589500035837999672072543043602840788957717961509455167482434710307026091446215722898802581825451803257070188608721131280795122334262883686223215037756666225039825343359745688844239002654981983854879482922068947216898310996983658468140228542433306603398508864458040010349339704275671864433837704860378616227717385456230658746790140867233276367187499e-308"]:
        got, want = parse(lib, t), float(t)
        assert bits(got) == bits(want), (t[:40], got.hex(), want.hex())
