"""Pins the zone-test half of the CPU oracle (orc_sign_det2x2, orc_polygon_contains, orc_zone_test)
against exact rational arithmetic (fractions.Fraction) and against the reference's own fixtures:
the 'workarea' zone of the construction dataset (service-tenant-management/dockerimage/datasets/
construction/scripts/device-management/content/initializer/deviceModel.groovy:106-112) and the MQTT
test location (sitewhere-communication/src/test/java/com/sitewhere/communication/test/mqtt/MqttTests.java:114-115).
CPU only."""
import math
import random
from fractions import Fraction

import numpy as np

import oracle
from sitewhere_b200 import synth as S

# (lat, lon) pairs exactly as the groovy script lists them
WORKAREA = [(34.10260138703638, -84.24412965774536), (34.101837372446774, -84.24243450164795),
            (34.101517550337825, -84.24091100692749), (34.10154953265732, -84.23856675624847),
            (34.10153176473365, -84.23575580120087), (34.10409030732968, -84.23689305782318),
            (34.104996439280704, -84.23700034618376), (34.10606246444614, -84.23700034618376),
            (34.107691680235604, -84.23690915107727)]
MQTT_LOCATION = (34.10469794977326, -84.23966646194458)  # (lat, lon)


def exact_sign(x1, y1, x2, y2):
    d = Fraction(x1) * Fraction(y2) - Fraction(y1) * Fraction(x2)
    return (d > 0) - (d < 0)


def jts_contains_exact(xs, ys, px, py):
    """JTS 1.13 Polygon.contains(Point) with the crossing side decided in exact rationals.  The
    translation p1 - p is done in double (as RayCrossingCounter does) before the exact sign."""
    minx, maxx, miny, maxy = min(xs), max(xs), min(ys), max(ys)
    if not (minx <= px <= maxx and miny <= py <= maxy):
        return False
    if oracle.polygon_is_rectangle(xs, ys):
        return not (px in (minx, maxx) or py in (miny, maxy))
    crossings = 0
    for i in range(1, len(xs)):
        p1x, p1y, p2x, p2y = xs[i], ys[i], xs[i - 1], ys[i - 1]
        if p1x < px and p2x < px:
            continue
        if px == p2x and py == p2y:
            return False
        if p1y == py and p2y == py:
            if min(p1x, p2x) <= px <= max(p1x, p2x):
                return False
            continue
        if (p1y > py and p2y <= py) or (p2y > py and p1y <= py):
            x1, y1, x2, y2 = p1x - px, p1y - py, p2x - px, p2y - py
            s = exact_sign(x1, y1, x2, y2)
            if s == 0:
                return False
            if y2 < y1:
                s = -s
            if s > 0:
                crossings += 1
    return crossings % 2 == 1


def closed(lonlat):
    xs = [p[1] for p in lonlat] + [lonlat[0][1]]
    ys = [p[0] for p in lonlat] + [lonlat[0][0]]
    return xs, ys


def test_sign_det_exact_random_and_degenerate():
    rnd = random.Random(7)
    for _ in range(20000):
        mode = rnd.random()
        if mode < 0.4:
            v = [rnd.uniform(-360, 360) for _ in range(4)]
        elif mode < 0.7:  # nearly collinear: (x1,y1) ~ k*(x2,y2)
            x2, y2, k = rnd.uniform(-1, 1), rnd.uniform(-1, 1), rnd.uniform(-3, 3)
            v = [x2 * k, y2 * k, x2, y2]
            if rnd.random() < 0.5:
                v[0] = math.nextafter(v[0], math.inf)
        elif mode < 0.85:  # exactly collinear through powers of two
            x2, y2 = rnd.uniform(-1, 1), rnd.uniform(-1, 1)
            k = 2.0 ** rnd.randint(-20, 20)
            v = [x2 * k, y2 * k, x2, y2]
        else:  # tiny / huge magnitudes
            v = [rnd.uniform(-1, 1) * 10.0 ** rnd.randint(-150, 150) for _ in range(4)]
        assert oracle.sign_det2x2(*v) == exact_sign(*v), v
    for _ in range(4000):  # products that underflow to (sub)zero or overflow in double arithmetic
        if rnd.random() < 0.5:
            v = [rnd.choice([5e-324, 1e-310, 3e-200, 0.5, 0.01]) * rnd.choice([1, -1, 3, 0.5]) for _ in range(4)]
        else:
            v = [rnd.uniform(-1, 1) * 10.0 ** rnd.choice([-320, -300, -160, 0, 155, 290, 307]) for _ in range(4)]
        assert oracle.sign_det2x2(*v) == exact_sign(*v), v
    assert oracle.sign_det2x2(0.0, 0.0, 1.0, 1.0) == 0
    assert oracle.sign_det2x2(5e-324, 5e-324, 5e-324, 5e-324) == 0
    assert oracle.sign_det2x2(5e-324, 0.0, 0.0, 5e-324) == 1
    assert oracle.sign_det2x2(1e308, 1e308, 1e308, math.nextafter(1e308, 0)) == -1


def test_reference_fixture_workarea():
    xs, ys = closed(WORKAREA)
    lat, lon = MQTT_LOCATION
    # the MQTT test location lies inside the construction site's work area
    assert jts_contains_exact(xs, ys, lon, lat) is True
    assert oracle.polygon_contains(xs, ys, lon, lat) is True
    assert not oracle.polygon_is_rectangle(xs, ys)
    # every vertex is on the boundary -> not contained; far away -> not contained
    for (la, lo) in WORKAREA:
        assert oracle.polygon_contains(xs, ys, lo, la) is False
    assert oracle.polygon_contains(xs, ys, -84.0, 34.0) is False
    # the zone's own envelope as a rectangle zone: isRectangle path
    rx = [min(xs), max(xs), max(xs), min(xs), min(xs)]
    ry = [min(ys), min(ys), max(ys), max(ys), min(ys)]
    assert oracle.polygon_is_rectangle(rx, ry)
    assert oracle.polygon_contains(rx, ry, lon, lat) is True
    assert oracle.polygon_contains(rx, ry, min(xs), lat) is False        # on the border
    assert oracle.polygon_contains(rx, ry, lon, max(ys)) is False
    # edge midpoint of the two vertices that share a longitude (vertical edge) is on the boundary
    assert oracle.polygon_contains(xs, ys, -84.23700034618376, (34.104996439280704 + 34.10606246444614) / 2) is False


def test_is_rectangle_rules():
    assert oracle.polygon_is_rectangle([0, 1, 1, 0, 0], [0, 0, 1, 1, 0])
    assert oracle.polygon_is_rectangle([0, 0, 1, 1, 0], [0, 1, 1, 0, 0])          # other winding
    assert not oracle.polygon_is_rectangle([0, 1, 1, 0, 0], [0, 0.1, 1, 1, 0])    # not axis aligned
    assert not oracle.polygon_is_rectangle([0, 1, 0, 1, 0], [0, 0, 1, 1, 0])      # bow-tie order: x and y change together
    assert not oracle.polygon_is_rectangle([0, 1, 1, 0.5, 0, 0], [0, 0, 1, 1, 1, 0])  # 6 points
    assert not oracle.polygon_is_rectangle([0, 1, 0, 0], [0, 0, 1, 0])            # triangle


def test_contains_random_polygons_vs_exact():
    rnd = random.Random(11)
    vo, zlon, zlat, cond = S.star_zones(3, nvert=16, seed=5)
    for t in range(9):
        xs = list(zlon[vo[t]:vo[t + 1]]) + [zlon[vo[t]]]
        ys = list(zlat[vo[t]:vo[t + 1]]) + [zlat[vo[t]]]
        pts = [(rnd.uniform(min(xs), max(xs)), rnd.uniform(min(ys), max(ys))) for _ in range(300)]
        # adversarial: vertices, points sharing a vertex's y or x, edge midpoints, envelope corners
        for i in range(len(xs) - 1):
            pts += [(xs[i], ys[i]), (xs[i] + 1e-3, ys[i]), (xs[i] - 1e-3, ys[i]), (xs[i], ys[i] + 1e-4),
                    ((xs[i] + xs[i + 1]) / 2, (ys[i] + ys[i + 1]) / 2)]
        pts += [(min(xs), min(ys)), (max(xs), max(ys)), (min(xs), ys[0])]
        for (px, py) in pts:
            assert oracle.polygon_contains(xs, ys, px, py) == jts_contains_exact(xs, ys, px, py), (t, px, py)


def test_axis_aligned_shapes_boundaries():
    # an L-shaped (non-rectangle) axis-aligned polygon: horizontal and vertical edges, collinear hits
    xs = [0, 4, 4, 2, 2, 0, 0]
    ys = [0, 0, 2, 2, 4, 4, 0]
    grid = [i / 2 for i in range(-1, 10)]
    for px in grid:
        for py in grid:
            assert oracle.polygon_contains(xs, ys, px, py) == jts_contains_exact(xs, ys, px, py), (px, py)
    assert oracle.polygon_contains(xs, ys, 1.0, 1.0)
    assert not oracle.polygon_contains(xs, ys, 3.0, 3.0)
    assert not oracle.polygon_contains(xs, ys, 2.0, 3.0)   # on a vertical edge
    assert not oracle.polygon_contains(xs, ys, 3.0, 2.0)   # on a horizontal edge
    assert oracle.polygon_contains(xs, ys, 1.0, 2.0)       # ray passes through vertex (2,2)/(4,2) edge


def test_zone_test_alert_order_and_conditions():
    vo, zlon, zlat, cond = S.rect_zones(4)
    lon, lat = S.locations_for_zones(500, vo, zlon, zlat, seed=3, boundary_rate=0.05)
    pairs, total = oracle.zone_test(vo, zlon, zlat, cond, lon, lat)
    assert total == len(pairs)
    # ordered by (location, test)
    assert np.all(np.diff(pairs.astype(np.int64)) > 0)
    got = set(int(p) for p in pairs)
    for i in range(500):
        for t in range(16):
            xs = list(zlon[vo[t]:vo[t + 1]]) + [zlon[vo[t]]]
            ys = list(zlat[vo[t]:vo[t + 1]]) + [zlat[vo[t]]]
            inside = jts_contains_exact(xs, ys, lon[i], lat[i])
            fires = (cond[t] == 0) == inside
            assert (((i << 32) | t) in got) == fires
    # meta filter: only Location rows with lat/lon are tested
    meta = np.zeros(500, np.uint64)
    meta[::2] = 1 | ((64 | 128) << 8)      # Location, HAS_V0|HAS_V1
    meta[1::2] = 0 | (64 << 8)             # Measurement
    p2, t2 = oracle.zone_test(vo, zlon, zlat, cond, lon, lat, meta=meta)
    assert set(int(p) >> 32 for p in p2) <= set(range(0, 500, 2))
    assert [int(p) for p in p2] == [int(p) for p in pairs if (int(p) >> 32) % 2 == 0]
    # NaN coordinates are outside every zone: only Outside-condition tests fire
    p3, _ = oracle.zone_test(vo, zlon, zlat, cond, np.array([np.nan]), np.array([34.5]))
    assert [int(p) & 0xFFFFFFFF for p in p3] == [t for t in range(16) if cond[t] == 1]
