"""GPU parity: swgpu_set_zones + swgpu_zone_test through the C ABI vs the CPU oracle (index-free
JTS loop), bit for bit, including boundary / vertex / collinear cases."""
import numpy as np
import pytest

import oracle
from sitewhere_b200 import synth as S
import sitewhere_b200 as sw
from tests.test_oracle_zone import WORKAREA, MQTT_LOCATION

pytestmark = pytest.mark.gpu


def run_both(ctx, zones, lon, lat, meta=None):
    vo, zlon, zlat, cond = zones
    ctx.set_zones(vo, zlon, zlat, cond)
    got = ctx.zone_test(lon, lat, meta)
    want, total = oracle.zone_test(vo, zlon, zlat, cond, lon, lat, meta=meta)
    assert got.shape[0] == total
    assert np.array_equal(got, want)
    return got


def adversarial_points(vo, zlon, zlat, rng, n_extra=2000):
    xs, ys = [], []
    T = len(vo) - 1
    for t in rng.choice(T, size=min(T, 64), replace=False):
        a, b = int(vo[t]), int(vo[t + 1])
        zx, zy = zlon[a:b], zlat[a:b]
        xs += list(zx); ys += list(zy)                                   # vertices
        xs += list((zx + np.roll(zx, -1)) / 2); ys += list((zy + np.roll(zy, -1)) / 2)  # edge midpoints
        xs += list(zx + 1e-3); ys += list(zy)                            # ray through a vertex
        xs += list(zx - 1e-3); ys += list(zy)
        xs += [zx.min(), zx.max(), zx.min(), zx.max()]; ys += [zy.min(), zy.max(), zy[0], zy[1]]  # envelope
        xs += list(np.nextafter(zx, np.inf)); ys += list(zy)             # one ulp off a vertex
    xs += list(rng.uniform(zlon.min() - 0.1, zlon.max() + 0.1, n_extra))
    ys += list(rng.uniform(zlat.min() - 0.1, zlat.max() + 0.1, n_extra))
    xs += [np.nan, 0.0, np.inf, -np.inf, zlon[0]]
    ys += [34.5, np.nan, 34.5, 34.5, np.nan]
    return np.array(xs), np.array(ys)


def test_c2_star_zones(ctx):
    zones = S.star_zones(32)           # 1,024 concave 64-vertex zones
    lon, lat = S.locations_for_zones(20_000, *zones[:3], seed=S.SEED_BASE + 1, boundary_rate=0.01)
    got = run_both(ctx, zones, lon, lat)
    assert got.shape[0] > 1000


def test_c1_rectangles_alternating_conditions(ctx):
    zones = S.rect_zones(4)
    lon, lat = S.locations_for_zones(5000, *zones[:3], seed=3, boundary_rate=0.05)
    run_both(ctx, zones, lon, lat)


def test_adversarial_boundaries(ctx):
    rng = np.random.default_rng(1)
    for zones in (S.star_zones(6, nvert=16, seed=2), S.star_zones(5, nvert=64, seed=3, cond_mode="alternate"),
                  S.rect_zones(3), S.star_zones(4, nvert=5, seed=9, cond_mode="outside")):
        lon, lat = adversarial_points(zones[0], zones[1], zones[2], rng)
        run_both(ctx, zones, lon, lat)


def test_reference_fixture(ctx):
    vo = np.array([0, len(WORKAREA)], np.uint32)
    zlon = np.array([p[1] for p in WORKAREA])
    zlat = np.array([p[0] for p in WORKAREA])
    lon = np.array([MQTT_LOCATION[1]] + list(zlon) + [-84.0])
    lat = np.array([MQTT_LOCATION[0]] + list(zlat) + [34.0])
    got = run_both(ctx, (vo, zlon, zlat, np.array([0], np.uint8)), lon, lat)
    assert got.tolist() == [0]            # only the MQTT location is inside; vertices are boundary
    proc = sw.ZoneTestRuleProcessor(ctx, [sw.ZoneTest("workarea", "Inside", "zone.in", "Warning", "in work area"),
                                          sw.ZoneTest("workarea", "Outside", "zone.out", "Error", "left work area")],
                                    {"workarea": WORKAREA})
    alerts = proc.onLocation(None, dict(latitude=MQTT_LOCATION[0], longitude=MQTT_LOCATION[1]))
    assert [a["type"] for a in alerts] == ["zone.in"]
    alerts = proc.onLocation(None, dict(latitude=34.0, longitude=-84.0))
    assert [a["type"] for a in alerts] == ["zone.out"]


def test_overlapping_zones_and_overflow_path(ctx):
    """20 nested squares (not rectangles in isRectangle's sense: 5th vertex inserted) all containing the
    centre: more containing zones than the kernel's register set holds"""
    rng = np.random.default_rng(4)
    vo, zlon, zlat, cond = [0], [], [], []
    for k in range(20):
        r = 1.0 + k * 0.1
        zlon += [-r, 0.0, r, r, -r]
        zlat += [-r, -r, -r, r, r]
        vo.append(len(zlon))
        cond.append(k % 3 == 0)
    zones = (np.array(vo, np.uint32), np.array(zlon), np.array(zlat), np.array(cond, np.uint8))
    lon = np.concatenate([[0.0, 0.5, 1.05, -1.0, 0.0], rng.uniform(-3.5, 3.5, 3000)])
    lat = np.concatenate([[0.0, 0.5, 0.0, -1.0, -1.0], rng.uniform(-3.5, 3.5, 3000)])
    run_both(ctx, zones, lon, lat)


def test_meta_filter_capacity_and_empty(ctx):
    zones = S.star_zones(8, cond_mode="alternate")
    lon, lat = S.locations_for_zones(3000, *zones[:3], seed=8)
    meta = np.zeros(3000, np.uint64)
    meta[::3] = 1 | ((64 | 128) << 8)
    meta[1::3] = 1 | (64 << 8)            # a Location without longitude is skipped
    run_both(ctx, zones, lon, lat, meta)
    assert ctx.zone_test(np.zeros(0), np.zeros(0)).shape[0] == 0
    ctx.set_zones(np.zeros(1, np.uint32), np.zeros(0), np.zeros(0), np.zeros(0, np.uint8))   # no tests at all
    assert ctx.zone_test(lon, lat).shape[0] == 0


def test_degenerate_zones(ctx):
    """zero-area and collinear rings, duplicate vertices, a single-vertex zone"""
    vo = np.array([0, 1, 3, 7, 11], np.uint32)
    zlon = np.array([1.0,  0.0, 2.0,  0.0, 1.0, 1.0, 0.0,  0.0, 2.0, 4.0, 2.0])
    zlat = np.array([1.0,  0.0, 2.0,  0.0, 0.0, 1.0, 1.0,  0.0, 0.0, 0.0, 0.0])
    cond = np.array([0, 0, 0, 1], np.uint8)
    rng = np.random.default_rng(3)
    lon = np.concatenate([zlon, rng.uniform(-1, 5, 500), [1.0, 0.5, 3.0]])
    lat = np.concatenate([zlat, rng.uniform(-1, 3, 500), [1.0, 0.5, 0.0]])
    run_both(ctx, (vo, zlon, zlat, cond), lon, lat)


def _comb(x0, y0, teeth, tall):
    """a comb: `teeth` tiny teeth along the bottom (all their edges fall into the lowest y-slab) and one
    tall spike, so the slabs are very uneven and the table keeps {start,end} ranges for this zone"""
    xs, ys = [x0], [y0]
    for k in range(teeth):
        xs += [x0 + k + 0.5, x0 + k + 1.0]
        ys += [y0 + 0.01, y0]
    xs += [x0 + teeth, x0 + teeth / 2.0]
    ys += [y0 - 1.0, y0 + tall]
    xs += [x0]
    ys += [y0 - 1.0]
    return xs, ys


def test_uneven_slabs_use_ranges(ctx):
    rng = np.random.default_rng(11)
    vo, zlon, zlat = [0], [], []
    for j in range(3):
        xs, ys = _comb(10.0 * j, 5.0 * j, 120, 400.0)
        zlon += xs; zlat += ys
        vo.append(len(zlon))
    star = S.star_zones(3, nvert=32, lon0=0.0, lon1=30.0, lat0=0.0, lat1=30.0, seed=5)
    zlon += list(star[1]); zlat += list(star[2])
    vo += [vo[-1] + int(v) for v in star[0][1:]]
    zones = (np.array(vo, np.uint32), np.array(zlon), np.array(zlat), (np.arange(len(vo) - 1) % 2).astype(np.uint8))
    lon, lat = adversarial_points(zones[0], zones[1], zones[2], rng, n_extra=6000)
    run_both(ctx, zones, lon, lat)


def test_coordinates_beyond_float_range_and_float_envelope_margin(ctx):
    """the cell records hold the envelope as outward-rounded floats: zones beyond the float range, and
    points between the exact and the rounded border, must still match"""
    rng = np.random.default_rng(12)
    vo = np.array([0, 4, 8, 12, 15, 20, 23], np.uint32)
    big = 1e300
    tiny = 1e-310                           # subnormal: products of differences underflow
    zlon = np.array([-big, big, big, -big,   0.1, 0.3, 0.3, 0.1,   1e-300, 3e-300, 2e-300, 0.5e-300,   1e39, 2e39, 1.5e39,
                     -big, big, big, 0.0, -big,   0.0, tiny, 0.0])
    zlat = np.array([-big, -big, big, big,   0.1, 0.1, 0.7, 0.7,   1e-300, 1e-300, 3e-300, 2e-300,    1e39, 1e39, 3e39,
                     -big, -big, big, 1.5 * big, big,   0.0, 0.0, tiny])
    cond = np.array([0, 1, 0, 0, 0, 0], np.uint8)
    xs = [0.0, 0.1, np.nextafter(0.1, 0), np.nextafter(0.1, 1), 0.3, np.nextafter(0.3, 1), np.nextafter(0.3, 0), 0.2,
          1.5e-300, 2e-300, 0.0, big, -big, np.nextafter(big, 0), 1.5e39, 1.2e39, 1e39,
          0.0, big / 2, big / 2, np.nextafter(big / 2, 0), 0.9 * big,
          5e-324, 5e-324, tiny / 2, tiny / 4, np.nextafter(tiny / 2, 0), tiny / 2, tiny, 0.0]
    ys = [0.0, 0.1, 0.4, 0.4, 0.7, 0.4, 0.4, np.nextafter(0.7, 0),
          1.5e-300, 2e-300, 1e-300, big, 0.0, 0.0, 1.5e39, 1.0000001e39, 1e39,
          1.5 * big, 1.25 * big, 1.2 * big, 1.25 * big, 1.04 * big,
          5e-324, 0.0, tiny / 2, tiny / 4, tiny / 2, np.nextafter(tiny / 2, 0), 0.0, tiny]
    lon = np.concatenate([xs, rng.uniform(0.09, 0.31, 2000), rng.uniform(-2, 2, 500) * 1e39,
                          rng.uniform(-1.1, 1.1, 500) * big, rng.uniform(-0.1, 1.1, 500) * tiny])
    lat = np.concatenate([ys, rng.uniform(0.09, 0.71, 2000), rng.uniform(-1, 4, 500) * 1e39,
                          rng.uniform(0.9, 1.6, 500) * big, rng.uniform(-0.1, 1.1, 500) * tiny])
    run_both(ctx, (vo, zlon, zlat, cond), lon, lat)
