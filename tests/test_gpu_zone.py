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
    """zero-area and collinear rings, duplicate vertices; zones JTS cannot construct are refused"""
    vo = np.array([0, 3, 7, 11], np.uint32)
    zlon = np.array([0.0, 2.0, 0.0,  0.0, 1.0, 1.0, 0.0,  0.0, 2.0, 4.0, 2.0])
    zlat = np.array([0.0, 2.0, 0.0,  0.0, 0.0, 1.0, 1.0,  0.0, 0.0, 0.0, 0.0])
    cond = np.array([0, 0, 1], np.uint8)
    rng = np.random.default_rng(3)
    lon = np.concatenate([zlon, rng.uniform(-1, 5, 500), [1.0, 0.5, 3.0]])
    lat = np.concatenate([zlat, rng.uniform(-1, 3, 500), [1.0, 0.5, 0.0]])
    run_both(ctx, (vo, zlon, zlat, cond), lon, lat)
    # 1 or 2 bounds close to a ring of 2 or 3 points: LinearRing's constructor throws in the reference
    from sitewhere_b200._lib import SwgpuError
    for nb in (1, 2):
        with pytest.raises(SwgpuError, match="at least 3"):
            ctx.set_zones(np.array([0, nb], np.uint32), zlon[:nb], zlat[:nb], np.zeros(1, np.uint8))
    want, total = oracle.zone_test(vo, zlon, zlat, cond, lon[:10], lat[:10])
    assert np.array_equal(ctx.zone_test(lon[:10], lat[:10]), want)   # the previous table is still in place


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


# ---- the walk's choice of ray against the reference's single +x ray ------------------------------------------
# JTS 1.13 RayCrossingCounter casts one ray (+x) and takes the sign of the determinant of the coordinates
# translated IN DOUBLE (oracle/sw_oracle.c ring_interior).  The kernel walks whichever of four axis-parallel
# rays leaves the envelope soonest (zone.cu build_zone_blob explains when that is provably the same answer
# and flags the other zones ZC_XONLY).  Here every ray is forced in turn on points a few ulps around edge
# midpoints, points on edges and vertices of oblique edges.
def _ulp_shift(x, k):
    """x moved by k units in the last place (k may be an array)"""
    i = np.ascontiguousarray(x, np.float64).view(np.int64)
    lo = np.int64(-(2 ** 63))
    o = np.where(i >= 0, i, lo - i) + np.asarray(k, np.int64)
    return np.where(o >= 0, o, lo - o).astype(np.int64).view(np.float64)


def _near_edge_points(vo, zlon, zlat, rng, kmax=8):
    T = len(vo) - 1
    ks = np.arange(-kmax, kmax + 1)
    xs, ys = [], []
    for t in range(T):
        a, b = int(vo[t]), int(vo[t + 1])
        x1, y1 = zlon[a:b], zlat[a:b]
        x2, y2 = np.roll(x1, -1), np.roll(y1, -1)
        tt = rng.uniform(0.05, 0.95, size=x1.shape[0])
        for bx, by in (((x1 + x2) / 2, (y1 + y2) / 2),            # edge midpoints
                       (x1 + tt * (x2 - x1), y1 + tt * (y2 - y1)),  # a point on (next to) each edge
                       (x1, y1)):                                   # the vertices
            xs.append(_ulp_shift(bx[:, None], ks[None, :]).ravel()); ys.append(np.repeat(by, ks.size))
            xs.append(np.repeat(bx, ks.size)); ys.append(_ulp_shift(by[:, None], ks[None, :]).ravel())
            kd = rng.integers(-kmax, kmax + 1, size=(bx.size, 4, 2))
            xs.append(_ulp_shift(bx[:, None], kd[:, :, 0]).ravel()); ys.append(_ulp_shift(by[:, None], kd[:, :, 1]).ravel())
    return np.concatenate(xs), np.concatenate(ys)


def _mismatches(ctx, zones, lon, lat, force):
    ctx.set_option("zone.force_ray", force)
    try:
        got = ctx.zone_test(lon, lat, max_pairs=max(4 * lon.shape[0], 1024))
    finally:
        ctx.set_option("zone.force_ray", 0)
    want, total = oracle.zone_test(*zones, lon, lat, max_pairs=max(4 * lon.shape[0], 1024))
    if got.shape[0] == total and np.array_equal(got, want):
        return 0
    return int(np.setxor1d(got, want).shape[0])


def test_near_edge_all_directions(ctx):
    rng = np.random.default_rng(77)
    report = {}
    # (1) zones where every vertex - point difference is exact (Sterbenz): all four rays must agree with +x
    zones = S.star_zones(26, nvert=16, seed=31)      # 676 zones, 10,816 oblique edges
    ctx.set_zones(*zones)
    lon, lat = _near_edge_points(zones[0], zones[1], zones[2], rng)
    assert lon.shape[0] > 1_000_000
    for force in (0, 1, 2, 3, 4):
        bad = _mismatches(ctx, zones, lon, lat, force)
        report[f"exact_translation_zones/ray{force}"] = bad
        assert bad == 0, (force, bad)
    # (2) zones that straddle the axes (differences round): the kernel restricts itself to +-x, which test
    #     the same edges with the same determinant as the reference.  The +-y rays are measured, not asserted.
    zones = S.star_zones(12, nvert=16, lon0=-1.0, lon1=3.0, lat0=-0.5, lat1=2.0, seed=32)
    ctx.set_zones(*zones)
    lon, lat = _near_edge_points(zones[0], zones[1], zones[2], rng)
    for force in (0, 1, 2):
        bad = _mismatches(ctx, zones, lon, lat, force)
        report[f"rounded_translation_zones/ray{force}"] = bad
        assert bad == 0, (force, bad)
    for force in (3, 4):
        report[f"rounded_translation_zones/ray{force} (not used by the kernel)"] = _mismatches(ctx, zones, lon, lat, force)
    # (3) thin slivers with nearly collinear vertices around the origin: the geometry in which a rounded
    #     translation can mis-order a point and an edge
    vo, zx, zy = [0], [], []
    for k in range(200):
        x0, y0 = rng.uniform(-2, 2, 2)
        dx, dy = rng.uniform(0.2, 1.5, 2) * rng.choice([-1, 1], 2)
        eps = rng.uniform(1e-17, 1e-13)
        zx += [x0, x0 + dx, x0 + 2 * dx, x0 + dx]
        zy += [y0, y0 + dy, y0 + 2 * dy * (1 + eps), y0 + dy * (1 + 1e-9)]
        vo.append(len(zx))
    zones = (np.array(vo, np.uint32), np.array(zx), np.array(zy), np.zeros(200, np.uint8))
    ctx.set_zones(*zones)
    lon, lat = _near_edge_points(zones[0], zones[1], zones[2], rng)
    for force in (0, 1, 2):
        bad = _mismatches(ctx, zones, lon, lat, force)
        report[f"slivers/ray{force}"] = bad
        assert bad == 0, (force, bad)
    for force in (3, 4):
        report[f"slivers/ray{force} (not used by the kernel)"] = _mismatches(ctx, zones, lon, lat, force)
    print("near-edge report:", report)
    import json, os
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(out):
        json.dump(report, open(os.path.join(out, "near_edge_report.json"), "w"), indent=1)
