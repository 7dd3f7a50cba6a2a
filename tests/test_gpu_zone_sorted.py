"""Larger zone-test batches (many tiles: the count -> scan -> emit split across CTAs, ragged last tile,
both scan kernels).  Same parity bar as the small batches: identical, identically ordered alert pairs."""
import numpy as np
import pytest

import oracle
from sitewhere_b200 import synth as S

pytestmark = pytest.mark.gpu


def run_both(ctx, zones, lon, lat, meta=None):
    ctx.set_zones(*zones)
    got = ctx.zone_test(lon, lat, meta)
    want, total = oracle.zone_test(*zones, lon, lat, meta=meta)
    assert got.shape[0] == total
    assert np.array_equal(got, want)
    return total


def test_sorted_path_inside_conditions(ctx):
    zones = S.star_zones(16)
    lon, lat = S.locations_for_zones(150_000, *zones[:3], seed=21, boundary_rate=0.01)
    lon[::5000] = np.nan                       # not in any envelope
    lat[7::5000] = 1e9
    assert run_both(ctx, zones, lon, lat) > 10_000


def test_sorted_path_with_meta_and_outside_conditions(ctx):
    zones = S.star_zones(6, nvert=16, cond_mode="alternate")     # 18 Outside-condition tests
    lon, lat = S.locations_for_zones(60_000, *zones[:3], seed=22, boundary_rate=0.02)
    meta = np.zeros(60_000, np.uint64)
    meta[::3] = 1 | ((64 | 128) << 8)
    meta[1::3] = 1 | (64 << 8)
    run_both(ctx, zones, lon, lat, meta)


def test_sorted_path_overlapping_zones(ctx):
    rng = np.random.default_rng(4)
    vo, zlon, zlat, cond = [0], [], [], []
    for k in range(12):
        r = 1.0 + k * 0.1
        zlon += [-r, 0.0, r, r, -r]
        zlat += [-r, -r, -r, r, r]
        vo.append(len(zlon))
        cond.append(0)
    zones = (np.array(vo, np.uint32), np.array(zlon), np.array(zlat), np.array(cond, np.uint8))
    lon = rng.uniform(-2.5, 2.5, 70_000)
    lat = rng.uniform(-2.5, 2.5, 70_000)
    run_both(ctx, zones, lon, lat)


def test_sorted_path_rectangles_and_single_cell_grid(ctx):
    zones = S.rect_zones(4)
    lon, lat = S.locations_for_zones(40_000, *zones[:3], seed=23, boundary_rate=0.05)
    run_both(ctx, zones, lon, lat)
    one = (np.array([0, 4], np.uint32), np.array([0.0, 1.0, 1.0, 0.2]), np.array([0.0, 0.0, 1.0, 0.9]), np.array([1], np.uint8))
    rng = np.random.default_rng(5)
    run_both(ctx, one, rng.uniform(-0.5, 1.5, 40_000), rng.uniform(-0.5, 1.5, 40_000))


def test_sequential_scan_kernel_matches(ctx):
    """batches above 67 M locations scan their tile totals with the single-CTA kernel; force it here"""
    zones = S.star_zones(8, nvert=16, cond_mode="alternate")
    lon, lat = S.locations_for_zones(70_000, *zones[:3], seed=21, boundary_rate=0.01)
    ctx.set_zones(*zones)
    want = ctx.zone_test(lon, lat, max_pairs=70_000 * 40)
    ctx.set_option("zone.scan_sequential", 1)
    try:
        got = ctx.zone_test(lon, lat, max_pairs=70_000 * 40)
    finally:
        ctx.set_option("zone.scan_sequential", 0)
    assert np.array_equal(got, want) and len(got) > 70_000
