"""swgpu_zone_test_noalloc_dev — the zone test for callers that cannot allocate device memory
while a batch is in flight (a consumer thread sharing the GPU with other contexts, where cudaMalloc
would stall every stream; or a process whose memory budget is already committed to the state table).

Called through the C ABI on device buffers, twice in a row on different batches, and compared with
swgpu_zone_test_dev and with the CPU oracle's index-free loop: the alert pairs and their count must be
identical (bit-exact, as everywhere in this suite).  The second call also shows that the shared
in-library scratch is cleared between calls.  One more run fills the fixed scratch completely
(SWGPU_NOALLOC_MAX_LOCATIONS locations).  This file is collected last so that a failure here cannot
hide any other test.
"""
import ctypes as C

import numpy as np
import pytest

import oracle
from sitewhere_b200 import _lib as L
from sitewhere_b200 import synth as S

pytestmark = pytest.mark.gpu


def _run(ctx, fn, lon, lat, meta, dev):
    import torch

    n = lon.shape[0]
    d_lon = torch.from_numpy(lon).to(dev)
    d_lat = torch.from_numpy(lat).to(dev)
    d_meta = torch.from_numpy(meta.view(np.int64)).to(dev) if meta is not None else None
    cap = max(40 * n, 1024)      # 64 zones, half with the Outside condition: ~32 alerts per location
    d_pairs = torch.full((cap,), -1, dtype=torch.int64, device=dev)
    d_np = torch.full((1,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    fn(d_lon, d_lat, d_meta, n, None, d_pairs, d_np)
    ctx.sync()
    torch.cuda.synchronize()
    total = int(d_np.item())
    assert 0 <= total <= cap
    return total, d_pairs[:total].cpu().numpy().view(np.uint64)


def test_noalloc_matches_allocating_entry_point_twice_in_a_row(ctx):
    import torch

    dev = torch.device("cuda", ctx.device)
    zones = S.star_zones(8, nvert=64, cond_mode="alternate")
    ctx.set_zones(*zones)
    batches = []
    lon, lat = S.locations_for_zones(50_000, *zones[:3], seed=901, boundary_rate=0.01)
    batches.append((lon, lat, None))
    lon2, lat2 = S.locations_for_zones(30_011, *zones[:3], seed=902, boundary_rate=0.02)
    meta2 = np.zeros(30_011, np.uint64)
    meta2[::2] = 1 | ((64 | 128) << 8)          # every other row is a Location with lat/lon
    batches.append((lon2, lat2, meta2))
    for lon_, lat_, meta_ in batches:            # two calls in a row through the new entry point
        want_n, want = _run(ctx, ctx.zone_test_dev, lon_, lat_, meta_, dev)
        got_n, got = _run(ctx, ctx.zone_test_noalloc_dev, lon_, lat_, meta_, dev)
        assert got_n == want_n and want_n > 0
        assert np.array_equal(got, want)
        ref, ref_n = oracle.zone_test(*zones, lon_, lat_, meta=meta_)     # and against the checker itself
        assert got_n == ref_n and np.array_equal(got, ref)


def test_noalloc_at_its_documented_limit(ctx):
    """n = SWGPU_NOALLOC_MAX_LOCATIONS: 65,536 tiles, every byte of the in-library scratch in use"""
    import torch

    dev = torch.device("cuda", ctx.device)
    zones = S.star_zones(8, nvert=64)                      # Inside conditions only: ~0.3 alerts per location
    ctx.set_zones(*zones)
    n = L.NOALLOC_MAX_LOCATIONS
    lon, lat = S.locations_for_zones(n, *zones[:3], seed=903, boundary_rate=1e-4)
    d_lon = torch.from_numpy(lon).to(dev)
    d_lat = torch.from_numpy(lat).to(dev)
    res = []
    for fn in (ctx.zone_test_dev, ctx.zone_test_noalloc_dev):
        d_pairs = torch.full((n,), -1, dtype=torch.int64, device=dev)
        d_np = torch.full((1,), -1, dtype=torch.int64, device=dev)
        fn(d_lon, d_lat, None, n, None, d_pairs, d_np)
        ctx.sync()
        total = int(d_np.item())
        assert 0 < total <= n
        res.append(d_pairs[:total].cpu().numpy().view(np.uint64))
    assert np.array_equal(res[0], res[1])
    # the oracle on the head and the tail of the batch (the last tile is the one that publishes the total)
    k = 100_000
    loc = res[1] >> np.uint64(32)
    for a, b in ((0, k), (n - k, n)):
        ref, ref_n = oracle.zone_test(*zones, lon[a:b], lat[a:b])
        got = res[1][(loc >= a) & (loc < b)]
        assert got.shape[0] == ref_n
        assert np.array_equal(got - (np.uint64(a) << np.uint64(32)), ref)


def test_noalloc_rejects_batches_beyond_its_fixed_scratch(ctx):
    import torch

    dev = torch.device("cuda", ctx.device)
    zones = S.rect_zones(2)
    ctx.set_zones(*zones)
    small = torch.zeros(8, dtype=torch.float64, device=dev)
    pairs = torch.zeros(8, dtype=torch.int64, device=dev)
    npairs = torch.zeros(1, dtype=torch.int64, device=dev)
    rc = ctx.lib.swgpu_zone_test_noalloc_dev(ctx.h, small.data_ptr(), small.data_ptr(), None,
                                             L.NOALLOC_MAX_LOCATIONS + 1, None, 8, pairs.data_ptr(), npairs.data_ptr())
    assert rc == L.E_CAPACITY
    assert b"exceed" in ctx.lib.swgpu_last_error(ctx.h)
    # and an empty batch is fine
    rc = ctx.lib.swgpu_zone_test_noalloc_dev(ctx.h, None, None, None, 0, None, 8, pairs.data_ptr(), npairs.data_ptr())
    ctx.sync()
    assert rc == 0 and int(npairs.item()) == 0
