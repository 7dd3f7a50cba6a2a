"""Experiment: zone test time vs the ORDER of the locations (random / sorted by grid cell / sorted by
coarse 16x16 region).  Tells how much of the zone kernel's time is table misses (L1 locality)."""
import sys, time
import numpy as np, torch
sys.path.insert(0, '.')
import sitewhere_b200 as sw
from sitewhere_b200 import synth as S

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
ctx = sw.Context()
zones = S.star_zones(32)
ctx.set_zones(*zones)
lon, lat = S.locations_for_zones(n, *zones[:3], seed=5)
x0, x1, y0, y1 = zones[1].min(), zones[1].max(), zones[2].min(), zones[2].max()
def run(order, label):
    lo = torch.from_numpy(lon[order]).cuda(); la = torch.from_numpy(lat[order]).cuda()
    pairs = torch.zeros(n, dtype=torch.int64, device='cuda'); npairs = torch.zeros(1, dtype=torch.int64, device='cuda')
    st = torch.cuda.ExternalStream(ctx.stream)
    with torch.cuda.stream(st):
        for _ in range(3): ctx.zone_test_dev(lo, la, None, n, None, pairs, npairs)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(5): ctx.zone_test_dev(lo, la, None, n, None, pairs, npairs)
        e1.record(st)
    ctx.sync(); torch.cuda.synchronize()
    print(label, "%.3f ms" % (e0.elapsed_time(e1) / 5), int(npairs.item()))
ident = np.arange(n)
for g, label in [(0, "random order"), (64, "sorted by 64x64 cell"), (16, "sorted by 16x16 region"), (8, "sorted by 8x8 region")]:
    if g == 0:
        run(ident, label); continue
    cx = np.clip(((lon - x0) / (x1 - x0) * g).astype(np.int64), 0, g - 1)
    cy = np.clip(((lat - y0) / (y1 - y0) * g).astype(np.int64), 0, g - 1)
    run(np.argsort(cy * g + cx, kind="stable"), label)
