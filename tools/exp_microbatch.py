"""Experiment: latency of one 64 KB micro-batch (1,040 protobuf locations) through swgpu_decode_zone_test."""
import sys, time
import numpy as np
sys.path.insert(0, '.')
import sitewhere_b200 as sw
from sitewhere_b200 import synth as S
from sitewhere_b200.api import pinned_array, pinned_cols

ctx = sw.Context()
zones = S.star_zones(32)
ctx.set_zones(*zones)
per = 65536 // 63
nb = 2000
lon, lat = S.locations_for_zones(per * 8, *zones[:3], seed=3)
data, rec_off, _ = S.pb_locations(per * 8, 1_000_000, lon, lat, seed=4)
bufs = []
for k in range(8):
    d = pinned_array((per * 63,), np.uint8); d[:] = data[k * per * 63:(k + 1) * per * 63]
    o = pinned_array((per + 1,), np.uint64); o[:] = rec_off[k * per:(k + 1) * per + 1] - rec_off[k * per]
    bufs.append((d, o))
cols = pinned_cols(per, ["meta", "event_date", "v0", "v1", "v2", "token_span"])
st = pinned_array((per,), np.uint8); fr = pinned_array((per + 1,), np.uint32); pr = pinned_array((per,), np.uint64)
for k in range(50):
    ctx.decode_zone_test("pb", *bufs[k % 8], cols, st, fr, pr)
ts = []
for k in range(nb):
    t0 = time.perf_counter()
    ctx.decode_zone_test("pb", *bufs[k % 8], cols, st, fr, pr)
    ts.append(time.perf_counter() - t0)
ts = np.array(ts) * 1e6
print("per batch us: p50 %.1f p90 %.1f p99 %.1f max %.1f -> %.2f M events/s per calling thread" % (np.percentile(ts, 50), np.percentile(ts, 90), np.percentile(ts, 99), ts.max(), per / ts.mean()))
