import sys; sys.path.insert(0,'.')
import numpy as np, torch
import sitewhere_b200 as sw
from sitewhere_b200 import synth as S
zones = S.star_zones(4)
lon, lat = S.locations_for_zones(3000, *zones[:3], seed=1)
data, rec_off, _ = S.pb_locations(3000, 100, lon, lat, seed=1)
free0 = torch.cuda.mem_get_info()[0]
for k in range(150):
    ctx = sw.Context()
    ctx.set_zones(*zones)
    r = ctx.decode_zone_test("pb", data, rec_off, chunk_records=(0, 700)[k % 2])
    ids = np.arange(r[3], dtype=np.uint64)
    ctx.decode_merge("pb", data, rec_off, 0, ids, ids)
    ctx.zone_test(lon, lat)
    ctx.close() if hasattr(ctx, "close") else None
    del ctx
import gc; gc.collect()
free1 = torch.cuda.mem_get_info()[0]
print("device memory delta after 150 context lifetimes: %.1f MB" % ((free0 - free1) / 1e6))
