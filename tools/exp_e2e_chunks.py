"""Experiment: end-to-end time of swgpu_decode_zone_test vs chunk size (pinned host buffers)."""
import sys, time
import numpy as np, torch
sys.path.insert(0, '.')
import sitewhere_b200 as sw
from sitewhere_b200 import synth as S
from sitewhere_b200.api import pinned_array, pinned_cols
import bench

n = 10_000_000
ctx = sw.Context()
zones, data, rec_off = bench.make_workload(n, 0, 1)[:3]
ctx.set_zones(*zones)
total = int(rec_off[-1])
p_data = pinned_array((total,), np.uint8); p_data[:] = data
p_off = pinned_array((n + 1,), np.uint64); p_off[:] = rec_off
cols = pinned_cols(n, ["meta", "event_date", "v0", "v1", "v2", "token_span"])
p_status = pinned_array((n,), np.uint8); p_first = pinned_array((n + 1,), np.uint32); p_pairs = pinned_array((n,), np.uint64)
for chunk in [65536, 131072, 262144, 524288, 851968, 1048576, 2097152, 0]:
    for _ in range(2):
        ctx.decode_zone_test("pb", p_data, p_off, cols, p_status, p_first, p_pairs, chunk_records=chunk)
    t0 = time.perf_counter()
    for _ in range(5):
        ctx.decode_zone_test("pb", p_data, p_off, cols, p_status, p_first, p_pairs, chunk_records=chunk)
    torch.cuda.synchronize()
    print(chunk, "%.2f ms" % ((time.perf_counter() - t0) / 5 * 1e3))
