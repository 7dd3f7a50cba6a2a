"""Small driver for profiling the merge kernels (and the JSON kernel) alone under ncu."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import sitewhere_b200 as sw
from sitewhere_b200 import synth as S

what = sys.argv[1] if len(sys.argv) > 1 else "merge"
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
ctx = sw.Context(0, stream=stream.cuda_stream)
with torch.cuda.stream(stream):
    if what == "merge":
        n = 4_000_000
        data, rec_off, _ = S.pb_mixed(n, 400_000, seed=5)
    else:
        data, rec_off, _ = S.json_measurements(200_000, 50_000, seed=5)
        lens = np.diff(rec_off.astype(np.int64)); data = np.tile(data, 5)
        rec_off = np.concatenate([[0], np.cumsum(np.tile(lens, 5))]).astype(np.uint64)
        n = rec_off.shape[0] - 1
    d_bytes = torch.empty(data.shape[0] + 64, dtype=torch.uint8, device=dev)
    d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
    d_off = torch.from_numpy(rec_off.view(np.int64)).to(dev)
    cols = sw.DeviceCols(n, dev, optional=False)
    d_status = torch.empty(n, dtype=torch.uint8, device=dev)
    d_first = torch.empty(n + 1, dtype=torch.int32, device=dev)
    d_nrows = torch.zeros(1, dtype=torch.int32, device=dev)
    if what == "merge":
        ctx.decode_pb_dev(d_bytes, d_off, n, cols, d_status, d_first, d_nrows)
        ids = torch.arange(n, dtype=torch.int64, device=dev)
        ctx.state_reserve(8_000_000, 8_000_000 * 48)
        for k in range(4):
            ctx.merge_dev(d_bytes, d_off, n, cols, n, d_nrows, k * n, ids, ids)
    else:
        for k in range(4):
            ctx.decode_json_dev(d_bytes, d_off, n, cols, d_status, d_first, d_nrows)
    torch.cuda.synchronize()
    print(what, "ok", int(d_nrows.item()))
