"""Driver for profiling the fused decode+merge kernel under ncu: 2 batches of 10 M mixed events into a table that
already holds their keys."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import sitewhere_b200 as sw
from sitewhere_b200 import synth as S
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
ctx = sw.Context(0, stream=stream.cuda_stream)
n = 10_000_000
with torch.cuda.stream(stream):
    bat = []
    for b in range(2):
        data, rec_off, _ = S.pb_mixed(n, 1_000_000, seed=5 + b)
        d_bytes = torch.empty(data.shape[0] + 64, dtype=torch.uint8, device=dev)
        d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
        bat.append((d_bytes, torch.from_numpy(rec_off.view(np.int64)).to(dev)))
    ids = torch.arange(n, dtype=torch.int64, device=dev)
    st = torch.empty(n, dtype=torch.uint8, device=dev)
    rows = torch.zeros(1, dtype=torch.int32, device=dev)
    ctx.state_reserve(32_000_000, 1 << 20)
    seq = 0
    for rep in range(3):
        for b in range(2):
            ctx.decode_merge_dev("pb", bat[b][0], bat[b][1], n, seq, ids, ids, st, rows)
            seq += n << 12
    torch.cuda.synchronize()
    print("ok", int(rows.item()), ctx.state_size())
