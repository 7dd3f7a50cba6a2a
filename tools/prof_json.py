"""Times the JSON decode kernel with the fast structural pass on and off (C1 shape, 1 M records), for ncu or alone."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import sitewhere_b200 as sw
from sitewhere_b200 import synth as S
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
ctx = sw.Context(0, stream=stream.cuda_stream)
data, rec_off, _ = S.json_measurements(200_000, 50_000, seed=5)
if len(sys.argv) > 1 and sys.argv[1] == "shuffled":     # the same requests in another member order, ", " / ": " separators
    import json, random
    rnd = random.Random(3)
    raw = data.tobytes()
    recs = []
    for i in range(200_000):
        o = json.loads(raw[int(rec_off[i]): int(rec_off[i + 1])], parse_float=lambda t: t)
        req = list(o["request"].items()); rnd.shuffle(req)
        top = list(o.items()); rnd.shuffle(top)
        def dump(items):
            return "{" + ", ".join(json.dumps(k) + ": " + (dump(req) if k == "request" else (v if isinstance(v, str) and k == "value" else json.dumps(v))) for k, v in items) + "}"
        recs.append(dump(top).encode())
    data, rec_off = S.pack_records(recs)
lens = np.diff(rec_off.astype(np.int64)); data = np.tile(data, 5)
rec_off = np.concatenate([[0], np.cumsum(np.tile(lens, 5))]).astype(np.uint64)
n = rec_off.shape[0] - 1
with torch.cuda.stream(stream):
    d_bytes = torch.empty(data.shape[0] + 64, dtype=torch.uint8, device=dev)
    d_bytes[: data.shape[0]].copy_(torch.from_numpy(data))
    d_off = torch.from_numpy(rec_off.view(np.int64)).to(dev)
    cols = sw.DeviceCols(n, dev, optional=False)
    d_status = torch.empty(n, dtype=torch.uint8, device=dev)
    d_first = torch.empty(n + 1, dtype=torch.int32, device=dev)
    d_nrows = torch.zeros(1, dtype=torch.int32, device=dev)
    for fast in (1, 0, 1):
        ctx.set_option("json.fast_path", fast)
        for _ in range(2):
            ctx.decode_json_dev(d_bytes, d_off, n, cols, d_status, d_first, d_nrows)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(5):
            ctx.decode_json_dev(d_bytes, d_off, n, cols, d_status, d_first, d_nrows)
        b.record(stream)
        torch.cuda.synchronize()
        ctx.set_option("json.stats", 1)
        ctx.decode_json_dev(d_bytes, d_off, n, cols, d_status, d_first, d_nrows)
        print("  stats:", ctx.json_stats(), flush=True)
        ctx.set_option("json.stats", 0)
        print("json.fast_path =", fast, ": %.4f ms per %d records" % (a.elapsed_time(b) / 5, n), int(d_nrows.item()), flush=True)
