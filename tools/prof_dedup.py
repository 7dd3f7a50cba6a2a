"""Times swgpu_dedup_probe_dev: 10 M rows carrying 36-byte alternate ids (UUID strings), half of them repeats of ids of
the same or earlier batches; the ids are the records themselves (meta = HAS_ALT_ID | record index, span = whole record)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import sitewhere_b200 as sw
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
ctx = sw.Context(0, stream=stream.cuda_stream)
n, L = 10_000_000, 36
rng = np.random.default_rng(5)
hexd = np.frombuffer(b"0123456789abcdef", np.uint8)
class Cols: pass
with torch.cuda.stream(stream):
    ctx.dedup_reserve(40_000_000)
    tot_seen = 0
    for b in range(4):
        ids = rng.integers(0, 5_000_000 * (b + 1), size=n)            # about half repeat
        nib = (ids[:, None] * 2654435761 >> (np.arange(32)[None, :] % 16 * 2)) & 15
        rec = np.full((n, L), ord("-"), np.uint8)
        rec[:, [i for i in range(36) if i not in (8, 13, 18, 23)]] = hexd[nib]
        d_bytes = torch.from_numpy(rec.reshape(-1)).to(dev)
        d_off = torch.from_numpy(np.arange(n + 1, dtype=np.int64) * L).to(dev)
        c = Cols()
        c.meta = torch.from_numpy(((np.arange(n, dtype=np.uint64) << np.uint64(32)) | np.uint64((1 << 4) << 8)).view(np.int64)).to(dev)
        c.alternate_id_span = torch.full((n,), L << 32, dtype=torch.int64, device=dev)
        d_dup = torch.zeros(n, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ctx.dedup_probe_dev(d_bytes, d_off, c, n, None, d_dup)
        e1.record(stream)
        torch.cuda.synchronize()
        seen = int((d_dup == 1).sum().item())
        print("batch %d: %.3f ms per %d rows, %.2f G rows/s, %d seen, set holds %d ids" %
              (b, e0.elapsed_time(e1), n, n / e0.elapsed_time(e1) / 1e6, seen, ctx.dedup_size()), flush=True)
        tot_seen += seen
        assert tot_seen + ctx.dedup_size() == n * (b + 1)
