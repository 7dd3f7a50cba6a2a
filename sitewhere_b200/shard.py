"""Sharding of a micro-batch across GPUs (one process per GPU).

Records are routed by Kafka's DefaultPartitioner rule on the device token —
``toPositive(murmur2(utf8(token))) % n_gpus`` — which is how the reference spreads keyed records over
partitions (sitewhere-microservice/.../kafka/MicroserviceKafkaProducer.java:89-92 sends ``(key = deviceToken,
value)`` without a partition; key chosen at service-event-sources/.../EventSourcesManager.java:163).  All events
of a device therefore reach the same GPU, so decode, zone test and device-state merge need no data-path
collective.  The only collective is the replication of the packed zone table on configuration change
(``broadcast_zone_table``), and the summing of per-shard counters at the end of a run.

Host-side logic only (numpy + torch.distributed); runs under gloo on CPU for tests and under NCCL on B200.
"""
from __future__ import annotations

import numpy as np

from . import synth


def shard_of_tokens(tok: np.ndarray, world: int) -> np.ndarray:
    """tok: uint8[n, L] fixed-width tokens -> shard index per record (Kafka DefaultPartitioner)."""
    if world <= 1:
        return np.zeros(tok.shape[0], dtype=np.int64)
    h = synth.murmur2_rows(tok)
    return ((h & np.uint32(0x7FFFFFFF)) % np.uint32(world)).astype(np.int64)


def split_batch(data: np.ndarray, rec_off: np.ndarray, shard: np.ndarray, world: int):
    """Stable split of a packed batch into per-shard packed batches (record order kept inside a shard,
    as Kafka keeps order inside a partition).  Returns list of (bytes, rec_off, original_index)."""
    out = []
    lens = np.diff(rec_off.astype(np.int64))
    for r in range(world):
        idx = np.nonzero(shard == r)[0]
        l = lens[idx]
        off = np.zeros(idx.size + 1, dtype=np.uint64)
        np.cumsum(l, out=off[1:].view(np.int64))
        buf = np.empty(int(off[-1]), dtype=np.uint8)
        src = rec_off[idx].astype(np.int64)
        # gather each record's bytes (vectorised by expanding to per-byte indices)
        if idx.size:
            rep = np.repeat(np.arange(idx.size), l)
            within = np.arange(int(off[-1])) - np.repeat(off[:-1].astype(np.int64), l)
            buf[:] = data[src[rep] + within]
        out.append((buf, off, idx))
    return out


def broadcast_zone_table(ctx, dist, device, rank: int, src: int = 0):
    """Replicate rank ``src``'s packed zone table to every rank: one broadcast of its size and one of
    its bytes (NCCL over NVLink on B200; gloo in CPU tests passes a fake ctx)."""
    import torch

    nbytes = ctx.zone_table_bytes() if rank == src else 0
    sz = torch.tensor([nbytes], dtype=torch.int64, device=device)
    dist.broadcast(sz, src)
    nbytes = int(sz.item())
    blob = torch.empty(nbytes, dtype=torch.uint8, device=device)
    if rank == src:
        ctx.zone_table_export(blob)
    dist.broadcast(blob, src)
    if rank != src:
        # NCCL broadcasts are ordered against torch's CURRENT stream only and return before the bytes have
        # landed; the import copies on the context's own (non-blocking) stream.  Wait for the receive first.
        if getattr(device, "type", str(device)) == "cuda":
            torch.cuda.current_stream(device).synchronize()
        ctx.zone_table_import(blob, nbytes)
    return nbytes


def sum_counters(dist, device, **counters) -> dict:
    """All-reduce (sum) of per-shard counters (events, rows, alerts, state keys)."""
    import torch

    names = sorted(counters)
    t = torch.tensor([int(counters[k]) for k in names], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return {k: int(v) for k, v in zip(names, t.tolist())}
