"""N > 1 host logic on CPU: two processes under gloo.  Checks that (1) Kafka-rule sharding sends every
device to exactly one shard and agrees with the scalar partitioner of the C ABI and the oracle,
(2) the per-shard oracle results recombine to the unsharded result (decode rows, alerts, state),
(3) the zone-table broadcast and counter reduction plumbing works across ranks."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, tmpdir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    import oracle
    from sitewhere_b200 import shard as SH
    from sitewhere_b200 import synth as S

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    dev = torch.device("cpu")
    zones = S.star_zones(6, nvert=16)
    data, rec_off, truth = S.pb_mixed(6000, 300, seed=77, zones=zones)
    shard = SH.shard_of_tokens(S.token_bytes(truth["dev"]), world)
    parts = SH.split_batch(data, rec_off, shard, world)
    my_bytes, my_off, my_idx = parts[rank]

    # this rank's shard through the oracle
    cols, status, first_row, total = oracle.decode_pb(my_bytes, my_off)
    pairs, n_alerts = oracle.zone_test(*zones, cols.v1[:total], cols.v0[:total], meta=cols.meta[:total])
    st = oracle.State()
    ids = my_idx.astype(np.uint64)          # event id = original record index
    st.merge(my_bytes, my_off, cols, 0, ids, ids)
    snap = st.snapshot()

    # a fake context: the zone "table" is just bytes, to exercise the broadcast plumbing
    class FakeCtx:
        def __init__(self):
            self.blob = bytes(range(200)) * 3 if rank == 0 else None

        def zone_table_bytes(self):
            return len(self.blob)

        def zone_table_export(self, t):
            t.copy_(torch.frombuffer(bytearray(self.blob), dtype=torch.uint8))

        def zone_table_import(self, t, n):
            self.blob = bytes(t[:n].tolist())

    fc = FakeCtx()
    n = SH.broadcast_zone_table(fc, dist, dev, rank)
    assert n == 600 and fc.blob == bytes(range(200)) * 3

    tot = SH.sum_counters(dist, dev, events=len(my_idx), rows=total, alerts=n_alerts, keys=snap["seq"].shape[0])
    np.savez(os.path.join(tmpdir, f"r{rank}.npz"), idx=my_idx, alerts=n_alerts, rows=total, keys=snap["seq"].shape[0],
             key_bytes=snap["key_bytes"], key_off=snap["key_off"], id_lo=snap["id_lo"],
             tot_events=tot["events"], tot_alerts=tot["alerts"], tot_keys=tot["keys"],
             alert_rec=(my_idx[(cols.meta[:total] >> np.uint64(32)).astype(np.int64)][(pairs >> np.uint64(32)).astype(np.int64)]
                        if len(pairs) else np.zeros(0, np.int64)),
             alert_test=(pairs & np.uint64(0xFFFFFFFF)).astype(np.int64))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_recombines(tmp_path):
    import torch.multiprocessing as mp

    import oracle
    from sitewhere_b200 import synth as S
    from sitewhere_b200 import shard as SH
    import sitewhere_b200 as sw

    world = 2
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r = [np.load(os.path.join(tmp_path, f"r{k}.npz")) for k in range(world)]

    # unsharded reference
    zones = S.star_zones(6, nvert=16)
    data, rec_off, truth = S.pb_mixed(6000, 300, seed=77, zones=zones)
    cols, status, first_row, total = oracle.decode_pb(data, rec_off)
    pairs, n_alerts = oracle.zone_test(*zones, cols.v1[:total], cols.v0[:total], meta=cols.meta[:total])
    st = oracle.State()
    ids = np.arange(total, dtype=np.uint64)
    st.merge(data, rec_off, cols, 0, ids, ids)
    snap = st.snapshot()

    # (1) a partition of the records, by the Kafka rule; scalar helpers agree
    all_idx = np.sort(np.concatenate([x["idx"] for x in r]))
    assert np.array_equal(all_idx, np.arange(6000))
    toks = S.token_bytes(truth["dev"])
    shard = SH.shard_of_tokens(toks, world)
    for i in range(0, 6000, 97):
        t = bytes(toks[i])
        assert shard[i] == oracle.partition(t, world) == sw.partition(t, world)
    # (2) counters and results recombine
    assert int(r[0]["tot_events"]) == 6000 and int(r[0]["tot_alerts"]) == n_alerts
    assert sum(int(x["rows"]) for x in r) == total
    got_alerts = sorted(zip(np.concatenate([x["alert_rec"] for x in r]).tolist(),
                            np.concatenate([x["alert_test"] for x in r]).tolist()))
    want_alerts = sorted(zip((cols.meta[:total] >> np.uint64(32)).astype(np.int64)[(pairs >> np.uint64(32)).astype(np.int64)].tolist(),
                             (pairs & np.uint64(0xFFFFFFFF)).astype(np.int64).tolist()))
    assert got_alerts == want_alerts
    # device state: every key lives on exactly one shard, and its winner is the global winner
    assert int(r[0]["tot_keys"]) == snap["seq"].shape[0]
    merged = {}
    for x in r:
        kb = x["key_bytes"].tobytes()
        for k in range(int(x["keys"])):
            key = kb[int(x["key_off"][k]): int(x["key_off"][k + 1])]
            assert key not in merged
            merged[key] = int(x["id_lo"][k])
    kb = snap["key_bytes"].tobytes()
    for k in range(snap["seq"].shape[0]):
        key = kb[int(snap["key_off"][k]): int(snap["key_off"][k + 1])]
        assert merged[key] == int(snap["id_lo"][k])
