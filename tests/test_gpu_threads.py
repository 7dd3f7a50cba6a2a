"""Several contexts on one GPU, driven from several host threads at once (what the Java side does: one context
per receiver / consumer thread).  Every thread must get the oracle's answers: the library keeps no state outside
a context except three mutex-guarded process-wide caches."""
import threading

import numpy as np
import pytest

import oracle
import sitewhere_b200 as sw
from sitewhere_b200 import synth as S

pytestmark = pytest.mark.gpu


def test_contexts_in_parallel_threads():
    zones = S.star_zones(8, cond_mode="alternate")
    work = []
    for t in range(4):
        data, rec_off, _ = S.pb_mixed(20_000 + 1000 * t, 300, seed=900 + t, zones=zones)
        ocols, ostatus, ofirst, ototal = oracle.decode_pb(data, rec_off)
        want, total = oracle.zone_test(*zones, ocols.v1[:ototal], ocols.v0[:ototal], meta=ocols.meta[:ototal])
        ost = oracle.State()
        ids = np.arange(ototal, dtype=np.uint64)
        ost.merge(data, rec_off, ocols, 0, ids, ids)
        work.append((data, rec_off, ostatus, ofirst, ototal, want, ost.snapshot(), ids))
    errors = []

    def run(t):
        try:
            ctx = sw.Context()
            ctx.set_zones(*zones)
            data, rec_off, ostatus, ofirst, ototal, want, osnap, ids = work[t]
            for it in range(12):
                chunk = (0, 3000, 7000)[it % 3]
                cols, status, first_row, n_rows, pairs = ctx.decode_zone_test("pb", data, rec_off, chunk_records=chunk)
                assert n_rows == ototal and np.array_equal(status, ostatus) and np.array_equal(first_row, ofirst)
                assert np.array_equal(pairs, want)
                small = ctx.decode_zone_test("pb", data[: int(rec_off[500])], rec_off[:501])       # small-batch path
                assert small[3] == int(ofirst[500])
                res = ctx.decode_pb(data, rec_off)
                ctx.state_clear()
                ctx.merge(data, rec_off, res.cols, 0, ids, ids)
                snap = ctx.state_snapshot()
                for k in ("seq", "id_lo", "event_date", "key_bytes"):
                    assert np.array_equal(snap[k], osnap[k]), k
            ctx.close() if hasattr(ctx, "close") else None
        except BaseException as e:   # noqa: BLE001 - reported to the main thread
            errors.append((t, repr(e)))

    threads = [threading.Thread(target=run, args=(t,)) for t in range(4)]
    for th in threads:
        th.start()
    for th in threads:
        th.join(timeout=300)
    assert not any(th.is_alive() for th in threads), "a thread is stuck"
    assert not errors, errors
