"""Shared helpers for the parity tests: compare GPU columns with oracle columns bit for bit."""
import numpy as np

import oracle
from sitewhere_b200 import _lib as L


def assert_cols_equal(gpu_cols, orc_cols, n, optional=True):
    names = L.COL_NAMES if optional else L.MANDATORY_COLS
    for name in names:
        g = getattr(gpu_cols, name)
        o = getattr(orc_cols, name)
        if g is None:
            continue
        g = np.asarray(g[:n])
        o = np.asarray(o[:n])
        # compare raw bits (doubles included: -0.0 vs 0.0 and NaN payloads must match)
        gb = g.view(np.uint64)
        ob = o.view(np.uint64)
        if not np.array_equal(gb, ob):
            bad = np.nonzero(gb != ob)[0]
            i = int(bad[0])
            raise AssertionError(f"column {name}: {bad.size} rows differ; first row {i}: gpu={gb[i]:#x} oracle={ob[i]:#x}")


def check_decode_parity(res, data, rec_off, fn="decode_pb"):
    ocols, ostatus, ofirst, ototal = getattr(oracle, fn)(data, rec_off)
    n = len(rec_off) - 1
    assert res.n_rows == ototal, (res.n_rows, ototal)
    assert np.array_equal(res.status[:n], ostatus[:n]), np.nonzero(res.status[:n] != ostatus[:n])[0][:10]
    assert np.array_equal(res.first_row, ofirst)
    assert_cols_equal(res.cols, ocols, ototal)
    return ocols, ostatus
