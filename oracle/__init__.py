"""ctypes binding of the CPU oracle (oracle/_build/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, as the checker and the timed CPU baseline.  The product
package (sitewhere_b200/) never imports this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")

COL_NAMES = ["meta", "token_hash", "name_hash", "event_date", "v0", "v1", "v2", "token_span",
             "name_span", "originator_span", "message_span", "metadata_span", "alternate_id_span"]
COL_DTYPES = {"event_date": np.int64, "v0": np.float64, "v1": np.float64, "v2": np.float64}


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".c", ".h"))]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


class _ColsStruct(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in COL_NAMES]


ID_NAMES = ["event_hi", "event_lo", "assignment_hi", "assignment_lo", "device_hi", "device_lo"]


class _IdsStruct(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ID_NAMES]


class EventIds:
    """ids carried by enriched events, one element per decoded row (orc_event_ids)"""

    def __init__(self, capacity: int):
        for name in ID_NAMES:
            setattr(self, name, np.zeros(max(capacity, 1), dtype=np.uint64))

    def struct(self) -> _IdsStruct:
        s = _IdsStruct()
        for name in ID_NAMES:
            setattr(s, name, getattr(self, name).ctypes.data)
        return s


class Cols:
    """Decoded rows as numpy columns (same layout as include/swgpu.h swgpu_decoded_cols)."""

    def __init__(self, capacity: int):
        self.capacity = capacity
        self.n = 0
        for name in COL_NAMES:
            setattr(self, name, np.zeros(max(capacity, 1), dtype=COL_DTYPES.get(name, np.uint64)))

    def struct(self) -> _ColsStruct:
        s = _ColsStruct()
        for name in COL_NAMES:
            setattr(s, name, getattr(self, name).ctypes.data)
        return s

    def trimmed(self) -> dict:
        return {name: getattr(self, name)[: self.n] for name in COL_NAMES}


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.orc_murmur2.restype = C.c_uint32
        L.orc_murmur2.argtypes = [C.c_void_p, C.c_uint32]
        L.orc_partition.restype = C.c_uint32
        L.orc_partition.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        L.orc_hash64.restype = C.c_uint64
        L.orc_hash64.argtypes = [C.c_void_p, C.c_uint32]
        for fn in ("orc_decode_pb", "orc_decode_json"):
            if hasattr(L, fn):
                f = getattr(L, fn)
                f.restype = C.c_uint64
                f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.POINTER(_ColsStruct),
                              C.c_void_p, C.c_void_p]
        L.orc_decode_enriched.restype = C.c_uint64
        L.orc_decode_enriched.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.POINTER(_ColsStruct),
                                          C.POINTER(_IdsStruct), C.c_void_p, C.c_void_p]
        L.orc_state_merge_enriched.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(_ColsStruct),
                                               C.POINTER(_IdsStruct), C.c_uint32, C.c_uint64, C.c_void_p]
        if hasattr(L, "orc_json_number_to_double"):
            L.orc_json_number_to_double.restype = C.c_int
            L.orc_json_number_to_double.argtypes = [C.c_char_p, C.c_uint32, C.POINTER(C.c_double)]
            L.orc_parse_date.restype = C.c_int
            L.orc_parse_date.argtypes = [C.c_char_p, C.c_uint32, C.POINTER(C.c_int64)]
        L.orc_sign_det2x2.restype = C.c_int
        L.orc_sign_det2x2.argtypes = [C.c_double] * 4
        L.orc_polygon_contains.restype = C.c_int
        L.orc_polygon_contains.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_double, C.c_double]
        L.orc_polygon_is_rectangle.restype = C.c_int
        L.orc_polygon_is_rectangle.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.orc_zone_test.restype = C.c_uint64
        L.orc_zone_test.argtypes = [C.c_uint32] + [C.c_void_p] * 7 + [C.c_uint32, C.c_uint64, C.c_void_p]
        L.orc_state_new.restype = C.c_void_p
        L.orc_state_free.argtypes = [C.c_void_p]
        L.orc_state_merge.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(_ColsStruct), C.c_uint32,
                                      C.c_uint64, C.c_void_p, C.c_void_p]
        L.orc_state_merge_seq.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(_ColsStruct), C.c_uint32,
                                          C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_state_set_presence_now.argtypes = [C.c_void_p, C.c_uint64]
        L.orc_presence_scan.restype = C.c_uint64
        L.orc_presence_scan.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_state_size.restype = C.c_uint64
        L.orc_state_size.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        L.orc_state_snapshot.argtypes = [C.c_void_p] * 8
        L.orc_encode_inbound.restype = C.c_uint64
        L.orc_encode_inbound.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.POINTER(_ColsStruct), C.c_uint32, C.c_void_p,
                                         C.c_uint32, C.c_int64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _u8(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.uint8)


def murmur2(key: bytes) -> int:
    b = _u8(np.frombuffer(key, dtype=np.uint8)) if key else np.zeros(1, np.uint8)
    return int(lib().orc_murmur2(b.ctypes.data, len(key)))


def partition(key: bytes, n: int) -> int:
    b = _u8(np.frombuffer(key, dtype=np.uint8)) if key else np.zeros(1, np.uint8)
    return int(lib().orc_partition(b.ctypes.data, len(key), n))


def hash64(key: bytes) -> int:
    b = _u8(np.frombuffer(key, dtype=np.uint8)) if key else np.zeros(1, np.uint8)
    return int(lib().orc_hash64(b.ctypes.data, len(key)))


def _decode(fn, data, rec_off, max_rows=None):
    data = _u8(data)
    rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
    n = rec_off.shape[0] - 1
    if max_rows is None:
        max_rows = max(int(data.shape[0]) // 8 + n, n)  # fan-out can never exceed this
    cols = Cols(max_rows)
    status = np.zeros(max(n, 1), dtype=np.uint8)
    first_row = np.zeros(n + 1, dtype=np.uint32)
    dptr = data.ctypes.data if data.size else None
    s = cols.struct()
    total = getattr(lib(), fn)(dptr, rec_off.ctypes.data, n, max_rows, C.byref(s), status.ctypes.data,
                               first_row.ctypes.data)
    cols.n = min(int(total), max_rows)
    return cols, status[:n], first_row, int(total)


def decode_pb(data, rec_off, max_rows=None):
    return _decode("orc_decode_pb", data, rec_off, max_rows)


def decode_json(data, rec_off, max_rows=None):
    return _decode("orc_decode_json", data, rec_off, max_rows)


def decode_enriched(data, rec_off, max_rows=None):
    """GEnrichedEventPayload records -> (cols, ids, status, first_row, total)"""
    data = _u8(data)
    rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
    n = rec_off.shape[0] - 1
    if max_rows is None:
        max_rows = max(n, 1)
    cols = Cols(max_rows)
    ids = EventIds(max_rows)
    status = np.zeros(max(n, 1), dtype=np.uint8)
    first_row = np.zeros(n + 1, dtype=np.uint32)
    s, si = cols.struct(), ids.struct()
    total = lib().orc_decode_enriched(data.ctypes.data if data.size else None, rec_off.ctypes.data, n, max_rows, C.byref(s),
                                      C.byref(si), status.ctypes.data, first_row.ctypes.data)
    cols.n = min(int(total), max_rows)
    return cols, ids, status[:n], first_row, int(total)


def encode_inbound(fmt: str, data, rec_off, cols: "Cols", n_rows: int, source_id: bytes, now_ms: int, out_cap=None):
    """the GInboundEventPayload bytes each decoded row is forwarded as -> (out bytes, out_off[n_rows + 1], enc_status)"""
    data = _u8(data)
    rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
    if out_cap is None:
        out_cap = int(data.shape[0]) * 2 + (64 + len(source_id)) * max(n_rows, 1)
    out = np.zeros(max(out_cap, 1), dtype=np.uint8)
    out_off = np.zeros(n_rows + 1, dtype=np.uint64)
    st = np.zeros(max(n_rows, 1), dtype=np.uint8)
    src = np.frombuffer(source_id, dtype=np.uint8) if source_id else np.zeros(1, np.uint8)
    s = cols.struct()
    total = lib().orc_encode_inbound({"pb": 0, "json": 1}[fmt], data.ctypes.data if data.size else None, rec_off.ctypes.data,
                                     C.byref(s), n_rows, src.ctypes.data, len(source_id), int(now_ms), out.ctypes.data,
                                     out_cap, out_off.ctypes.data, st.ctypes.data)
    return out[: min(int(total), out_cap)], out_off, st[:n_rows]


def json_number_to_double(tok: str):
    out = C.c_double()
    b = tok.encode()
    rc = lib().orc_json_number_to_double(b, len(b), C.byref(out))
    return rc, out.value


def parse_date(s: str):
    out = C.c_int64()
    b = s.encode()
    rc = lib().orc_parse_date(b, len(b), C.byref(out))
    return rc, out.value


def sign_det2x2(x1, y1, x2, y2) -> int:
    return int(lib().orc_sign_det2x2(x1, y1, x2, y2))


def polygon_contains(xs, ys, px, py) -> bool:
    xs = np.ascontiguousarray(xs, dtype=np.float64)
    ys = np.ascontiguousarray(ys, dtype=np.float64)
    return bool(lib().orc_polygon_contains(xs.ctypes.data, ys.ctypes.data, xs.shape[0], px, py))


def polygon_is_rectangle(xs, ys) -> bool:
    xs = np.ascontiguousarray(xs, dtype=np.float64)
    ys = np.ascontiguousarray(ys, dtype=np.float64)
    return bool(lib().orc_polygon_is_rectangle(xs.ctypes.data, ys.ctypes.data, xs.shape[0]))


def zone_test(vtx_off, zlon, zlat, cond, plon, plat, meta=None, max_pairs=None):
    vtx_off = np.ascontiguousarray(vtx_off, dtype=np.uint32)
    zlon = np.ascontiguousarray(zlon, dtype=np.float64)
    zlat = np.ascontiguousarray(zlat, dtype=np.float64)
    cond = _u8(cond)
    plon = np.ascontiguousarray(plon, dtype=np.float64)
    plat = np.ascontiguousarray(plat, dtype=np.float64)
    n = plon.shape[0]
    T = vtx_off.shape[0] - 1
    if max_pairs is None:
        max_pairs = max(1, n * T)
    pairs = np.zeros(max_pairs, dtype=np.uint64)
    mptr = None
    if meta is not None:
        meta = np.ascontiguousarray(meta, dtype=np.uint64)
        mptr = meta.ctypes.data
    total = lib().orc_zone_test(T, vtx_off.ctypes.data, zlon.ctypes.data, zlat.ctypes.data, cond.ctypes.data,
                                plon.ctypes.data, plat.ctypes.data, mptr, n, max_pairs, pairs.ctypes.data)
    return pairs[: min(int(total), max_pairs)], int(total)


class State:
    def __init__(self):
        self._h = lib().orc_state_new()

    def __del__(self):
        try:
            lib().orc_state_free(self._h)
        except Exception:
            pass

    def merge(self, data, rec_off, cols: Cols, seq_base: int, id_hi, id_lo, seq=None):
        """Sequential fold of the rows; row r has sequence seq_base + r, or seq_base + seq[r] when given."""
        if seq is not None:
            return self._merge_seq(data, rec_off, cols, seq_base, id_hi, id_lo, seq)
        data = _u8(data)
        rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
        id_hi = np.ascontiguousarray(id_hi, dtype=np.uint64)
        id_lo = np.ascontiguousarray(id_lo, dtype=np.uint64)
        s = cols.struct()
        lib().orc_state_merge(self._h, data.ctypes.data, rec_off.ctypes.data, C.byref(s), cols.n, seq_base,
                              id_hi.ctypes.data, id_lo.ctypes.data)

    def _merge_seq(self, data, rec_off, cols, seq_base, id_hi, id_lo, seq):
        data = _u8(data)
        rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
        id_hi = np.ascontiguousarray(id_hi, dtype=np.uint64)
        id_lo = np.ascontiguousarray(id_lo, dtype=np.uint64)
        seq = np.ascontiguousarray(seq, dtype=np.uint64)
        s = cols.struct()
        lib().orc_state_merge_seq(self._h, data.ctypes.data, rec_off.ctypes.data, C.byref(s), cols.n, seq_base,
                                  seq.ctypes.data, id_hi.ctypes.data, id_lo.ctypes.data)

    def merge_by_record(self, data, rec_off, cols: Cols, total: int, seq_base: int, id_hi_rec, id_lo_rec, shift: int = 12):
        """The fold swgpu_decode_merge performs: ids per RECORD, row k of record i has sequence
        seq_base + (i << shift) + min(k, 2**shift - 1)  (include/swgpu.h SWGPU_SEQ_RECORD_SHIFT)."""
        rec = (cols.meta[:total].view(np.uint64) >> np.uint64(32)).astype(np.int64)
        first = np.r_[True, rec[1:] != rec[:-1]] if total else np.zeros(0, bool)
        start = np.maximum.accumulate(np.where(first, np.arange(total), 0)) if total else np.zeros(0, np.int64)
        k = np.minimum(np.arange(total) - start, (1 << shift) - 1)
        seq = (rec.astype(np.uint64) << np.uint64(shift)) + k.astype(np.uint64)
        old_n = cols.n
        cols.n = total
        try:
            self._merge_seq(data, rec_off, cols, seq_base, np.asarray(id_hi_rec, np.uint64)[rec],
                            np.asarray(id_lo_rec, np.uint64)[rec], seq)
        finally:
            cols.n = old_n

    def merge_enriched(self, data, rec_off, cols: Cols, ids: EventIds, total: int, seq_base: int, shift: int = 12):
        """The fold swgpu_decode_merge performs on enriched records: key token = the assignment id's 16 bytes, event
        id from the payload, sequence seq_base + (record << shift)."""
        data = _u8(data)
        rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
        rec = (cols.meta[:total].view(np.uint64) >> np.uint64(32))
        seq = np.ascontiguousarray(rec << np.uint64(shift), dtype=np.uint64)
        s, si = cols.struct(), ids.struct()
        lib().orc_state_merge_enriched(self._h, data.ctypes.data, rec_off.ctypes.data, C.byref(s), C.byref(si), total,
                                       seq_base, seq.ctypes.data)

    def set_presence_now(self, now_ms):
        """events folded from now on carry this timestamp as their device's lastInteractionDate (None: off)"""
        lib().orc_state_set_presence_now(self._h, 0 if now_ms is None else int(now_ms) + 1)

    def presence_scan(self, now_ms: int, interval_ms: int):
        """-> (lastInteractionDate uint64[k], [device key bytes]) of the devices reported missing, marked as such"""
        kb = C.c_uint64()
        n = int(lib().orc_state_size(self._h, C.byref(kb)))
        last = np.zeros(max(n, 1), np.uint64)
        off = np.zeros(n + 1, np.uint64)
        keys = np.zeros(max(int(kb.value), 1), np.uint8)
        k = int(lib().orc_presence_scan(self._h, int(now_ms), int(interval_ms), n, last.ctypes.data, off.ctypes.data,
                                        keys.ctypes.data))
        raw = keys.tobytes()
        return last[:k], [raw[int(off[i]): int(off[i + 1])] for i in range(k)]

    def snapshot(self) -> dict:
        kb = C.c_uint64()
        n = int(lib().orc_state_size(self._h, C.byref(kb)))
        out = dict(seq=np.zeros(n, np.uint64), id_hi=np.zeros(n, np.uint64), id_lo=np.zeros(n, np.uint64),
                   event_date=np.zeros(n, np.int64), value=np.zeros(n, np.float64),
                   key_off=np.zeros(n + 1, np.uint64), key_bytes=np.zeros(max(int(kb.value), 1), np.uint8))
        lib().orc_state_snapshot(self._h, *[out[k].ctypes.data for k in
                                            ("seq", "id_hi", "id_lo", "event_date", "value", "key_off", "key_bytes")])
        out["key_bytes"] = out["key_bytes"][: int(kb.value)]
        return out


class AltIdSet:
    """CPU restatement of the alternate-id probe (test infrastructure only).  Reference:
    AlternateIdDeduplicator.isDuplicate, service-event-sources/.../deduplicator/AlternateIdDeduplicator.java:42-57 — a
    request with an alternateId is a duplicate when the event store already holds an event with that id.  Restated as
    the set of ids shown so far, walked in row order (a plain Python loop: small cases only): 0 = new (or no id),
    1 = seen before, 2 = ask the store: the span's bytes are not the id's characters (escapes / coerced token), or the
    id is longer than the 72 bytes the device set keeps."""

    def __init__(self):
        self.seen = set()

    def probe(self, data, rec_off, cols, total):
        data = _u8(data)
        out = np.zeros(total, dtype=np.uint8)
        for i in range(total):
            m = int(cols.meta[i])
            fl = (m >> 8) & 0xffff
            if not fl & (1 << 4):            # SWGPU_F_HAS_ALT_ID
                continue
            sp = int(cols.alternate_id_span[i])
            off, ln = sp & 0xffffffff, sp >> 32
            if fl & ((1 << 9) | (1 << 12)) or ln > 72:    # SWGPU_F_STR_ESCAPED, SWGPU_F_NAME_COERCED; SWGPU_DEDUP_MAX_ID_BYTES
                out[i] = 2
                continue
            base = int(rec_off[m >> 32])
            key = bytes(data[base + off: base + off + ln])
            if key in self.seen:
                out[i] = 1
            else:
                self.seen.add(key)
        return out
