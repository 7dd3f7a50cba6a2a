"""Host side of the hot path above the C ABI.

Two layers:
  * :class:`Context` — a thin, typed wrapper over ``include/swgpu.h`` (numpy for host buffers,
    torch tensors for device-resident buffers; torch is plumbing only).
  * mirrors of the reference's plug-in classes with the same names, argument meaning and error
    behaviour, so the parity tests read like tests of the reference:
      ``ProtobufDeviceEventDecoder.decode``  (service-event-sources/.../decoder/protobuf/ProtobufDeviceEventDecoder.java:75)
      ``JsonDeviceRequestDecoder.decode``    (service-event-sources/.../decoder/json/JsonDeviceRequestDecoder.java:48)
      ``ZoneTestRuleProcessor.onLocation``   (service-rule-processing/.../geospatial/ZoneTestRuleProcessor.java:47)
      ``DeviceStateProcessingLogic.process`` (service-device-state/.../processing/DeviceStateProcessingLogic.java:65)
    each with a batch form, which is what a GPU needs (the Java binding micro-batches callers).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

from . import _lib as L

_COL_DT = {"event_date": np.int64, "v0": np.float64, "v1": np.float64, "v2": np.float64}


class EventDecodeException(Exception):
    """com.sitewhere.sources.spi.EventDecodeException (service-event-sources/.../spi/EventDecodeException.java:17)."""


class DecoderRuntimeException(RuntimeError):
    """An unchecked exception of the reference decoder (e.g. the NullPointerException when
    parseDelimitedFrom returns null): it is NOT routed to the failed-decode topic."""


class PassThrough(Exception):
    """Non-event request (registration, acknowledgement, stream): outside the GPU path."""


# --------------------------------------------------------------------------------------------- columns
class HostCols:
    """Decoded rows in host memory (numpy), layout of swgpu_decoded_cols."""

    def __init__(self, capacity: int, optional: bool = True):
        self.capacity = int(capacity)
        self.n = 0
        for name in L.COL_NAMES:
            if name in L.MANDATORY_COLS or optional:
                setattr(self, name, np.zeros(max(self.capacity, 1), dtype=_COL_DT.get(name, np.uint64)))
            else:
                setattr(self, name, None)

    def struct(self) -> L.DecodedColsStruct:
        s = L.DecodedColsStruct()
        for name in L.COL_NAMES:
            a = getattr(self, name)
            setattr(s, name, a.ctypes.data if a is not None else None)
        return s

    def col(self, name):
        a = getattr(self, name)
        return None if a is None else a[: self.n]

    # field views ---------------------------------------------------------------------------------
    @property
    def kind(self):
        return (self.meta[: self.n] & np.uint64(0xFF)).astype(np.uint8)

    @property
    def flags(self):
        return ((self.meta[: self.n] >> np.uint64(8)) & np.uint64(0xFFFF)).astype(np.uint32)

    @property
    def rec(self):
        return (self.meta[: self.n] >> np.uint64(32)).astype(np.uint32)


class DeviceCols:
    """Decoded rows in device memory (torch tensors, dtype int64/float64 views of the 8-byte columns)."""

    def __init__(self, capacity: int, device, optional: bool = False):
        import torch

        self.capacity = int(capacity)
        self.device = device
        for name in L.COL_NAMES:
            if name in L.MANDATORY_COLS or optional:
                dt = torch.float64 if name in ("v0", "v1", "v2") else torch.int64
                setattr(self, name, torch.empty(max(self.capacity, 1), dtype=dt, device=device))
            else:
                setattr(self, name, None)

    def struct(self) -> L.DecodedColsStruct:
        s = L.DecodedColsStruct()
        for name in L.COL_NAMES:
            t = getattr(self, name)
            setattr(s, name, t.data_ptr() if t is not None else None)
        return s

    def bytes_per_row(self) -> int:
        return 8 * sum(1 for n in L.COL_NAMES if getattr(self, n) is not None)


@dataclass
class DecodeResult:
    cols: HostCols
    status: np.ndarray      # uint8[n]
    first_row: np.ndarray   # uint32[n+1]
    n_rows: int


# --------------------------------------------------------------------------------------------- context
class Context:
    """One swgpu context (one GPU, one stream)."""

    def __init__(self, device: int = 0, stream: Optional[int] = None):
        self.lib = L.load()
        h = C.c_void_p()
        rc = self.lib.swgpu_create(int(device), C.c_void_p(stream) if stream else None, C.byref(h))
        if rc != 0:
            raise L.SwgpuError(rc, self.lib.swgpu_last_error(None).decode())
        self.h = h
        self.device = int(device)

    def close(self):
        if getattr(self, "h", None):
            self.lib.swgpu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int, allow_capacity: bool = False):
        if rc == 0 or (allow_capacity and rc == L.E_CAPACITY):
            return rc
        raise L.SwgpuError(rc, self.lib.swgpu_last_error(self.h).decode())

    def sync(self):
        self._check(self.lib.swgpu_sync(self.h))

    def set_option(self, name: str, value: int):
        """swgpu_set_option: pick between code paths that give identical results (used by the tests)."""
        self._check(self.lib.swgpu_set_option(self.h, name.encode(), int(value)))

    @property
    def stream(self) -> int:
        return int(self.lib.swgpu_stream(self.h) or 0)

    # ---- zones ----------------------------------------------------------------------------------
    def set_zones(self, vtx_off, lon, lat, cond):
        vtx_off = np.ascontiguousarray(vtx_off, dtype=np.uint32)
        lon = np.ascontiguousarray(lon, dtype=np.float64)
        lat = np.ascontiguousarray(lat, dtype=np.float64)
        cond = np.ascontiguousarray(cond, dtype=np.uint8)
        T = vtx_off.shape[0] - 1
        if T < 0 or cond.shape[0] != T or (T and (lon.shape[0] < vtx_off[-1] or lat.shape[0] < vtx_off[-1])):
            raise ValueError("inconsistent zone arrays")
        self._check(self.lib.swgpu_set_zones(self.h, T, vtx_off.ctypes.data, lon.ctypes.data, lat.ctypes.data,
                                             cond.ctypes.data))
        self.n_tests = T

    def zone_table_bytes(self) -> int:
        v = C.c_uint64()
        self._check(self.lib.swgpu_zone_table_bytes(self.h, C.byref(v)))
        return int(v.value)

    def zone_table_export(self, dst_tensor):
        self._check(self.lib.swgpu_zone_table_export_dev(self.h, dst_tensor.data_ptr(),
                                                         dst_tensor.numel() * dst_tensor.element_size()))

    def zone_table_import(self, src_tensor, nbytes: int):
        self._check(self.lib.swgpu_zone_table_import_dev(self.h, src_tensor.data_ptr(), int(nbytes)))

    # ---- decode (host buffers) --------------------------------------------------------------------
    def _decode(self, fn, data, rec_off, max_rows=None, optional_cols=True) -> DecodeResult:
        data = np.ascontiguousarray(data, dtype=np.uint8)
        rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
        n = rec_off.shape[0] - 1
        if n < 0:
            raise ValueError("rec_off needs n+1 entries")
        if max_rows is None:
            max_rows = max(n, 1)
        cols = HostCols(max_rows, optional=optional_cols)
        status = np.zeros(max(n, 1), dtype=np.uint8)
        first_row = np.zeros(n + 1, dtype=np.uint32)
        out_rows = C.c_uint32()
        s = cols.struct()
        rc = fn(self.h, data.ctypes.data if data.size else None, rec_off.ctypes.data, n, max_rows, C.byref(s),
                status.ctypes.data, first_row.ctypes.data, C.byref(out_rows))
        if rc == L.E_CAPACITY:  # measurement fan-out: retry with the exact row count
            return self._decode(fn, data, rec_off, max_rows=int(out_rows.value), optional_cols=optional_cols)
        self._check(rc)
        cols.n = int(out_rows.value)
        return DecodeResult(cols, status[:n], first_row, cols.n)

    def decode_pb(self, data, rec_off, max_rows=None, optional_cols=True) -> DecodeResult:
        return self._decode(self.lib.swgpu_decode_pb, data, rec_off, max_rows, optional_cols)

    def decode_json(self, data, rec_off, max_rows=None, optional_cols=True) -> DecodeResult:
        return self._decode(self.lib.swgpu_decode_json, data, rec_off, max_rows, optional_cols)

    def decode_enriched(self, data, rec_off, max_rows=None, optional_cols=True):
        """swgpu_decode_enriched: GEnrichedEventPayload records -> (DecodeResult, ids dict of uint64 arrays)"""
        data = np.ascontiguousarray(data, dtype=np.uint8)
        rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
        n = rec_off.shape[0] - 1
        if max_rows is None:
            max_rows = max(n, 1)
        cols = HostCols(max_rows, optional=optional_cols)
        ids = {k: np.zeros(max_rows, np.uint64) for k in L.ID_NAMES}
        si = L.EventIdsStruct()
        for k in L.ID_NAMES:
            setattr(si, k, ids[k].ctypes.data)
        status = np.zeros(max(n, 1), dtype=np.uint8)
        first_row = np.zeros(n + 1, dtype=np.uint32)
        out_rows = C.c_uint32()
        s = cols.struct()
        self._check(self.lib.swgpu_decode_enriched(self.h, data.ctypes.data if data.size else None, rec_off.ctypes.data, n,
                                                   max_rows, C.byref(s), C.byref(si), status.ctypes.data,
                                                   first_row.ctypes.data, C.byref(out_rows)))
        cols.n = int(out_rows.value)
        return DecodeResult(cols, status[:n], first_row, cols.n), {k: v[: cols.n] for k, v in ids.items()}

    # ---- decode (device buffers; async on the context stream) ---------------------------------------
    def decode_pb_dev(self, bytes_t, off_t, n, cols: DeviceCols, status_t, first_row_t, n_rows_t):
        s = cols.struct()
        self._check(self.lib.swgpu_decode_pb_dev(self.h, bytes_t.data_ptr(), off_t.data_ptr(), int(n), cols.capacity,
                                                 C.byref(s), status_t.data_ptr(), first_row_t.data_ptr(),
                                                 n_rows_t.data_ptr()))

    def decode_json_dev(self, bytes_t, off_t, n, cols: DeviceCols, status_t, first_row_t, n_rows_t):
        s = cols.struct()
        self._check(self.lib.swgpu_decode_json_dev(self.h, bytes_t.data_ptr(), off_t.data_ptr(), int(n),
                                                   cols.capacity, C.byref(s), status_t.data_ptr(),
                                                   first_row_t.data_ptr(), n_rows_t.data_ptr()))

    def decode_enriched_dev(self, bytes_t, off_t, n, cols: DeviceCols, status_t, first_row_t, n_rows_t, ids_struct=None):
        """swgpu_decode_enriched_dev (ids_struct: an EventIdsStruct of device pointers, or None = ids not wanted)"""
        s = cols.struct()
        self._check(self.lib.swgpu_decode_enriched_dev(self.h, bytes_t.data_ptr(), off_t.data_ptr(), int(n), cols.capacity,
                                                       C.byref(s), C.byref(ids_struct) if ids_struct is not None else None,
                                                       status_t.data_ptr(), first_row_t.data_ptr(), n_rows_t.data_ptr()))

    # ---- zone test ----------------------------------------------------------------------------------
    def zone_test(self, lon, lat, meta=None, max_pairs=None):
        lon = np.ascontiguousarray(lon, dtype=np.float64)
        lat = np.ascontiguousarray(lat, dtype=np.float64)
        n = lon.shape[0]
        if max_pairs is None:
            max_pairs = max(n, 1024)
        pairs = np.zeros(max(max_pairs, 1), dtype=np.uint64)
        npairs = C.c_uint64()
        mptr = None
        if meta is not None:
            meta = np.ascontiguousarray(meta, dtype=np.uint64)
            mptr = meta.ctypes.data
        rc = self.lib.swgpu_zone_test(self.h, lon.ctypes.data, lat.ctypes.data, mptr, n, max_pairs,
                                      pairs.ctypes.data, C.byref(npairs))
        if rc == L.E_CAPACITY:
            return self.zone_test(lon, lat, meta, max_pairs=int(npairs.value))
        self._check(rc)
        return pairs[: int(npairs.value)]

    def zone_test_dev(self, lon_t, lat_t, meta_t, n, n_rows_t, pairs_t, n_pairs_t):
        self._check(self.lib.swgpu_zone_test_dev(
            self.h, lon_t.data_ptr(), lat_t.data_ptr(), meta_t.data_ptr() if meta_t is not None else None, int(n),
            n_rows_t.data_ptr() if n_rows_t is not None else None, pairs_t.numel(), pairs_t.data_ptr(),
            n_pairs_t.data_ptr()))

    def decode_zone_test_dev(self, fmt, bytes_t, off_t, n, cols, status_t, first_row_t, n_rows_t, pairs_t, n_pairs_t):
        """swgpu_decode_zone_test_dev: decode + zone test on device buffers, one call"""
        s = cols.struct()
        self._check(self.lib.swgpu_decode_zone_test_dev(
            self.h, L.FMT[fmt] if isinstance(fmt, str) else int(fmt), bytes_t.data_ptr(), off_t.data_ptr(), int(n),
            cols.capacity, C.byref(s), status_t.data_ptr(), first_row_t.data_ptr(), n_rows_t.data_ptr(), pairs_t.numel(),
            pairs_t.data_ptr(), n_pairs_t.data_ptr()))

    def dedup_probe_dev(self, bytes_t, off_t, cols, n_rows, n_rows_t, dup_t):
        """swgpu_dedup_probe_dev: per decoded row 0 = new alternate id (or none), 1 = seen before, 2 = ask the store"""
        self._check(self.lib.swgpu_dedup_probe_dev(self.h, bytes_t.data_ptr(), off_t.data_ptr(), cols.meta.data_ptr(),
                                                   cols.alternate_id_span.data_ptr(), int(n_rows),
                                                   n_rows_t.data_ptr() if n_rows_t is not None else None, dup_t.data_ptr()))

    def dedup_probe(self, ids) -> np.ndarray:
        """swgpu_dedup_probe: list of alternate ids (bytes) -> uint8 verdicts (0 new, 1 seen, 2 ask the store)"""
        ids = [bytes(x) for x in ids]
        n = len(ids)
        off = np.zeros(n + 1, dtype=np.uint64)
        np.cumsum([len(x) for x in ids], out=off[1:].view(np.int64)) if n else None
        data = np.frombuffer(b"".join(ids), dtype=np.uint8).copy() if n and off[-1] else np.zeros(1, np.uint8)
        dup = np.zeros(max(n, 1), dtype=np.uint8)
        self._check(self.lib.swgpu_dedup_probe(self.h, data.ctypes.data, off.ctypes.data, n, dup.ctypes.data))
        return dup[:n]

    def dedup_size(self) -> int:
        v = C.c_uint64()
        self._check(self.lib.swgpu_dedup_size(self.h, C.byref(v)))
        return int(v.value)

    def dedup_clear(self):
        self._check(self.lib.swgpu_dedup_clear(self.h))

    def dedup_reserve(self, max_ids: int):
        self._check(self.lib.swgpu_dedup_reserve(self.h, int(max_ids)))

    def encode_inbound_dev(self, fmt, bytes_t, off_t, cols, n_rows, n_rows_t, source_id: bytes, now_ms, out_t, out_off_t,
                           enc_status_t, out_bytes_t):
        """swgpu_encode_inbound_dev: the GInboundEventPayload bytes of every decoded row, back to back in `out_t`"""
        s = cols.struct()
        src = (C.c_uint8 * max(len(source_id), 1)).from_buffer_copy(source_id or b"\0")
        self._check(self.lib.swgpu_encode_inbound_dev(
            self.h, L.FMT[fmt], bytes_t.data_ptr(), off_t.data_ptr(), C.byref(s), int(n_rows),
            n_rows_t.data_ptr() if n_rows_t is not None else None, src, len(source_id), int(now_ms), out_t.data_ptr(),
            out_t.numel(), out_off_t.data_ptr(), enc_status_t.data_ptr(), out_bytes_t.data_ptr()))

    def zone_stats(self) -> dict:
        """swgpu_zone_stats: work counters of the last zone test run with option zone.stats = 1."""
        out = (C.c_uint64 * 4)()
        self._check(self.lib.swgpu_zone_stats(self.h, out))
        return dict(cell_records=int(out[0]), edge_records=int(out[1]), straddle_tests=int(out[2]), locations=int(out[3]))

    def json_stats(self) -> dict:
        """swgpu_json_stats: recogniser hits of the last JSON decode run with option json.stats = 1."""
        out = (C.c_uint64 * 2)()
        self._check(self.lib.swgpu_json_stats(self.h, out))
        return dict(template_records=int(out[0]), second_pass=int(out[1]))

    # ---- fused host-buffer path (chunked, overlapped) -------------------------------------------------
    def decode_zone_test(self, fmt: str, data, rec_off, cols: "HostCols" = None, status=None, first_row=None,
                         pairs=None, chunk_records: int = 0):
        """swgpu_decode_zone_test: decode + zone test of a host-resident batch in one call.
        Returns (cols, status, first_row, n_rows, pairs[:n_pairs]).  Pass buffers from
        :func:`pinned_array` to let copies and kernels overlap."""
        data = np.ascontiguousarray(data, dtype=np.uint8)
        rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
        n = rec_off.shape[0] - 1
        if cols is None:
            cols = HostCols(max(n, 1))
        if status is None:
            status = np.zeros(max(n, 1), dtype=np.uint8)
        if first_row is None:
            first_row = np.zeros(n + 1, dtype=np.uint32)
        if pairs is None:
            pairs = np.zeros(max(n, 1024), dtype=np.uint64)
        out_rows, n_pairs = C.c_uint32(), C.c_uint64()
        s = cols.struct()
        rc = self.lib.swgpu_decode_zone_test(
            self.h, L.FMT[fmt], data.ctypes.data if data.size else None, rec_off.ctypes.data, n,
            cols.capacity, C.byref(s), status.ctypes.data, first_row.ctypes.data, C.byref(out_rows), pairs.shape[0],
            pairs.ctypes.data, C.byref(n_pairs), int(chunk_records))
        if rc == L.E_CAPACITY:
            big = HostCols(max(int(out_rows.value), 1))
            return self.decode_zone_test(fmt, data, rec_off, big, status, first_row,
                                         np.zeros(max(int(n_pairs.value), 1), np.uint64), chunk_records)
        self._check(rc)
        cols.n = int(out_rows.value)
        return cols, status[:n], first_row, cols.n, pairs[: int(n_pairs.value)]

    def zone_test_noalloc_dev(self, lon_t, lat_t, meta_t, n, n_rows_t, pairs_t, n_pairs_t):
        """swgpu_zone_test_noalloc_dev: same as zone_test_dev, but the library allocates nothing
        (fixed in-library scratch, at most NOALLOC_MAX_LOCATIONS locations per call)."""
        self._check(self.lib.swgpu_zone_test_noalloc_dev(
            self.h, lon_t.data_ptr(), lat_t.data_ptr(), meta_t.data_ptr() if meta_t is not None else None, int(n),
            n_rows_t.data_ptr() if n_rows_t is not None else None, pairs_t.numel(), pairs_t.data_ptr(),
            n_pairs_t.data_ptr()))

    # ---- merge ----------------------------------------------------------------------------------------
    def state_reserve(self, max_keys: int, max_key_bytes: int):
        self._check(self.lib.swgpu_state_reserve(self.h, int(max_keys), int(max_key_bytes)))

    def merge(self, data, rec_off, cols: HostCols, seq_base: int, id_hi, id_lo):
        data = np.ascontiguousarray(data, dtype=np.uint8)
        rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
        id_hi = np.ascontiguousarray(id_hi, dtype=np.uint64)
        id_lo = np.ascontiguousarray(id_lo, dtype=np.uint64)
        s = cols.struct()
        self._check(self.lib.swgpu_merge(self.h, data.ctypes.data, rec_off.ctypes.data, rec_off.shape[0] - 1,
                                         C.byref(s), cols.n, int(seq_base), id_hi.ctypes.data, id_lo.ctypes.data))

    def decode_merge(self, fmt: str, data, rec_off, seq_base: int, id_hi, id_lo, status=None) -> int:
        """swgpu_decode_merge: decode a host-resident batch and fold it into the device state in one call
        (ids per RECORD).  Returns the number of rows merged; ``status`` (uint8[n]) is filled when given."""
        data = np.ascontiguousarray(data, dtype=np.uint8)
        rec_off = np.ascontiguousarray(rec_off, dtype=np.uint64)
        n = rec_off.shape[0] - 1
        hi_p = lo_p = None
        if fmt != "enriched":           # (an enriched payload carries its own event id)
            id_hi = np.ascontiguousarray(id_hi, dtype=np.uint64)
            id_lo = np.ascontiguousarray(id_lo, dtype=np.uint64)
            if id_hi.shape[0] < n or id_lo.shape[0] < n:
                raise ValueError("one event id per record is required")
            hi_p, lo_p = id_hi.ctypes.data, id_lo.ctypes.data
        rows = C.c_uint32()
        self._check(self.lib.swgpu_decode_merge(
            self.h, L.FMT[fmt], data.ctypes.data if data.size else None, rec_off.ctypes.data, n,
            int(seq_base), hi_p, lo_p, status.ctypes.data if status is not None else None,
            C.byref(rows)))
        return int(rows.value)

    def decode_merge_dev(self, fmt: str, bytes_t, off_t, n, seq_base, id_hi_t, id_lo_t, status_t, n_rows_t):
        """swgpu_decode_merge_dev: the fused decode + merge kernel on device-resident buffers (async)."""
        self._check(self.lib.swgpu_decode_merge_dev(self.h, L.FMT[fmt], bytes_t.data_ptr(), off_t.data_ptr(),
                                                    int(n), int(seq_base),
                                                    id_hi_t.data_ptr() if id_hi_t is not None else None,
                                                    id_lo_t.data_ptr() if id_lo_t is not None else None,
                                                    status_t.data_ptr(), n_rows_t.data_ptr()))

    def merge_dev(self, bytes_t, off_t, n_records, cols: DeviceCols, n_rows, n_rows_t, seq_base, id_hi_t, id_lo_t):
        s = cols.struct()
        self._check(self.lib.swgpu_merge_dev(self.h, bytes_t.data_ptr(), off_t.data_ptr(), int(n_records), C.byref(s),
                                             int(n_rows), n_rows_t.data_ptr() if n_rows_t is not None else None,
                                             int(seq_base), id_hi_t.data_ptr(), id_lo_t.data_ptr()))

    def state_size(self):
        a, b = C.c_uint64(), C.c_uint64()
        self._check(self.lib.swgpu_state_size(self.h, C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)

    def presence_scan(self, now_ms: int, interval_ms: int, max_devices: int = None):
        """swgpu_presence_scan -> (lastInteractionDate int64[k], [device key bytes], n_found)"""
        if max_devices is None:
            max_devices, kb = self.state_size()
        else:
            kb = max_devices * 64
        last = np.zeros(max(max_devices, 1), np.int64)
        off = np.zeros(max_devices + 1, np.uint64)
        keys = np.zeros(max(kb, 1), np.uint8)
        rep, found = C.c_uint64(), C.c_uint64()
        rc = self.lib.swgpu_presence_scan(self.h, int(now_ms), int(interval_ms), max_devices, keys.shape[0], last.ctypes.data,
                                          off.ctypes.data, keys.ctypes.data, C.byref(rep), C.byref(found))
        self._check(rc, allow_capacity=True)
        k = int(rep.value)
        raw = keys.tobytes()
        return last[:k], [raw[int(off[i]): int(off[i + 1])] for i in range(k)], int(found.value)

    def state_clear(self):
        self._check(self.lib.swgpu_state_clear(self.h))

    def state_snapshot(self, sort: bool = True) -> dict:
        n, kb = self.state_size()
        out = dict(seq=np.zeros(max(n, 1), np.uint64), id_hi=np.zeros(max(n, 1), np.uint64),
                   id_lo=np.zeros(max(n, 1), np.uint64), event_date=np.zeros(max(n, 1), np.int64),
                   value=np.zeros(max(n, 1), np.float64), key_off=np.zeros(n + 1, np.uint64),
                   key_bytes=np.zeros(max(kb, 1), np.uint8))
        s = L.StateRowsStruct()
        for k in out:
            setattr(s, k, out[k].ctypes.data)
        got = C.c_uint64()
        self._check(self.lib.swgpu_state_snapshot(self.h, n, kb, C.byref(s), C.byref(got)))
        n = int(got.value)
        for k in ("seq", "id_hi", "id_lo", "event_date", "value"):
            out[k] = out[k][:n]
        out["key_off"] = out["key_off"][: n + 1]
        out["key_bytes"] = out["key_bytes"][: int(out["key_off"][n])] if n else out["key_bytes"][:0]
        if sort and n:
            kb_ = out["key_bytes"].tobytes()
            ko = out["key_off"]
            keys = [kb_[int(ko[i]): int(ko[i + 1])] for i in range(n)]
            order = sorted(range(n), key=lambda i: keys[i])
            for k in ("seq", "id_hi", "id_lo", "event_date", "value"):
                out[k] = out[k][order]
            lens = np.array([len(keys[i]) for i in order], dtype=np.uint64)
            out["key_off"] = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
            out["key_bytes"] = np.frombuffer(b"".join(keys[i] for i in order), dtype=np.uint8).copy()
        return out

    def state_restore(self, snap: dict):
        n = snap["seq"].shape[0]
        arrs = {k: np.ascontiguousarray(snap[k]) for k in snap}
        s = L.StateRowsStruct()
        for k in arrs:
            setattr(s, k, arrs[k].ctypes.data if arrs[k].size else None)
        self._check(self.lib.swgpu_state_restore(self.h, C.byref(s), n))


class _Pinned:
    """Owner of one swgpu_alloc_pinned block."""

    def __init__(self, nbytes: int):
        self.lib = L.load()
        self.ptr = self.lib.swgpu_alloc_pinned(max(int(nbytes), 1))
        if not self.ptr:
            raise MemoryError(f"swgpu_alloc_pinned({nbytes}) failed")
        self.nbytes = int(nbytes)

    def __del__(self):
        try:
            if self.ptr:
                self.lib.swgpu_free_pinned(self.ptr)
                self.ptr = None
        except Exception:
            pass


def pinned_array(shape, dtype) -> np.ndarray:
    """numpy array in page-locked host memory (swgpu_alloc_pinned), as the Java binding's staging
    segments are.  The block is freed when the array (and its views) are garbage-collected."""
    dt = np.dtype(dtype)
    n = int(np.prod(shape))
    owner = _Pinned(n * dt.itemsize)
    buf = (C.c_uint8 * max(n * dt.itemsize, 1)).from_address(owner.ptr)
    arr = np.frombuffer(buf, dtype=dt, count=n).reshape(shape)
    buf._owner = owner  # the array references buf, buf references the owner: freed with the last view
    return arr


def pinned_cols(capacity: int, names=None) -> "HostCols":
    """HostCols whose (selected) columns live in pinned memory; unselected columns are None."""
    hc = HostCols.__new__(HostCols)
    hc.capacity = int(capacity)
    hc.n = 0
    for name in L.COL_NAMES:
        if names is None or name in names:
            setattr(hc, name, pinned_array((max(capacity, 1),), _COL_DT.get(name, np.uint64)))
        else:
            setattr(hc, name, None)
    return hc


def murmur2(key: bytes) -> int:
    b = np.frombuffer(key, dtype=np.uint8) if key else np.zeros(1, np.uint8)
    return int(L.load().swgpu_murmur2(b.ctypes.data, len(key)))


def partition(key: bytes, n_partitions: int) -> int:
    b = np.frombuffer(key, dtype=np.uint8) if key else np.zeros(1, np.uint8)
    return int(L.load().swgpu_partition(b.ctypes.data, len(key), n_partitions))


# --------------------------------------------------------------------------------------------- mirrors
@dataclass
class DecodedDeviceRequest:
    """ES/DecodedDeviceRequest.java:17-29 with the request POJO flattened into a dict."""
    deviceToken: Optional[str]
    originator: Optional[str]
    request: dict


_KIND_NAMES = {0: "Measurement", 1: "Location", 2: "Alert"}
_LEVELS = ["Info", "Warning", "Error", "Critical"]
_SOURCES = ["Device", "System"]


def _span(rec: bytes, s: int) -> bytes:
    off, ln = s & 0xFFFFFFFF, s >> 32
    return rec[off: off + ln]


def rows_to_requests(res: DecodeResult, data: np.ndarray, rec_off: np.ndarray, i: int, json_mode=False):
    """Rebuild the reference's result list for record i from the columnar rows (what the Java binding
    does when it completes a caller's future)."""
    st = int(res.status[i])
    if st == L.ST_DECODE_EXCEPTION:
        raise EventDecodeException("Unable to decode message.")
    if st == L.ST_RUNTIME_EXCEPTION:
        raise DecoderRuntimeException("NullPointerException")
    if st == L.ST_PASSTHROUGH:
        raise PassThrough()
    rec = bytes(data[int(rec_off[i]): int(rec_off[i + 1])])
    out = []
    c = res.cols
    for r in range(int(res.first_row[i]), int(res.first_row[i + 1])):
        m = int(c.meta[r])
        kind, fl = m & 0xFF, (m >> 8) & 0xFFFF

        def text(b: bytes):
            if json_mode and (fl & L.F_STR_ESCAPED):
                import json as _json
                return _json.loads(b'"' + b + b'"')
            return b.decode("utf-8", errors="replace")
        req = {"eventType": _KIND_NAMES.get((m >> 28) & 7) if (m >> 27) & 1 else None,
               "updateState": bool(fl & L.F_UPDATE_STATE),
               "eventDate": int(c.event_date[r]) if fl & L.F_HAS_DATE else None}
        if kind == L.KIND_MEASUREMENT:
            req["name"] = text(_span(rec, int(c.name_span[r]))) if fl & L.F_HAS_NAME else None
            req["value"] = float(c.v0[r])
        elif kind == L.KIND_LOCATION:
            req["latitude"] = float(c.v0[r]) if fl & L.F_HAS_V0 else None
            req["longitude"] = float(c.v1[r]) if fl & L.F_HAS_V1 else None
            req["elevation"] = float(c.v2[r]) if fl & L.F_HAS_V2 else None
        elif kind == L.KIND_ALERT:
            req["type"] = text(_span(rec, int(c.name_span[r]))) if fl & L.F_HAS_NAME else None
            req["message"] = (text(_span(rec, int(c.message_span[r])))
                              if (fl & L.F_HAS_MESSAGE and c.message_span is not None) else None)
            req["level"] = _LEVELS[(m >> 24) & 3]
            req["source"] = _SOURCES[(m >> 26) & 1]
        tok = None if fl & L.F_NULL_TOKEN else text(_span(rec, int(c.token_span[r])))
        orig = (text(_span(rec, int(c.originator_span[r])))
                if (fl & L.F_HAS_ORIGINATOR and c.originator_span is not None) else None)
        out.append(DecodedDeviceRequest(tok, orig, req))
    return out


class _BatchDecoder:
    _fn = None
    _json = False

    def __init__(self, ctx: Context):
        self.ctx = ctx

    def decode_batch(self, payloads) -> tuple[DecodeResult, np.ndarray, np.ndarray]:
        from .synth import pack_records
        data, rec_off = pack_records([bytes(p) for p in payloads])
        res = getattr(self.ctx, self._fn)(data, rec_off)
        return res, data, rec_off

    def decode(self, payload: bytes, metadata=None):
        """IDeviceEventDecoder<byte[]>.decode(payload, metadata) (ES/spi/IDeviceEventDecoder.java:39):
        list of decoded requests, or EventDecodeException."""
        res, data, rec_off = self.decode_batch([payload])
        return rows_to_requests(res, data, rec_off, 0, json_mode=self._json)


class ProtobufDeviceEventDecoder(_BatchDecoder):
    _fn = "decode_pb"


class JsonDeviceRequestDecoder(_BatchDecoder):
    _fn = "decode_json"
    _json = True


@dataclass
class ZoneTest:
    """RP/processors/geospatial/ZoneTest.java:19-34."""
    zoneToken: str
    condition: str = "Inside"   # ZoneContainment
    alertType: str = ""
    alertLevel: str = "Error"
    alertMessage: str = ""


class ZoneTestRuleProcessor:
    """RP/processors/geospatial/ZoneTestRuleProcessor.java:33-93.  ``zones`` maps zone token ->
    list of (latitude, longitude) bounds, what getDeviceManagement().getZoneByToken returns."""

    def __init__(self, ctx: Context, zone_tests: list[ZoneTest], zones: dict):
        self.ctx = ctx
        self.zone_tests = list(zone_tests)
        vtx_off = [0]
        lon, lat, cond = [], [], []
        for t in self.zone_tests:
            if t.zoneToken not in zones:
                raise KeyError(f"Invalid zone token in ZoneTestRuleProcessor: {t.zoneToken}")
            for (la, lo) in zones[t.zoneToken]:
                lat.append(la)
                lon.append(lo)
            vtx_off.append(len(lon))
            cond.append(0 if t.condition == "Inside" else 1)
        ctx.set_zones(np.array(vtx_off, np.uint32), np.array(lon, np.float64), np.array(lat, np.float64),
                      np.array(cond, np.uint8))

    def on_locations(self, lon, lat) -> np.ndarray:
        """Batch form: (location index << 32 | test index) for every alert the reference fires."""
        return self.ctx.zone_test(lon, lat)

    def onLocation(self, context, location: dict) -> list[dict]:
        """Returns the DeviceAlertCreateRequests the reference would pass to addDeviceAlerts, in order."""
        pairs = self.ctx.zone_test([location["longitude"]], [location["latitude"]])
        out = []
        for p in pairs:
            t = self.zone_tests[int(p) & 0xFFFFFFFF]
            out.append(dict(type=t.alertType, level=t.alertLevel, message=t.alertMessage, updateState=False))
        return out


class DeviceStateProcessingLogic:
    """DS/processing/DeviceStateProcessingLogic.java:41-210, batch-shaped like
    IDeviceStateProcessingLogic.process(List<ConsumerRecord>)."""

    def __init__(self, ctx: Context):
        self.ctx = ctx
        self.seq = 0

    def process(self, data, rec_off, cols: HostCols, id_hi, id_lo):
        self.ctx.merge(data, rec_off, cols, self.seq, id_hi, id_lo)
        self.seq += cols.n

    def process_records(self, fmt: str, data, rec_off, id_hi, id_lo) -> int:
        """process(List<ConsumerRecord>) on raw record payloads: one fused native call (decode + merge)."""
        rows = self.ctx.decode_merge(fmt, data, rec_off, self.seq, id_hi, id_lo)
        self.seq += (len(rec_off) - 1) << L.SEQ_RECORD_SHIFT      # SWGPU_SEQ_STRIDE(n)
        return rows

    def snapshot(self):
        return self.ctx.state_snapshot()
