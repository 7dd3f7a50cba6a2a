"""ctypes binding of libswgpu.so (include/swgpu.h).  There is no CPU implementation: if the
library is missing, or there is no sm_100a device, every call fails loudly."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "libswgpu.so")

COL_NAMES = ["meta", "token_hash", "name_hash", "event_date", "v0", "v1", "v2", "token_span",
             "name_span", "originator_span", "message_span", "metadata_span", "alternate_id_span"]
MANDATORY_COLS = COL_NAMES[:9]

OK, E_INVALID, E_CUDA, E_NOMEM, E_CAPACITY, E_STATE = 0, -1, -2, -3, -4, -5
ST_OK, ST_DECODE_EXCEPTION, ST_RUNTIME_EXCEPTION, ST_PASSTHROUGH = 0, 1, 2, 3
KIND_MEASUREMENT, KIND_LOCATION, KIND_ALERT = 0, 1, 2
NOALLOC_MAX_LOCATIONS = 1 << 24

F_UPDATE_STATE = 1 << 0
F_HAS_DATE = 1 << 1
F_HAS_ORIGINATOR = 1 << 2
F_HAS_METADATA = 1 << 3
F_HAS_ALT_ID = 1 << 4
F_NULL_TOKEN = 1 << 5
F_HAS_V0 = 1 << 6
F_HAS_V1 = 1 << 7
F_HAS_V2 = 1 << 8
F_STR_ESCAPED = 1 << 9
F_HAS_NAME = 1 << 10
F_HAS_MESSAGE = 1 << 11
F_NAME_COERCED = 1 << 12
F_NULL_REQUEST = 1 << 13
SEQ_RECORD_SHIFT = 12


class SwgpuError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"swgpu error {code}: {msg}")
        self.code = code


class DecodedColsStruct(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in COL_NAMES]


ID_NAMES = ["event_hi", "event_lo", "assignment_hi", "assignment_lo", "device_hi", "device_lo"]


class EventIdsStruct(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ID_NAMES]


class StateRowsStruct(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("seq", "id_hi", "id_lo", "event_date", "value", "key_off", "key_bytes")]


_lib = None

# name -> (restype, argtypes); the full list of entry points include/swgpu.h declares
_P, _U32, _U64, _I32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int32
_COLS = C.POINTER(DecodedColsStruct)
_ROWS = C.POINTER(StateRowsStruct)
FMT = {"pb": 0, "json": 1, "enriched": 2}
SIGNATURES = {
    "swgpu_create": (_I32, [_I32, _P, C.POINTER(_P)]),
    "swgpu_destroy": (None, [_P]),
    "swgpu_last_error": (C.c_char_p, [_P]),
    "swgpu_abi_version": (_I32, []),
    "swgpu_stream": (_P, [_P]),
    "swgpu_sync": (_I32, [_P]),
    "swgpu_set_option": (_I32, [_P, C.c_char_p, C.c_int64]),
    "swgpu_alloc_pinned": (_P, [C.c_size_t]),
    "swgpu_free_pinned": (None, [_P]),
    "swgpu_set_zones": (_I32, [_P, _U32, _P, _P, _P, _P]),
    "swgpu_zone_table_bytes": (_I32, [_P, C.POINTER(_U64)]),
    "swgpu_zone_table_export_dev": (_I32, [_P, _P, _U64]),
    "swgpu_zone_table_import_dev": (_I32, [_P, _P, _U64]),
    "swgpu_decode_pb": (_I32, [_P, _P, _P, _U32, _U32, _COLS, _P, _P, _P]),
    "swgpu_decode_json": (_I32, [_P, _P, _P, _U32, _U32, _COLS, _P, _P, _P]),
    "swgpu_decode_pb_dev": (_I32, [_P, _P, _P, _U32, _U32, _COLS, _P, _P, _P]),
    "swgpu_decode_json_dev": (_I32, [_P, _P, _P, _U32, _U32, _COLS, _P, _P, _P]),
    "swgpu_decode_enriched": (_I32, [_P, _P, _P, _U32, _U32, _COLS, C.POINTER(EventIdsStruct), _P, _P, _P]),
    "swgpu_decode_enriched_dev": (_I32, [_P, _P, _P, _U32, _U32, _COLS, C.POINTER(EventIdsStruct), _P, _P, _P]),
    "swgpu_zone_test": (_I32, [_P, _P, _P, _P, _U32, _U64, _P, _P]),
    "swgpu_zone_test_dev": (_I32, [_P, _P, _P, _P, _U32, _P, _U64, _P, _P]),
    "swgpu_zone_stats": (_I32, [_P, _P]),
    "swgpu_json_stats": (_I32, [_P, _P]),
    "swgpu_decode_zone_test": (_I32, [_P, _I32, _P, _P, _U32, _U32, _COLS, _P, _P, _P, _U64, _P, _P, _U32]),
    "swgpu_dedup_probe_dev": (_I32, [_P, _P, _P, _P, _P, _U32, _P, _P]),
    "swgpu_dedup_probe": (_I32, [_P, _P, _P, _U32, _P]),
    "swgpu_dedup_reserve": (_I32, [_P, _U64]),
    "swgpu_dedup_size": (_I32, [_P, _P]),
    "swgpu_dedup_clear": (_I32, [_P]),
    "swgpu_encode_inbound_dev": (_I32, [_P, _I32, _P, _P, _COLS, _U32, _P, _P, _U32, C.c_int64, _P, _U64, _P, _P, _P]),
    "swgpu_decode_zone_test_dev": (_I32, [_P, _I32, _P, _P, _U32, _U32, _COLS, _P, _P, _P, _U64, _P, _P]),
    "swgpu_reserve_batch": (_I32, [_P, _U32, _U64]),
    "swgpu_submit_decode_zone_test": (_I32, [_P, _I32, _P, _P, _U32, _U32, _COLS, _P, _P, _P, _U64, _P, _P, _U32, C.POINTER(_U64)]),
    "swgpu_wait": (_I32, [_P, _U64]),
    "swgpu_zone_test_noalloc_dev": (_I32, [_P, _P, _P, _P, _U32, _P, _U64, _P, _P]),
    "swgpu_state_reserve": (_I32, [_P, _U64, _U64]),
    "swgpu_merge": (_I32, [_P, _P, _P, _U32, _COLS, _U32, _U64, _P, _P]),
    "swgpu_decode_merge": (_I32, [_P, _I32, _P, _P, _U32, _U64, _P, _P, _P, C.POINTER(_U32)]),
    "swgpu_decode_merge_dev": (_I32, [_P, _I32, _P, _P, _U32, _U64, _P, _P, _P, _P]),
    "swgpu_merge_dev": (_I32, [_P, _P, _P, _U32, _COLS, _U32, _P, _U64, _P, _P]),
    "swgpu_state_size": (_I32, [_P, C.POINTER(_U64), C.POINTER(_U64)]),
    "swgpu_presence_scan": (_I32, [_P, C.c_int64, C.c_int64, _U64, _U64, _P, _P, _P, C.POINTER(_U64), C.POINTER(_U64)]),
    "swgpu_state_snapshot": (_I32, [_P, _U64, _U64, _ROWS, C.POINTER(_U64)]),
    "swgpu_state_restore": (_I32, [_P, _ROWS, _U64]),
    "swgpu_state_clear": (_I32, [_P]),
    "swgpu_murmur2": (_U32, [_P, _U32]),
    "swgpu_parse_double": (_I32, [C.c_char_p, _U32, C.POINTER(C.c_double)]),
    "swgpu_partition": (_U32, [_P, _U32, _U32]),
}


def load() -> C.CDLL:
    """Load libswgpu.so; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise RuntimeError(
            f"{SO_PATH} is missing: build it with `python -m sitewhere_b200.build` "
            "(nvcc, sm_100a).  sitewhere_b200 has no CPU path.")
    lib = C.CDLL(SO_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.swgpu_abi_version() != 2:
        raise RuntimeError("libswgpu.so ABI version mismatch")
    _lib = lib
    return lib
