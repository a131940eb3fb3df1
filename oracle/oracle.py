"""ctypes binding of the CPU oracle (oracle/beholder_oracle.c).

TEST INFRASTRUCTURE ONLY — imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Never imported by beholder_b200.  PARITY UNPINNED
(see beholder_oracle.h): the reference has no tests/golden vectors and cannot run here.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbeholder_oracle.so")

MAX_FIELDS, MAX_ENUM, MAX_PARTS, HIST_MAX = 16, 64, 3, 64
EV_OK, EV_DROP_DECODE, EV_DROP_ENUM, EV_DROP_NAN = 0, 1, 2, 3


class Field(C.Structure):
    _fields_ = [("number", C.c_uint32), ("kind", C.c_uint32), ("role", C.c_uint32), ("reserved", C.c_uint32)]


class SchemaC(C.Structure):
    _fields_ = [("n_fields", C.c_uint32), ("fields", Field * MAX_FIELDS), ("n_enum", C.c_uint32),
                ("enum_number", C.c_int32 * MAX_ENUM), ("enum_name", (C.c_char * 32) * MAX_ENUM)]


class Msg(C.Structure):
    _fields_ = [("outcome", C.c_int32), ("status", C.c_int32), ("value", C.c_double), ("n_parts", C.c_uint32),
                ("part_is_len", C.c_uint32 * MAX_PARTS), ("part_field", C.c_uint32 * MAX_PARTS),
                ("part_pos", C.c_uint32 * MAX_PARTS), ("part_len", C.c_uint32 * MAX_PARTS),
                ("part_val", C.c_uint32 * MAX_PARTS)]


class AggConfig(C.Structure):
    _fields_ = [("ops", C.c_uint32), ("hist_buckets", C.c_uint32), ("hist_e0", C.c_int32),
                ("reference_key", C.c_uint32), ("window_ms", C.c_uint64), ("hash_seed", C.c_uint64)]


RECORD_DTYPE = np.dtype([("key_hash", "<u8"), ("window_id", "<u8"), ("count", "<u8"),
                         ("sum", "<f8"), ("min", "<f8"), ("max", "<f8")])


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "beholder_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return LIB_PATH


def _load():
    build()
    lib = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    lib.bho_decode.argtypes = [C.POINTER(SchemaC), vp, C.c_uint32, C.POINTER(Msg)]
    lib.bho_handler_new.restype = vp
    lib.bho_handler_new.argtypes = [C.POINTER(SchemaC), C.c_int]
    lib.bho_handler_free.argtypes = [vp]
    lib.bho_handle_progress.argtypes = [vp, vp, C.c_uint32]
    lib.bho_handler_count.restype = C.c_double
    lib.bho_handler_count.argtypes = [vp, C.c_char_p]
    lib.bho_handler_count_at.restype = C.c_double
    lib.bho_handler_count_at.argtypes = [vp, C.c_uint32]
    lib.bho_handler_acks.restype = C.c_uint64
    lib.bho_handler_acks.argtypes = [vp]
    lib.bho_handler_dropped.restype = C.c_uint64
    lib.bho_handler_dropped.argtypes = [vp, C.c_int]
    lib.bho_handler_metrics_text.restype = C.c_size_t
    lib.bho_handler_metrics_text.argtypes = [vp, C.c_char_p, C.c_size_t]
    lib.bho_agg_new.restype = vp
    lib.bho_agg_new.argtypes = [C.POINTER(SchemaC), C.POINTER(AggConfig)]
    lib.bho_agg_free.argtypes = [vp]
    lib.bho_agg_push_raw.argtypes = [vp, vp, vp, vp, C.c_uint64]
    lib.bho_agg_push_columns.argtypes = [vp, vp, vp, vp, C.c_uint64]
    lib.bho_agg_merge.argtypes = [vp, vp]
    lib.bho_agg_size.restype = C.c_uint64
    lib.bho_agg_size.argtypes = [vp]
    lib.bho_agg_snapshot.restype = C.c_uint64
    lib.bho_agg_snapshot.argtypes = [vp, vp, vp, C.c_uint64]
    lib.bho_agg_dropped.restype = C.c_uint64
    lib.bho_agg_dropped.argtypes = [vp, C.c_int]
    lib.bho_agg_hash_collisions.restype = C.c_uint64
    lib.bho_agg_hash_collisions.argtypes = [vp]
    lib.bho_key_hash.restype = C.c_uint64
    lib.bho_key_hash.argtypes = [C.POINTER(Msg), vp, C.c_uint64, C.c_int]
    lib.bho_hash_bytes.restype = C.c_uint64
    lib.bho_hash_bytes.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint64]
    lib.bho_hist_bucket.restype = C.c_uint32
    lib.bho_hist_bucket.argtypes = [C.c_double, C.c_uint32, C.c_int32]
    lib.bho_f64_order.restype = C.c_uint64
    lib.bho_f64_order.argtypes = [C.c_double]
    lib.bho_decode_columns.argtypes = [C.POINTER(SchemaC), C.POINTER(AggConfig), vp, vp, C.c_uint64, vp, vp, vp, vp]
    lib.bho_agg_run_threads.restype = vp
    lib.bho_agg_run_threads.argtypes = [C.POINTER(SchemaC), C.POINTER(AggConfig), vp, vp, vp, C.c_uint64, C.c_int]
    lib.bho_handler_run_threads.argtypes = [C.POINTER(SchemaC), C.c_int, vp, vp, C.c_uint64, C.c_int, vp, vp]
    return lib


lib = _load()


def schema_c(schema) -> SchemaC:
    """from a beholder_b200.schema.Schema-like object (fields with number/kind/role; enum dict)"""
    s = SchemaC()
    s.n_fields = len(schema.fields)
    for i, f in enumerate(schema.fields):
        s.fields[i].number, s.fields[i].kind, s.fields[i].role = f.number, f.kind, f.role
    items = sorted(schema.enum.items(), key=lambda kv: kv[1])
    s.n_enum = len(items)
    for i, (name, num) in enumerate(items):
        s.enum_number[i] = num
        s.enum_name[i].value = name.encode()
    return s


def _p(a):
    if a is None:
        return None
    if isinstance(a, (bytes, bytearray)):
        return C.cast(C.c_char_p(bytes(a)), C.c_void_p)
    return a.ctypes.data


def decode(schema_c_, payload: bytes) -> Msg:
    m = Msg()
    buf = np.frombuffer(payload, dtype=np.uint8) if len(payload) else np.zeros(1, np.uint8)
    lib.bho_decode(C.byref(schema_c_), buf.ctypes.data, len(payload), C.byref(m))
    return m


class Handler:
    """the reference handler of index.js:127-155 (I/O stubbed)"""

    def __init__(self, schema, log_mode: int = 0):
        self.schema = schema
        self.sc = schema_c(schema)
        self.h = lib.bho_handler_new(C.byref(self.sc), log_mode)

    def handle(self, payload: bytes) -> int:
        buf = np.frombuffer(payload, dtype=np.uint8) if len(payload) else np.zeros(1, np.uint8)
        return lib.bho_handle_progress(self.h, buf.ctypes.data, len(payload))

    def handle_slab(self, slab: np.ndarray, offsets: np.ndarray) -> None:
        base = slab.ctypes.data
        for i in range(len(offsets) - 1):
            lib.bho_handle_progress(self.h, base + int(offsets[i]), int(offsets[i + 1] - offsets[i]))

    def counts(self) -> dict:
        out = {}
        for name in self.schema.enum:
            v = lib.bho_handler_count(self.h, name.lower().encode())
            if v:
                out[name.lower()] = int(v)
        return out

    def acks(self) -> int:
        return lib.bho_handler_acks(self.h)

    def dropped(self) -> dict:
        return {"decode": lib.bho_handler_dropped(self.h, EV_DROP_DECODE), "enum": lib.bho_handler_dropped(self.h, EV_DROP_ENUM)}

    def metrics_text(self) -> str:
        buf = C.create_string_buffer(1 << 16)
        n = lib.bho_handler_metrics_text(self.h, buf, len(buf))
        return buf.raw[:n].decode()

    def __del__(self):
        if getattr(self, "h", None):
            lib.bho_handler_free(self.h)
            self.h = None


def run_handler_threads(schema, slab, offsets, n_threads: int, log_mode: int = 0):
    sc = schema_c(schema)
    counts = np.zeros(MAX_ENUM, dtype=np.float64)
    dropped = np.zeros(4, dtype=np.uint64)
    lib.bho_handler_run_threads(C.byref(sc), log_mode, slab.ctypes.data, offsets.ctypes.data, len(offsets) - 1,
                                n_threads, counts.ctypes.data, dropped.ctypes.data)
    items = sorted(schema.enum.items(), key=lambda kv: kv[1])
    return {name.lower(): int(counts[i]) for i, (name, _) in enumerate(items) if counts[i]}, dropped


class Agg:
    """extension aggregator"""

    def __init__(self, schema, ops: int, window_ms: int, hash_seed: int, hist_buckets: int = 64, hist_e0: int = 991,
                 reference_key: bool = False, _handle=None):
        self.sc = schema_c(schema) if schema is not None else SchemaC()
        self.cfg = AggConfig(ops, hist_buckets, hist_e0, int(reference_key), window_ms, hash_seed)
        self.hist_buckets = hist_buckets
        self.a = _handle or lib.bho_agg_new(C.byref(self.sc), C.byref(self.cfg))

    def push_raw(self, slab, offsets, ts):
        lib.bho_agg_push_raw(self.a, slab.ctypes.data, offsets.ctypes.data, ts.ctypes.data if ts is not None else None,
                             len(offsets) - 1)

    def push_columns(self, kh, ts, val):
        lib.bho_agg_push_columns(self.a, kh.ctypes.data, ts.ctypes.data, val.ctypes.data if val is not None else None, len(kh))

    def merge(self, other: "Agg"):
        lib.bho_agg_merge(self.a, other.a)

    def snapshot(self, with_hist: bool = False):
        n = lib.bho_agg_size(self.a)
        out = np.zeros(n, dtype=RECORD_DTYPE)
        hist = np.zeros((n, self.hist_buckets), dtype=np.uint32) if with_hist else None
        lib.bho_agg_snapshot(self.a, out.ctypes.data, hist.ctypes.data if with_hist else None, n)
        return (out, hist) if with_hist else out

    def dropped(self) -> dict:
        return {"decode": lib.bho_agg_dropped(self.a, EV_DROP_DECODE), "enum": lib.bho_agg_dropped(self.a, EV_DROP_ENUM),
                "nan": lib.bho_agg_dropped(self.a, EV_DROP_NAN)}

    def hash_collisions(self) -> int:
        return lib.bho_agg_hash_collisions(self.a)

    @classmethod
    def run_threads(cls, schema, ops, window_ms, hash_seed, slab, offsets, ts, n_threads, hist_buckets=64, hist_e0=991,
                    reference_key=False):
        self = cls.__new__(cls)
        self.sc = schema_c(schema)
        self.cfg = AggConfig(ops, hist_buckets, hist_e0, int(reference_key), window_ms, hash_seed)
        self.hist_buckets = hist_buckets
        self.a = lib.bho_agg_run_threads(C.byref(self.sc), C.byref(self.cfg), slab.ctypes.data, offsets.ctypes.data,
                                         ts.ctypes.data if ts is not None else None, len(offsets) - 1, n_threads)
        return self

    def __del__(self):
        if getattr(self, "a", None):
            lib.bho_agg_free(self.a)
            self.a = None


def decode_columns(schema, ops, hash_seed, slab, offsets, reference_key=False):
    sc = schema_c(schema)
    cfg = AggConfig(ops, 64, 991, int(reference_key), 0, hash_seed)
    n = len(offsets) - 1
    kh = np.zeros(n, np.uint64); val = np.zeros(n, np.float64); st = np.zeros(n, np.int32); oc = np.zeros(n, np.uint8)
    lib.bho_decode_columns(C.byref(sc), C.byref(cfg), slab.ctypes.data, offsets.ctypes.data, n,
                           kh.ctypes.data, val.ctypes.data, st.ctypes.data, oc.ctypes.data)
    return kh, val, st, oc
