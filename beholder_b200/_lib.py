"""ctypes binding of libbeholder_gpu.so (include/beholder_gpu.h).

The library is the product: there is no Python or CPU fallback.  Importing this module
fails loudly when the shared object is missing or lacks a declared symbol.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# BEHOLDER_GPU_LIB: another build of the same library (A/B timing of kernel changes); default = the in-tree build
LIB_PATH = os.environ.get("BEHOLDER_GPU_LIB") or os.path.join(_HERE, "libbeholder_gpu.so")

BH_MAX_FIELDS = 16
BH_MAX_ENUM = 64
BH_HIST_MAX = 64
BH_ALL_WINDOWS = 0xFFFFFFFFFFFFFFFF
BH_STATIC_EXPORT_MAX_OWNERS = 1024
BH_NO_BATCH_DEADLINE = 0xFFFFFFFF
BH_COMPACT_BLOCK = 1 << 16

# bh_status
BH_OK = 0
BH_ERR_INVALID = -1
BH_ERR_NO_DEVICE = -2
BH_ERR_CUDA = -3
BH_ERR_NOMEM = -4
BH_ERR_TABLE_FULL = -5
BH_ERR_WINDOW_RING = -6
BH_ERR_TOO_LARGE = -7
BH_ERR_UNSUPPORTED = -8
BH_ERR_KEY_COLLISION = -9
STATUS_NAMES = {
    -9: "BH_ERR_KEY_COLLISION",
    0: "BH_OK", -1: "BH_ERR_INVALID", -2: "BH_ERR_NO_DEVICE", -3: "BH_ERR_CUDA", -4: "BH_ERR_NOMEM",
    -5: "BH_ERR_TABLE_FULL", -6: "BH_ERR_WINDOW_RING", -7: "BH_ERR_TOO_LARGE", -8: "BH_ERR_UNSUPPORTED",
}

# kinds / roles / ops / modes / outcomes
K_STRING, K_BYTES, K_INT32, K_UINT32, K_SINT32, K_BOOL, K_FLOAT, K_DOUBLE = range(1, 9)
R_KEY, R_STATUS, R_VALUE = 1, 2, 4
OP_COUNT, OP_SUM, OP_MIN, OP_MAX, OP_HIST = 1, 2, 4, 8, 16
F_GENERIC_KERNELS, F_NO_NUMA_BIND, F_PARTITIONED, F_NO_PARTITION, F_VERIFY_KEYS = 1, 2, 4, 8, 16   # bh_config.flags
MODE_REFERENCE, MODE_WINDOWED = 0, 1
EV_OK, EV_DROP_DECODE, EV_DROP_ENUM, EV_DROP_NAN = 0, 1, 2, 3


class BhField(C.Structure):
    _fields_ = [("number", C.c_uint32), ("kind", C.c_uint32), ("role", C.c_uint32), ("reserved", C.c_uint32)]


class BhConfig(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("device", C.c_int32), ("mode", C.c_uint32), ("ops", C.c_uint32),
        ("window_ms", C.c_uint64), ("window_ring", C.c_uint32), ("hist_buckets", C.c_uint32),
        ("hist_e0", C.c_int32), ("n_fields", C.c_uint32), ("fields", BhField * BH_MAX_FIELDS),
        ("n_enum", C.c_uint32), ("enum_numbers", C.c_int32 * BH_MAX_ENUM),
        ("table_capacity", C.c_uint64), ("hash_seed", C.c_uint64), ("slab_bytes", C.c_uint64),
        ("slab_events", C.c_uint32), ("n_slabs", C.c_uint32), ("message_bytes_hint", C.c_uint32), ("flags", C.c_uint32),
        ("max_batch_latency_ms", C.c_uint32), ("part_region_tuples", C.c_uint32), ("debug_hash_mask", C.c_uint64),
    ]


class BhRecord(C.Structure):
    _fields_ = [("key_hash", C.c_uint64), ("window_id", C.c_uint64), ("count", C.c_uint64),
                ("sum", C.c_double), ("min", C.c_double), ("max", C.c_double)]


class BhCounters(C.Structure):
    _fields_ = [("events_ok", C.c_uint64), ("dropped_decode", C.c_uint64), ("dropped_enum", C.c_uint64),
                ("dropped_nan", C.c_uint64), ("dropped_table_full", C.c_uint64), ("dropped_window", C.c_uint64),
                ("smem_spills", C.c_uint64), ("slow_decodes", C.c_uint64), ("key_collisions", C.c_uint64)]


class BhSynthConfig(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("n_status", C.c_uint32), ("seed", C.c_uint64), ("n_keys", C.c_uint64),
        ("t0_ms", C.c_uint64), ("rate_per_ms", C.c_uint32), ("jitter_ms", C.c_uint32),
        ("payload_bytes", C.c_uint32), ("value_f64", C.c_uint32), ("value_exp_lo", C.c_int32),
        ("value_exp_span", C.c_uint32), ("value_signed", C.c_uint32), ("malformed_ppm", C.c_uint32),
        ("unknown_enum_ppm", C.c_uint32), ("zipf", C.c_uint32),
    ]


COMMIT_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_uint64)   # bh_commit_cb
_vp, _u8p, _u32p, _u64p, _f64p, _i32p = (C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p)

# every symbol include/beholder_gpu.h declares: name -> (restype, argtypes)
PROTOTYPES = {
    "bh_create": (C.c_int, [C.POINTER(BhConfig), C.POINTER(C.c_void_p)]),
    "bh_destroy": (None, [C.c_void_p]),
    "bh_last_error": (C.c_char_p, [C.c_void_p]),
    "bh_device_count": (C.c_int, []),
    "bh_push": (C.c_int, [C.c_void_p, _u8p, C.c_uint32, C.c_uint64]),
    "bh_push_batch": (C.c_int, [C.c_void_p, _u8p, _u32p, _u64p, C.c_uint32]),
    "bh_push_batch_pinned": (C.c_int, [C.c_void_p, _u8p, _u32p, _u64p, C.c_uint32]),
    "bh_push_batch_compact": (C.c_int, [C.c_void_p, _u8p, C.c_void_p, _u64p, _u64p, C.c_void_p, C.c_uint32]),
    "bh_push_columns": (C.c_int, [C.c_void_p, _u64p, _u64p, _f64p, C.c_uint64]),
    "bh_flush": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "bh_poll_committed": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "bh_submit": (C.c_int, [C.c_void_p]),
    "bh_set_commit_callback": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "bh_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "bh_host_free": (C.c_int, [C.c_void_p]),
    "bh_host_register": (C.c_int, [C.c_void_p, C.c_size_t]),
    "bh_host_unregister": (C.c_int, [C.c_void_p]),
    "bh_snapshot": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "bh_snapshot_hist": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "bh_list_windows": (C.c_int, [C.c_void_p, _u64p, C.c_uint32, C.POINTER(C.c_uint32)]),
    "bh_close_window": (C.c_int, [C.c_void_p, C.c_uint64]),
    "bh_restore": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64]),
    "bh_get_counters": (C.c_int, [C.c_void_p, C.POINTER(BhCounters)]),
    "bh_process_device": (C.c_int, [C.c_void_p, _u8p, _u32p, _u64p, C.c_uint32, C.c_void_p]),
    "bh_process_columns_device": (C.c_int, [C.c_void_p, _u64p, _u64p, _f64p, C.c_uint64, C.c_void_p]),
    "bh_decode_device": (C.c_int, [C.c_void_p, _u8p, _u32p, C.c_uint32, _u64p, _f64p, _i32p, _u8p, C.c_void_p]),
    "bh_export_window_device": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, _u64p, C.c_void_p]),
    "bh_export_window_static": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, _u64p, C.c_void_p]),
    "bh_reset_device": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bh_close_window_device": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p]),
    "bh_drain_device": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bh_merge_records_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]),
    "bh_sync": (C.c_int, [C.c_void_p]),
    "bh_synth_event_size": (C.c_uint32, [C.POINTER(BhSynthConfig), C.c_uint64, _u64p]),
    "bh_synth_event_host": (C.c_uint32, [C.POINTER(BhSynthConfig), C.c_uint64, _u64p, _u8p, C.POINTER(C.c_uint64)]),
    "bh_synth_generate_device": (C.c_int, [C.c_int, C.POINTER(BhSynthConfig), C.c_uint64, C.c_uint32, _u64p, _u8p,
                                           C.c_uint64, _u32p, _u64p, C.POINTER(C.c_uint64), C.c_void_p]),
    "bh_synth_columns_device": (C.c_int, [C.c_int, C.POINTER(BhSynthConfig), C.c_uint64, C.c_uint64, C.c_uint64,
                                          _u64p, _u64p, _u64p, _f64p, C.c_void_p]),
}



class BhAmqpDelivery(C.Structure):
    _fields_ = [("delivery_tag", C.c_uint64), ("body_bytes", C.c_uint32), ("channel", C.c_uint16), ("redelivered", C.c_uint16)]


class BhAmqpStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("frames", "deliveries", "body_bytes", "heartbeats", "other_methods",
                                          "skipped_content", "filtered")]


PROTOTYPES.update({
    "bh_export_window_peers": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, _u64p, _u64p, C.c_uint32, C.c_uint64, _u64p,
                                         C.c_void_p]),
    "bh_amqp_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_char_p, C.c_uint64]),
    "bh_amqp_destroy": (None, [C.c_void_p]),
    "bh_amqp_feed": (C.c_int, [C.c_void_p, _u8p, C.c_uint64, _u8p, C.c_uint64, _u32p, C.POINTER(BhAmqpDelivery), C.c_uint32,
                               C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]),
    "bh_amqp_get_stats": (C.c_int, [C.c_void_p, C.POINTER(BhAmqpStats)]),
    "bh_amqp_last_error": (C.c_char_p, [C.c_void_p]),
})

PROTOTYPES.update({
    "bh_merge_blocks_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, _u64p, C.c_void_p]),
    "bh_stream": (C.c_void_p, [C.c_void_p]),
    "bh_stream_wait": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bh_create_group": (C.c_int, [C.POINTER(BhConfig), C.POINTER(BhConfig), C.POINTER(C.c_int), C.c_uint32, C.c_uint64, C.POINTER(C.c_void_p)]),
    "bh_group_destroy": (None, [C.c_void_p]),
    "bh_group_size": (C.c_uint32, [C.c_void_p]),
    "bh_group_local": (C.c_void_p, [C.c_void_p, C.c_uint32]),
    "bh_group_owned": (C.c_void_p, [C.c_void_p, C.c_uint32]),
    "bh_group_member": (C.c_void_p, [C.c_void_p, C.c_uint32]),
    "bh_group_close_window": (C.c_int, [C.c_void_p, C.c_uint64]),
    "bh_group_sync": (C.c_int, [C.c_void_p]),
    "bh_group_last_error": (C.c_char_p, []),
    "bh_member_create": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, C.POINTER(C.c_void_p)]),
    "bh_member_destroy": (None, [C.c_void_p]),
    "bh_member_ipc_handle": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bh_member_connect_ipc": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bh_member_connect_ptrs": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "bh_member_area": (C.c_void_p, [C.c_void_p]),
    "bh_member_close_window": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p]),
    "bh_member_check": (C.c_int, [C.c_void_p, _u64p]),
})

# test hooks that are not part of the public header
_EXTRA = {"bh_debug_div64": (C.c_uint64, [C.c_uint64, C.c_uint64]),
          "bh_debug_push_each": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32])}


class BeholderError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"{STATUS_NAMES.get(code, code)}: {msg}")
        self.code = code


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C beholder_b200/csrc`). beholder_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in {**PROTOTYPES, **_EXTRA}.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:  # pragma: no cover
            raise ImportError(f"{LIB_PATH} does not export {name}; rebuild it") from e
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(rc: int, ctx=None) -> int:
    if rc != BH_OK:
        msg = lib.bh_last_error(ctx)
        raise BeholderError(rc, msg.decode() if msg else "")
    return rc
