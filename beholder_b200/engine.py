"""Aggregator: thin object wrapper over the C-ABI context (bh_ctx).

All compute happens in libbeholder_gpu.so on the GPU.  numpy is used only to hold host
buffers; torch (optional) only to hold device buffers handed to the *_device entry points.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _lib as L
from .schema import Schema, fill_config, telemetry_progress

RECORD_DTYPE = np.dtype([("key_hash", "<u8"), ("window_id", "<u8"), ("count", "<u8"),
                         ("sum", "<f8"), ("min", "<f8"), ("max", "<f8")])
assert RECORD_DTYPE.itemsize == C.sizeof(L.BhRecord)

DEFAULT_HIST_E0 = 991   # bucket 32 <-> [1, 2)


def _ptr(a) -> Optional[int]:
    """address of a numpy array / torch tensor / int / None"""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    raise TypeError(type(a))


class Aggregator:
    """One GPU's group-by state: replaces the prom-client Counter of index.js:30-34 (REFERENCE
    mode) or holds the windowed count/sum/min/max/histogram tables (WINDOWED mode)."""

    def __init__(self, schema: Optional[Schema] = None, *, mode: int = L.MODE_WINDOWED,
                 key: Sequence[str] = ("mediaId", "status"), ops: int = L.OP_COUNT | L.OP_SUM | L.OP_MIN | L.OP_MAX,
                 window_ms: int = 10_000, window_ring: int = 0, table_capacity: int = 0,
                 hist_buckets: int = 64, hist_e0: int = DEFAULT_HIST_E0, hash_seed: int = 0xBE401DE5,
                 device: int = 0, slab_bytes: int = 0, slab_events: int = 0, n_slabs: int = 0,
                 enum_drop: bool = True, message_bytes_hint: int = 0, generic_kernels: bool = False,
                 max_batch_latency_ms: int = 0, flags: int = 0, part_region_tuples: int = 0, verify_keys: bool = False,
                 debug_hash_mask: int = 0, _handle=None):
        schema = schema or telemetry_progress()
        if mode == L.MODE_WINDOWED:
            schema = schema.with_roles(key=key)
        self.schema = schema
        cfg = L.BhConfig()
        cfg.struct_size = C.sizeof(cfg)
        cfg.device = device
        cfg.mode = mode
        cfg.ops = ops
        cfg.window_ms = window_ms if mode == L.MODE_WINDOWED else 0
        cfg.window_ring = window_ring
        cfg.hist_buckets = hist_buckets
        cfg.hist_e0 = hist_e0
        cfg.table_capacity = table_capacity
        cfg.hash_seed = hash_seed
        cfg.slab_bytes = slab_bytes
        cfg.slab_events = slab_events
        cfg.n_slabs = n_slabs
        cfg.message_bytes_hint = message_bytes_hint
        cfg.flags = (L.F_GENERIC_KERNELS if generic_kernels else 0) | (L.F_VERIFY_KEYS if verify_keys else 0) | flags
        cfg.debug_hash_mask = debug_hash_mask
        cfg.max_batch_latency_ms = max_batch_latency_ms
        cfg.part_region_tuples = part_region_tuples
        self._commit_cb = None
        fill_config(cfg, schema)
        if not enum_drop:
            cfg.n_enum = 0
        self.cfg = cfg
        self.mode = mode
        self.ops = ops | L.OP_COUNT if mode == L.MODE_WINDOWED else L.OP_COUNT
        self.hist_buckets = hist_buckets or 64
        self.device = device
        self._owns = _handle is None            # a context that belongs to a bh_group is destroyed by the group
        if _handle is not None:
            self._ctx = C.c_void_p(_handle)
        elif _handle is None and not getattr(self, "_config_only", False):
            self._ctx = C.c_void_p()
            L.check(L.lib.bh_create(C.byref(cfg), C.byref(self._ctx)))

    # ---- lifecycle ----
    @classmethod
    def config(cls, *args, **kw) -> "L.BhConfig":
        """the bh_config these arguments describe, without creating a context (for bh_create_group)"""
        self = cls.__new__(cls)
        self._config_only = True
        self._ctx = C.c_void_p()
        cls.__init__(self, *args, **kw)
        return self.cfg

    def close(self) -> None:
        if self._ctx and not getattr(self, "_owns", True):
            self._ctx = C.c_void_p()
            return
        if self._ctx:
            L.lib.bh_destroy(self._ctx)       # stops the completion thread before anything is freed
            self._ctx = C.c_void_p()
            self._commit_cb = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc: int) -> int:
        return L.check(rc, self._ctx)

    # ---- ingest ----
    def push(self, payload: bytes, ts_ms: int = 0) -> None:
        buf = (C.c_char * len(payload)).from_buffer_copy(payload) if payload else None
        self._ck(L.lib.bh_push(self._ctx, C.addressof(buf) if buf is not None else None, len(payload), ts_ms))

    def push_batch(self, slab, offsets, ts_ms=None, *, pinned: bool = False) -> None:
        n = len(offsets) - 1
        fn = L.lib.bh_push_batch_pinned if pinned else L.lib.bh_push_batch
        self._ck(fn(self._ctx, _ptr(slab), _ptr(offsets), _ptr(ts_ms), n))

    def push_batch_compact(self, slab, lengths, block_offsets, ts_block=None, ts_delta=None) -> None:
        """caller-pinned batch with compact sidecars (see compact_sidecars): 68 instead of 76 bytes per 64-byte event
        over PCIe"""
        self._ck(L.lib.bh_push_batch_compact(self._ctx, _ptr(slab), _ptr(lengths), _ptr(block_offsets), _ptr(ts_block),
                                             _ptr(ts_delta), len(lengths)))

    def push_columns(self, key_hash, ts_ms, value) -> None:
        self._ck(L.lib.bh_push_columns(self._ctx, _ptr(key_hash), _ptr(ts_ms), _ptr(value), len(key_hash)))

    def flush(self) -> int:
        seq = C.c_uint64()
        self._ck(L.lib.bh_flush(self._ctx, C.byref(seq)))
        return seq.value

    def poll_committed(self) -> int:
        seq = C.c_uint64()
        self._ck(L.lib.bh_poll_committed(self._ctx, C.byref(seq)))
        return seq.value

    def sync(self) -> None:
        self._ck(L.lib.bh_sync(self._ctx))

    def submit(self) -> None:
        """hand the partly filled ingest slab to the GPU now, without waiting for the result"""
        self._ck(L.lib.bh_submit(self._ctx))

    def set_commit_callback(self, fn) -> None:
        """fn(committed_seq) is called on the library's completion thread whenever more events have been
        aggregated (bh_set_commit_callback); None stops the thread"""
        cb = L.COMMIT_CB(lambda _user, seq: fn(seq)) if fn is not None else None
        self._ck(L.lib.bh_set_commit_callback(self._ctx, C.cast(cb, C.c_void_p) if cb is not None else None, None))
        self._commit_cb = cb                  # keep the trampoline alive as long as the library may call it

    # ---- device-resident entry points (pointers or torch tensors) ----
    def process_device(self, d_payload, d_offsets, d_ts, n: int, stream: int = 0) -> None:
        self._ck(L.lib.bh_process_device(self._ctx, _ptr(d_payload), _ptr(d_offsets), _ptr(d_ts), n, stream or None))

    def process_columns_device(self, d_kh, d_ts, d_val, n: int, stream: int = 0) -> None:
        self._ck(L.lib.bh_process_columns_device(self._ctx, _ptr(d_kh), _ptr(d_ts), _ptr(d_val), n, stream or None))

    def decode_device(self, d_payload, d_offsets, n: int, d_kh, d_val, d_status, d_outcome, stream: int = 0) -> None:
        self._ck(L.lib.bh_decode_device(self._ctx, _ptr(d_payload), _ptr(d_offsets), n, _ptr(d_kh), _ptr(d_val),
                                        _ptr(d_status), _ptr(d_outcome), stream or None))

    def export_window_device(self, window_id: int, n_owners: int, d_records, d_hist, cap: int, d_owner_counts,
                             owner_stride: int = 0, stream: int = 0) -> None:
        self._ck(L.lib.bh_export_window_device(self._ctx, window_id, n_owners, _ptr(d_records), _ptr(d_hist), cap,
                                               owner_stride, _ptr(d_owner_counts), stream or None))

    def export_window_peers(self, window_id: int, n_owners: int, d_peer_bases, my_rank: int, owner_stride: int,
                            d_owner_counts, d_peer_hist_bases=None, stream: int = 0) -> None:
        """export straight into the owners' receive buffers (peer-mapped memory): see dist.PeerWindowExchange"""
        self._ck(L.lib.bh_export_window_peers(self._ctx, window_id, n_owners, _ptr(d_peer_bases), _ptr(d_peer_hist_bases),
                                              my_rank, owner_stride, _ptr(d_owner_counts), stream or None))

    def export_window_static(self, window_id: int, n_owners: int, d_records, d_hist, cap: int, d_owner_counts,
                             owner_stride: int = 0, stream: int = 0) -> None:
        """export_window_device without any device allocation: the per-owner cursors sit in an 8 KB array
        built into the library (at most BH_STATIC_EXPORT_MAX_OWNERS owners; calls on one GPU must not overlap)"""
        self._ck(L.lib.bh_export_window_static(self._ctx, window_id, n_owners, _ptr(d_records), _ptr(d_hist), cap,
                                               owner_stride, _ptr(d_owner_counts), stream or None))

    def close_window_device(self, window_id: int, stream: int = 0) -> None:
        self._ck(L.lib.bh_close_window_device(self._ctx, window_id, stream or None))

    def drain_device(self, stream: int = 0) -> None:
        """partitioned aggregation: fold the logged tuples into the tables now (reads do it by themselves)"""
        self._ck(L.lib.bh_drain_device(self._ctx, stream or None))

    def reset_device(self, stream: int = 0) -> None:
        self._ck(L.lib.bh_reset_device(self._ctx, stream or None))

    def merge_records_device(self, d_records, d_hist, n: int, stream: int = 0) -> None:
        self._ck(L.lib.bh_merge_records_device(self._ctx, _ptr(d_records), _ptr(d_hist), n, stream or None))

    # ---- read side ----
    def snapshot(self, window_id: int = 0, with_hist: bool = False, cap: int = 0):
        """rows of one window, sorted by key_hash; (records, hist) when with_hist"""
        cap = cap or 1 << 16
        while True:
            out = np.zeros(cap, dtype=RECORD_DTYPE)
            hist = np.zeros((cap, self.hist_buckets), dtype=np.uint32) if with_hist else None
            n = C.c_uint64()
            self._ck(L.lib.bh_snapshot_hist(self._ctx, window_id, out.ctypes.data,
                                            hist.ctypes.data if hist is not None else None, cap, C.byref(n)))
            if n.value <= cap:
                break
            cap = int(n.value)
        out = out[: n.value]
        order = np.argsort(out["key_hash"], kind="stable")
        out = out[order]
        if with_hist:
            return out, hist[: n.value][order]
        return out

    def snapshot_all(self, with_hist: bool = False):
        recs, hists = [], []
        for w in self.list_windows():
            r = self.snapshot(w, with_hist)
            if with_hist:
                recs.append(r[0]); hists.append(r[1])
            else:
                recs.append(r)
        if not recs:
            empty = np.zeros(0, dtype=RECORD_DTYPE)
            return (empty, np.zeros((0, self.hist_buckets), np.uint32)) if with_hist else empty
        return (np.concatenate(recs), np.concatenate(hists)) if with_hist else np.concatenate(recs)

    # ---- snapshot / restore (the reference loses its counters on restart, index.js:32) ----
    def restore(self, records: np.ndarray, hist: Optional[np.ndarray] = None) -> None:
        records = np.ascontiguousarray(records, dtype=RECORD_DTYPE)
        if hist is not None:
            hist = np.ascontiguousarray(hist, dtype=np.uint32)
        self._ck(L.lib.bh_restore(self._ctx, records.ctypes.data, hist.ctypes.data if hist is not None else None, len(records)))

    def save(self, path: str) -> int:
        """every open window (rows + histograms if kept) into an .npz file; returns the row count"""
        with_hist = bool(self.ops & L.OP_HIST)
        snap = self.snapshot_all(with_hist=with_hist)
        recs, hist = snap if with_hist else (snap, np.zeros((0, self.hist_buckets), np.uint32))
        np.savez_compressed(path, records=recs, hist=hist, counters=np.array(list(self.counters(check=False).values()), dtype=np.uint64))
        return len(recs)

    def load(self, path: str) -> int:
        z = np.load(path)
        recs = z["records"].view(RECORD_DTYPE) if z["records"].dtype != RECORD_DTYPE else z["records"]
        self.restore(recs, z["hist"] if len(z["hist"]) else None)
        return len(recs)

    def list_windows(self):
        out = np.zeros(1024, dtype=np.uint64)
        n = C.c_uint32()
        self._ck(L.lib.bh_list_windows(self._ctx, out.ctypes.data, 1024, C.byref(n)))
        return [int(x) for x in out[: n.value]]

    def close_window(self, window_id: int) -> None:
        self._ck(L.lib.bh_close_window(self._ctx, window_id))

    def counters(self, check: bool = True) -> dict:
        k = L.BhCounters()
        rc = L.lib.bh_get_counters(self._ctx, C.byref(k))
        if check and rc not in (L.BH_OK,):
            self._ck(rc)
        return {f: getattr(k, f) for f, _ in L.BhCounters._fields_ if f != "reserved"}

    # ---- REFERENCE mode convenience: the Counter's label -> value map ----
    def status_counts(self) -> dict:
        """{lower-cased enum name: count} == beholder_progress_updates_total{status=...} (index.js:136-138)"""
        recs = self.snapshot(0)
        out = {}
        for r in recs:
            name = self.schema.enum_name(int(r["key_hash"]))
            if name is not None:
                out[name.lower()] = int(r["count"])
        return out


def compact_sidecars(offsets: np.ndarray, ts_ms: Optional[np.ndarray] = None):
    """(lengths u16[n], block_offsets u64[ceil(n / 65536) + 1], ts_block u64[...], ts_delta i16[n]) for
    bh_push_batch_compact from the u32 offsets (relative to the slab start) and u64 timestamps of a batch.  Raises
    when a message is longer than 65 535 bytes or a timestamp is further than 32 767 ms from its block's base."""
    B = L.BH_COMPACT_BLOCK
    n = len(offsets) - 1
    lens = np.diff(offsets.astype(np.int64))
    if n and (lens.min() < 0 or lens.max() > 0xFFFF):
        raise ValueError("message length outside 0..65535")
    rel = offsets.astype(np.uint64) - np.uint64(offsets[0])
    nb = (n + B - 1) // B
    boff = np.concatenate([rel[0:n:B], rel[n:n + 1]]).astype(np.uint64) if n else np.zeros(1, np.uint64)
    assert len(boff) == nb + 1
    if ts_ms is None:
        return lens.astype(np.uint16), boff, None, None
    t = ts_ms.astype(np.int64)
    # base of a block = the middle of its range, so that both ends are within reach of an i16
    tsb = np.array([(int(t[k * B: (k + 1) * B].min()) + int(t[k * B: (k + 1) * B].max())) // 2 for k in range(nb)], dtype=np.int64)
    delta = t - np.repeat(tsb, B)[:n]
    if n and (delta.min() < -32768 or delta.max() > 32767):
        raise ValueError("a timestamp is further than 32767 ms from its block's base: use push_batch")
    return lens.astype(np.uint16), boff, tsb.astype(np.uint64), delta.astype(np.int16)


class PinnedBuffer:
    """page-locked host memory from bh_host_alloc, viewed as a numpy array"""

    def __init__(self, nbytes: int):
        p = C.c_void_p()
        L.check(L.lib.bh_host_alloc(C.byref(p), max(nbytes, 1)))
        self.ptr = p.value
        self.nbytes = nbytes
        self._buf = (C.c_uint8 * max(nbytes, 1)).from_address(self.ptr)

    def view(self, dtype, count: Optional[int] = None, offset: int = 0) -> np.ndarray:
        return np.frombuffer(self._buf, dtype=dtype, count=-1 if count is None else count, offset=offset)

    def free(self) -> None:
        if self.ptr:
            self._buf = None
            L.lib.bh_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
