"""Seeded synthetic api.TelemetryProgress streams: wrappers over bh_synth_*.

The generator itself lives in libbeholder_gpu.so (csrc/bh_common.cuh, one definition compiled
for host and device).  `generate_host` runs the host build of it (no GPU needed);
`generate_device` runs the CUDA build into torch tensors.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import _lib as L


@dataclass
class SynthSpec:
    n_keys: int = 1000
    n_status: int = 6
    seed: int = 0xBE401DE5
    t0_ms: int = 1_700_000_000_000
    rate_per_ms: int = 500          # 100 M events -> 200 s of stream time -> 20+ ten-second windows
    jitter_ms: int = 2000
    payload_bytes: int = 64
    value_f64: bool = False
    value_exp_lo: int = -8
    value_exp_span: int = 24
    value_signed: bool = False
    malformed_ppm: int = 0
    unknown_enum_ppm: int = 0
    zipf: bool = False

    def to_c(self) -> "L.BhSynthConfig":
        c = L.BhSynthConfig()
        c.struct_size = C.sizeof(c)
        c.n_status = self.n_status
        c.seed = self.seed
        c.n_keys = self.n_keys
        c.t0_ms = self.t0_ms
        c.rate_per_ms = self.rate_per_ms
        c.jitter_ms = self.jitter_ms
        c.payload_bytes = self.payload_bytes
        c.value_f64 = int(self.value_f64)
        c.value_exp_lo = self.value_exp_lo
        c.value_exp_span = self.value_exp_span
        c.value_signed = int(self.value_signed)
        c.malformed_ppm = self.malformed_ppm
        c.unknown_enum_ppm = self.unknown_enum_ppm
        c.zipf = int(self.zipf)
        return c

    def max_event_bytes(self) -> int:
        return max(38 + 1 + 3 + (9 if self.value_f64 else 2) + 18, self.payload_bytes)


def zipf_cdf(n_keys: int, s: float = 1.0) -> np.ndarray:
    """u64 thresholds: key k is drawn when cdf[k-1] <= u < cdf[k]; cdf[-1] = 2^64-1."""
    w = 1.0 / np.arange(1, n_keys + 1, dtype=np.float64) ** s
    c = np.cumsum(w)
    c /= c[-1]
    t = np.minimum(np.floor(c * 18446744073709551616.0), 18446744073709549568.0).astype(np.uint64)
    t[-1] = np.uint64(0xFFFFFFFFFFFFFFFF)
    return t


def generate_host(spec: SynthSpec, first: int, n: int, cdf: Optional[np.ndarray] = None):
    """(slab u8[], offsets u32[n+1], ts u64[n]) built on the CPU, event by event."""
    c = spec.to_c()
    if spec.zipf and cdf is None:
        cdf = zipf_cdf(spec.n_keys)
    cdf_p = cdf.ctypes.data if cdf is not None else None
    cap = spec.max_event_bytes()
    slab = np.zeros(n * cap + 16, dtype=np.uint8)
    offsets = np.zeros(n + 1, dtype=np.uint32)
    ts = np.zeros(n, dtype=np.uint64)
    t = C.c_uint64()
    pos = 0
    base = slab.ctypes.data
    for i in range(n):
        ln = L.lib.bh_synth_event_host(C.byref(c), first + i, cdf_p, base + pos, C.byref(t))
        ts[i] = t.value
        pos += ln
        offsets[i + 1] = pos
    return slab[: (pos + 15) // 16 * 16].copy(), offsets, ts


def generate_device(spec: SynthSpec, first: int, n: int, device: int = 0, d_cdf=None, out=None):
    """Generate on the GPU into torch tensors (payload u8 [padded to 16], offsets i32-as-u32 [n+1],
    ts i64-as-u64 [n]).  `out` may hold preallocated (payload, offsets, ts) tensors."""
    import torch
    c = spec.to_c()
    dev = torch.device("cuda", device)
    if spec.zipf and d_cdf is None:
        d_cdf = torch.from_numpy(zipf_cdf(spec.n_keys).view(np.int64)).to(dev)
    if out is None:
        cap = (n * spec.max_event_bytes() + 15) // 16 * 16 + 16
        payload = torch.empty(cap, dtype=torch.uint8, device=dev)
        offsets = torch.empty(n + 1, dtype=torch.int32, device=dev)
        ts = torch.empty(n, dtype=torch.int64, device=dev)
    else:
        payload, offsets, ts = out
    total = C.c_uint64()
    stream = torch.cuda.current_stream(dev).cuda_stream
    L.check(L.lib.bh_synth_generate_device(device, C.byref(c), first, n, d_cdf.data_ptr() if d_cdf is not None else None,
                                           payload.data_ptr(), payload.numel(), offsets.data_ptr(), ts.data_ptr(),
                                           C.byref(total), stream or None))
    torch.cuda.current_stream(dev).synchronize()   # consumers may run on other (non-blocking) streams
    return payload, offsets, ts, int(total.value)


def columns_device(spec: SynthSpec, hash_seed: int, first: int, n: int, device: int = 0, d_cdf=None):
    import torch
    c = spec.to_c()
    dev = torch.device("cuda", device)
    if spec.zipf and d_cdf is None:
        d_cdf = torch.from_numpy(zipf_cdf(spec.n_keys).view(np.int64)).to(dev)
    kh = torch.empty(n, dtype=torch.int64, device=dev)
    ts = torch.empty(n, dtype=torch.int64, device=dev)
    val = torch.empty(n, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    L.check(L.lib.bh_synth_columns_device(device, C.byref(c), hash_seed, first, n,
                                          d_cdf.data_ptr() if d_cdf is not None else None,
                                          kh.data_ptr(), ts.data_ptr(), val.data_ptr(), stream or None))
    torch.cuda.current_stream(dev).synchronize()
    return kh, ts, val
