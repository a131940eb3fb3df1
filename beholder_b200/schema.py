"""Message schema of the progress path.

The reference loads `api.TelemetryProgress` from triton-core (`proto.load`, index.js:46) and
destructures `{ mediaId, status, progress, host }` (index.js:131).  The .proto itself is not in
the reference tree, so field numbers, the scalar type of `progress` and the value list of
`TelemetryStatusEntry` below are an ASSUMPTION (a recollection of tritonmedia/core), kept
configurable: pass your own `Schema` to `BeholderGPU` / `Aggregator` if the real .proto differs.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

from . import _lib as L


@dataclass(frozen=True)
class Field:
    name: str
    number: int
    kind: int            # _lib.K_*
    role: int = 0        # _lib.R_* bits


@dataclass
class Schema:
    fields: List[Field]
    enum: Dict[str, int] = field(default_factory=dict)   # TelemetryStatusEntry name -> number
    type_name: str = "api.TelemetryProgress"

    def enum_name(self, number: int) -> Optional[str]:
        """proto.enumToString(type, 'TelemetryStatusEntry', number) — index.js:134"""
        for n, v in self.enum.items():
            if v == number:
                return n
        return None

    def enum_number(self, name: str) -> Optional[int]:
        """proto.stringToEnum(type, 'TelemetryStatusEntry', name) — index.js:94"""
        return self.enum.get(name)

    def with_roles(self, key: Sequence[str] = (), status: Optional[str] = "status",
                   value: Optional[str] = "progress") -> "Schema":
        out = []
        for f in self.fields:
            role = 0
            if f.name in key:
                role |= L.R_KEY
            if status and f.name == status:
                role |= L.R_STATUS
            if value and f.name == value:
                role |= L.R_VALUE
            out.append(Field(f.name, f.number, f.kind, role))
        return Schema(out, dict(self.enum), self.type_name)


def telemetry_progress(progress_kind: int = L.K_INT32) -> Schema:
    """ASSUMED api.TelemetryProgress: mediaId=1 string, status=2 enum, progress=3, host=4 string."""
    return Schema(
        fields=[
            Field("mediaId", 1, L.K_STRING),
            Field("status", 2, L.K_INT32, L.R_STATUS),
            Field("progress", 3, progress_kind, L.R_VALUE),
            Field("host", 4, L.K_STRING),
        ],
        enum={"QUEUED": 0, "DOWNLOADING": 1, "CONVERTING": 2, "UPLOADING": 3, "DEPLOYED": 4, "ERRORED": 5},
    )


def telemetry_status() -> Schema:
    """ASSUMED api.TelemetryStatus (index.js:47,63-64 destructures { mediaId, status }): mediaId=1, status=2."""
    return Schema(
        fields=[Field("mediaId", 1, L.K_STRING), Field("status", 2, L.K_INT32, L.R_STATUS)],
        enum=dict(telemetry_progress().enum), type_name="api.TelemetryStatus",
    )


def fill_config(cfg: "L.BhConfig", schema: Schema) -> None:
    if len(schema.fields) > L.BH_MAX_FIELDS:
        raise ValueError("too many fields")
    cfg.n_fields = len(schema.fields)
    for i, f in enumerate(schema.fields):
        cfg.fields[i].number = f.number
        cfg.fields[i].kind = f.kind
        cfg.fields[i].role = f.role
    nums = sorted(set(schema.enum.values()))
    if len(nums) > L.BH_MAX_ENUM:
        raise ValueError("too many enum values")
    cfg.n_enum = len(nums)
    for i, v in enumerate(nums):
        cfg.enum_numbers[i] = v


# ---- minimal protobuf wire encoder (host side; used by tests, the stub queue and examples) ----
def _varint(v: int) -> bytes:
    v &= (1 << 64) - 1
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def encode(schema: Schema, **values) -> bytes:
    """Serialise a message of `schema` (proto3: absent/default fields are skipped unless given)."""
    import struct
    out = bytearray()
    for f in schema.fields:
        if f.name not in values or values[f.name] is None:
            continue
        v = values[f.name]
        if f.kind in (L.K_STRING, L.K_BYTES):
            b = v.encode() if isinstance(v, str) else bytes(v)
            out += _varint((f.number << 3) | 2) + _varint(len(b)) + b
        elif f.kind in (L.K_INT32, L.K_UINT32, L.K_BOOL):
            out += _varint((f.number << 3) | 0) + _varint(int(v))
        elif f.kind == L.K_SINT32:
            iv = int(v)
            out += _varint((f.number << 3) | 0) + _varint(((iv << 1) ^ (iv >> 31)) & 0xFFFFFFFF)
        elif f.kind == L.K_FLOAT:
            out += _varint((f.number << 3) | 5) + struct.pack("<f", float(v))
        elif f.kind == L.K_DOUBLE:
            out += _varint((f.number << 3) | 1) + struct.pack("<d", float(v))
        else:
            raise ValueError(f.kind)
    return bytes(out)


# ---- host-side decode of ONE message (handler glue only: which messages reach the side effects) ----
class _Reader:
    """protobufjs 6.8.8 Reader / BufferReader on a Python bytes object: uint32() takes 5 value bytes, assumes a
    longer varint is 10 bytes long and treats bytes past the end as continuation bytes; string() clamps to the
    end of the buffer (Node Buffer), bytes() and the fixed-width reads throw past it."""

    def __init__(self, buf: bytes):
        self.buf, self.pos, self.len = buf, 0, len(buf)

    def _peek(self, pos: int) -> int:
        return self.buf[pos] if pos < self.len else 0x80

    def uint32(self) -> int:
        value = 0
        for i in range(5):
            b = self._peek(self.pos); self.pos += 1
            value |= (b & (15 if i == 4 else 127)) << (7 * i)
            if b < 128:
                return value & 0xFFFFFFFF
        self.pos += 5
        if self.pos > self.len:
            self.pos = self.len
            raise IndexError("index out of range")
        return value & 0xFFFFFFFF

    def skip(self, n: Optional[int] = None) -> None:
        if n is None:
            while True:
                if self.pos >= self.len:
                    raise IndexError("index out of range")
                b = self.buf[self.pos]; self.pos += 1
                if not b & 128:
                    return
        if self.pos + n > self.len:
            raise IndexError("index out of range")
        self.pos += n

    def skip_type(self, wire: int) -> None:
        depth = 0
        while True:
            if wire == 0:
                self.skip()
            elif wire == 1:
                self.skip(8)
            elif wire == 2:
                self.skip(self.uint32())
            elif wire == 3:
                depth += 1
            elif wire == 5:
                self.skip(4)
            elif wire == 4 and depth > 0:
                depth -= 1
            else:
                raise ValueError("invalid wire type")
            if depth == 0:
                return
            wire = self.uint32() & 7


def decode_message(schema: Schema, buf: bytes) -> Optional[dict]:
    """What `proto.decode(type, content)` (index.js:129) yields for the fields of `schema`, or None where it
    throws.  Host code for the handler glue (the GPU decodes the same bytes for the counters): it decides whether a
    delivery reaches the per-event side effects of index.js:140-148, which the reference only runs for messages
    that decoded and whose status has an enum name."""
    import struct
    by_number = {f.number: f for f in schema.fields}
    out = {f.name: (b"" if f.kind in (L.K_STRING, L.K_BYTES) else 0) for f in schema.fields}     # strings as their raw bytes
    r = _Reader(bytes(buf))
    try:
        while r.pos < r.len:
            t = r.uint32()
            f = by_number.get(t >> 3)
            if f is None:
                r.skip_type(t & 7)
                continue
            if f.kind in (L.K_STRING, L.K_BYTES):
                n = r.uint32()
                end = r.pos + n
                if end > r.len:
                    if f.kind == L.K_BYTES:
                        raise IndexError("index out of range")
                    end = r.len
                out[f.name] = r.buf[r.pos:end]
                r.pos = end
            elif f.kind in (L.K_INT32, L.K_UINT32, L.K_SINT32, L.K_BOOL):
                u = r.uint32()
                if f.kind == L.K_SINT32:
                    u = ((u >> 1) ^ -(u & 1)) & 0xFFFFFFFF
                if f.kind == L.K_BOOL:
                    u = int(u != 0)
                out[f.name] = u if f.kind in (L.K_UINT32, L.K_BOOL) else (u - (1 << 32) if u & 0x80000000 else u)
            elif f.kind == L.K_FLOAT:
                if r.pos + 4 > r.len:
                    raise IndexError("index out of range")
                out[f.name] = struct.unpack_from("<f", r.buf, r.pos)[0]; r.pos += 4
            else:
                if r.pos + 8 > r.len:
                    raise IndexError("index out of range")
                out[f.name] = struct.unpack_from("<d", r.buf, r.pos)[0]; r.pos += 8
    except (IndexError, ValueError):
        return None
    return out
