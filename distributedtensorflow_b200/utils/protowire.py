"""Minimal protobuf wire-format encoder/decoder.

TensorFlow is not installable here, so the few TF protos this framework must read/write
(``Event``/``Summary`` for TensorBoard, ``BundleHeaderProto``/``BundleEntryProto`` for TF1
checkpoints, ``ClusterDef``, ``CheckpointState``) are hand-encoded from their field numbers.
The C++ host library has the same encoders for the hot paths; this module is the portable
fallback and the decoder used by the readers.
"""
from __future__ import annotations

import struct
from typing import Dict, Iterator, List, Tuple, Union

WT_VARINT, WT_FIXED64, WT_LEN, WT_FIXED32 = 0, 1, 2, 5


def varint(n: int) -> bytes:
    if n < 0:
        n += 1 << 64
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def tag(field: int, wt: int) -> bytes:
    return varint((field << 3) | wt)


def field_varint(field: int, value: int) -> bytes:
    return tag(field, WT_VARINT) + varint(value)


def field_bytes(field: int, value: bytes) -> bytes:
    return tag(field, WT_LEN) + varint(len(value)) + value


def field_string(field: int, value: str) -> bytes:
    return field_bytes(field, value.encode("utf-8"))


def field_fixed32(field: int, value: int) -> bytes:
    return tag(field, WT_FIXED32) + struct.pack("<I", value & 0xFFFFFFFF)


def field_float(field: int, value: float) -> bytes:
    return tag(field, WT_FIXED32) + struct.pack("<f", value)


def field_double(field: int, value: float) -> bytes:
    return tag(field, WT_FIXED64) + struct.pack("<d", value)


def field_fixed64(field: int, value: int) -> bytes:
    return tag(field, WT_FIXED64) + struct.pack("<Q", value & 0xFFFFFFFFFFFFFFFF)


def read_varint(buf: bytes, pos: int) -> Tuple[int, int]:
    result = 0
    shift = 0
    while True:
        if pos >= len(buf):
            raise ValueError("truncated varint")
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7
        if shift > 70:
            raise ValueError("varint too long")


Value = Union[int, bytes]


def iter_fields(buf: bytes) -> Iterator[Tuple[int, int, Value]]:
    """Yields (field number, wire type, value). LEN values are bytes; fixed values are raw ints."""
    pos = 0
    n = len(buf)
    while pos < n:
        key, pos = read_varint(buf, pos)
        field, wt = key >> 3, key & 7
        if wt == WT_VARINT:
            v, pos = read_varint(buf, pos)
        elif wt == WT_FIXED64:
            v = struct.unpack_from("<Q", buf, pos)[0]
            pos += 8
        elif wt == WT_LEN:
            ln, pos = read_varint(buf, pos)
            v = bytes(buf[pos:pos + ln])
            if len(v) != ln:
                raise ValueError("truncated length-delimited field")
            pos += ln
        elif wt == WT_FIXED32:
            v = struct.unpack_from("<I", buf, pos)[0]
            pos += 4
        else:
            raise ValueError(f"unsupported wire type {wt}")
        yield field, wt, v


def parse(buf: bytes) -> Dict[int, List[Value]]:
    out: Dict[int, List[Value]] = {}
    for f, _, v in iter_fields(buf):
        out.setdefault(f, []).append(v)
    return out


def as_float(v: int) -> float:
    return struct.unpack("<f", struct.pack("<I", v))[0]


def as_double(v: int) -> float:
    return struct.unpack("<d", struct.pack("<Q", v))[0]


def as_signed(v: int, bits: int = 64) -> int:
    return v - (1 << bits) if v >= 1 << (bits - 1) else v
