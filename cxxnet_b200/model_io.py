"""Binary stream helpers for the ``.model`` checkpoint format.

Little-endian, LP64.  Strings / vectors carry a uint64 length prefix (reference
``src/utils/io.h:44-76``); tensors are ``uint32 shape[dim]`` followed by contiguous float32
(mshadow ``SaveBinary``).
"""
from __future__ import annotations

import io
import struct
from typing import BinaryIO, List, Sequence

import numpy as np


class ModelFormatError(RuntimeError):
    pass


def _read_exact(f: BinaryIO, n: int) -> bytes:
    b = f.read(n)
    if len(b) != n:
        raise ModelFormatError("invalid model file: unexpected end of stream")
    return b


def write_i32(f: BinaryIO, v: int) -> None:
    f.write(struct.pack("<i", int(v)))


def read_i32(f: BinaryIO) -> int:
    return struct.unpack("<i", _read_exact(f, 4))[0]


def write_i64(f: BinaryIO, v: int) -> None:
    f.write(struct.pack("<q", int(v)))


def read_i64(f: BinaryIO) -> int:
    return struct.unpack("<q", _read_exact(f, 8))[0]


def write_str(f: BinaryIO, s) -> None:
    b = s if isinstance(s, (bytes, bytearray)) else s.encode("utf-8")
    f.write(struct.pack("<Q", len(b)))
    if b:
        f.write(b)


def read_bytes(f: BinaryIO) -> bytes:
    (n,) = struct.unpack("<Q", _read_exact(f, 8))
    return _read_exact(f, n) if n else b""


def read_str(f: BinaryIO) -> str:
    return read_bytes(f).decode("utf-8")


def write_i32_vec(f: BinaryIO, v: Sequence[int]) -> None:
    f.write(struct.pack("<Q", len(v)))
    if len(v):
        f.write(np.asarray(v, dtype="<i4").tobytes())


def read_i32_vec(f: BinaryIO) -> List[int]:
    (n,) = struct.unpack("<Q", _read_exact(f, 8))
    if n == 0:
        return []
    return np.frombuffer(_read_exact(f, 4 * n), dtype="<i4").astype(int).tolist()


def write_tensor(f: BinaryIO, arr) -> None:
    """mshadow SaveBinary: uint32 shape[dim] + float32 data."""
    a = np.ascontiguousarray(np.asarray(arr, dtype="<f4"))
    f.write(np.asarray(a.shape, dtype="<u4").tobytes())
    f.write(a.tobytes())


def read_tensor(f: BinaryIO, dim: int) -> np.ndarray:
    shape = np.frombuffer(_read_exact(f, 4 * dim), dtype="<u4").astype(np.int64)
    n = int(np.prod(shape)) if dim else 1
    data = np.frombuffer(_read_exact(f, 4 * n), dtype="<f4")
    return data.reshape([int(s) for s in shape]).copy()


def bytes_io(b: bytes = b"") -> io.BytesIO:
    return io.BytesIO(b)
