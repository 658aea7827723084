"""NumPy readers/writers for the three ECOZ2 binary formats on the VQ path.

They exist for tests, fixtures and the benchmark (writing synthetic `.prd`
shards, reading back `.cbook` / `.seq` results).  The product's own I/O is the
C++ code in `csrc/vq_io.cpp`; both follow the same reference layouts:

* `.prd`   ecoz2/src/lpc/prd.c:76-115   `<predictor>` | class | i32 P | i32 T | T x (1+P) f64
* `.cbook` ecoz2/src/vq/vq.c:22-59      `<codebook>`  | class | i32 P | i32 M | M x (1+P) f64 reflections
* `.seq`   ecoz2/src/hmm/symbol.c:17-36 `<sequence>`  | class | i32 T | i32 M | T x u16

Header fields: a 16-byte NUL-padded ident (ecoz2/src/utl/fileutil.c:121-139) and
a 96-byte class name, NUL-terminated then padded with '_'
(ecoz2/src/utl/fileutil.c:162-182).  Integers and floats are native-endian
(little-endian on every supported host; the Rust `.seq` reader assumes LE,
src/sequence/mod.rs:49-75).
"""
from __future__ import annotations

import os
import struct

import numpy as np

FILE_IDENT_LEN = 16        # ecoz2/src/include/utl.h:14, src/utl/mod.rs:19
MAX_CLASS_NAME_LEN = 96    # ecoz2/src/include/utl.h:16, src/utl/mod.rs:20
HEADER_LEN = FILE_IDENT_LEN + MAX_CLASS_NAME_LEN + 8


def _ident(ident: str) -> bytes:
    b = ident.encode()
    assert len(b) < FILE_IDENT_LEN
    return b + b"\0" * (FILE_IDENT_LEN - len(b))


def _class_name(name: str) -> bytes:
    b = name.encode()[: MAX_CLASS_NAME_LEN - 1]
    return b + b"\0" + b"_" * (MAX_CLASS_NAME_LEN - len(b) - 1)


def _read_header(f, ident: str):
    head = f.read(FILE_IDENT_LEN)
    if len(head) != FILE_IDENT_LEN or not head.startswith(ident.encode()):
        raise ValueError(f"not a {ident} file")
    cls = f.read(MAX_CLASS_NAME_LEN)
    if len(cls) != MAX_CLASS_NAME_LEN:
        raise ValueError("truncated class name")
    cls = cls[: MAX_CLASS_NAME_LEN - 1].split(b"\0", 1)[0].decode()
    a, b = struct.unpack("<ii", f.read(8))
    return cls, a, b


def write_prd(path: str, vectors: np.ndarray, class_name: str = "_") -> None:
    v = np.ascontiguousarray(vectors, dtype="<f8")
    T, D = v.shape
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "wb") as f:
        f.write(_ident("<predictor>"))
        f.write(_class_name(class_name))
        f.write(struct.pack("<ii", D - 1, T))
        f.write(v.tobytes())


def read_prd(path: str):
    with open(path, "rb") as f:
        cls, P, T = _read_header(f, "<predictor>")
        v = np.frombuffer(f.read(T * (P + 1) * 8), dtype="<f8")
    if v.size != T * (P + 1):
        raise ValueError("truncated predictor file")
    return cls, v.reshape(T, P + 1).copy()


def write_cbook(path: str, reflections: np.ndarray, class_name: str = "_") -> None:
    r = np.ascontiguousarray(reflections, dtype="<f8")
    M, D = r.shape
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "wb") as f:
        f.write(_ident("<codebook>"))
        f.write(_class_name(class_name))
        f.write(struct.pack("<ii", D - 1, M))
        f.write(r.tobytes())


def read_cbook(path: str):
    with open(path, "rb") as f:
        cls, P, M = _read_header(f, "<codebook>")
        r = np.frombuffer(f.read(M * (P + 1) * 8), dtype="<f8")
    if r.size != M * (P + 1):
        raise ValueError("truncated codebook file")
    return cls, r.reshape(M, P + 1).copy()


def write_seq(path: str, symbols: np.ndarray, M: int, class_name: str = "_") -> None:
    s = np.ascontiguousarray(symbols, dtype="<u2")
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "wb") as f:
        f.write(_ident("<sequence>"))
        f.write(_class_name(class_name))
        f.write(struct.pack("<ii", s.size, M))
        f.write(s.tobytes())


def read_seq(path: str):
    with open(path, "rb") as f:
        cls, T, M = _read_header(f, "<sequence>")
        s = np.frombuffer(f.read(T * 2), dtype="<u2")
    if s.size != T:
        raise ValueError("truncated sequence file")
    return cls, M, s.copy()


def write_prd_shards(directory: str, vectors: np.ndarray, class_name: str = "_",
                     shard_frames: int = 1_000_000, stem: str = "part") -> list[str]:
    """Write `vectors` as `<directory>/<stem>_%05d.prd` shards of <= shard_frames frames.

    The reference cannot load one `.prd` above 2^31 / ((P+1)*8) frames: `new_matrix`
    does its size arithmetic in `int` (ecoz2/src/utl/memutil.c:20-21).
    """
    paths = []
    for i, lo in enumerate(range(0, len(vectors), shard_frames)):
        p = os.path.join(directory, f"{stem}_{i:05d}.prd")
        write_prd(p, vectors[lo:lo + shard_frames], class_name)
        paths.append(p)
    return paths
