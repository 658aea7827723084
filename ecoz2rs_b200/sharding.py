"""Host-side sharding rules of the multi-GPU VQ path, as the C-ABI library implements them.

Frames are independent, so they are split in contiguous blocks, one per rank (the reference
splits them over OpenMP threads, ecoz2/src/vq/vq_learn_par.c:62; its serial merge :88-111
becomes ONE allreduce of the packed buffer).  Both rules live in the library -- `shard_range`
(csrc/vq_engine.h, used by Engine::frames_reserve) and `PackedLayout` (csrc/vq_kernels.cuh, the
buffer ncclAllReduce sums) -- and are reached here through `ecoz2b200_shard_range` /
`ecoz2b200_packed_layout`, host-only entry points that need no GPU.  `bench.py`, the multi-rank
GPU test and the world-size-2 gloo test all take their shard boundaries and buffer offsets from
these calls, i.e. from the code the engine itself runs.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import vq


def shard_range(n_total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block [lo, hi) of rank `rank`: ceil(n/world) frames per rank, the
    last ranks may get fewer (or none)."""
    lo, hi = C.c_long(0), C.c_long(0)
    rc = vq.lib().ecoz2b200_shard_range(n_total, rank, world, C.byref(lo), C.byref(hi))
    if rc:
        raise ValueError(f"shard_range({n_total}, {rank}, {world})")
    return lo.value, hi.value


def packed_offsets(K: int, D: int) -> dict:
    off = (C.c_long * 6)()
    rc = vq.lib().ecoz2b200_packed_layout(K, D - 1, off)
    if rc:
        raise ValueError(f"packed_layout({K}, {D})")
    return dict(zip(("cells", "counts", "discel", "dd", "inertia", "size"), off))


def packed_size(K: int, D: int) -> int:
    """[K*D cell sums | K counts | K cell distortions | DD | inertia]"""
    return packed_offsets(K, D)["size"]


def pack(cells: np.ndarray, counts: np.ndarray, discel: np.ndarray, DD: float,
         inertia: float = 0.0) -> np.ndarray:
    K, D = cells.shape
    o = packed_offsets(K, D)
    buf = np.zeros(o["size"], np.float64)
    buf[o["cells"]:o["cells"] + K * D] = cells.ravel()
    buf[o["counts"]:o["counts"] + K] = counts          # counts travel as f64: exact below 2^53
    buf[o["discel"]:o["discel"] + K] = discel
    buf[o["dd"]] = DD
    buf[o["inertia"]] = inertia
    return buf


def unpack(buf: np.ndarray, K: int, D: int):
    o = packed_offsets(K, D)
    cells = buf[o["cells"]:o["cells"] + K * D].reshape(K, D).copy()
    counts = np.rint(buf[o["counts"]:o["counts"] + K]).astype(np.int32)
    discel = buf[o["discel"]:o["discel"] + K].copy()
    return cells, counts, discel, float(buf[o["dd"]]), float(buf[o["inertia"]])
