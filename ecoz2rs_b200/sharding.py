"""Host-side sharding rules of the multi-GPU VQ path.

Frames are independent, so they are split in contiguous blocks, one per rank
(the reference splits them over OpenMP threads, ecoz2/src/vq/vq_learn_par.c:62;
its serial merge :88-111 becomes ONE allreduce of the packed buffer below).
The same two rules are implemented in C++ (Engine::frames_reserve and
vqb200::PackedLayout); this module is what bench.py and the multi-rank tests use.
"""
from __future__ import annotations

import numpy as np


def shard_range(n_total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block [lo, hi) of rank `rank`: ceil(n/world) frames per rank, the
    last ranks may get fewer (or none)."""
    per = (n_total + world - 1) // world
    lo = min(rank * per, n_total)
    return lo, min(lo + per, n_total)


def packed_size(K: int, D: int) -> int:
    """[K*D cell sums | K counts | K cell distortions | DD | inertia]"""
    return K * D + 2 * K + 2


def pack(cells: np.ndarray, counts: np.ndarray, discel: np.ndarray, DD: float,
         inertia: float = 0.0) -> np.ndarray:
    K, D = cells.shape
    buf = np.empty(packed_size(K, D), np.float64)
    buf[:K * D] = cells.ravel()
    buf[K * D:K * D + K] = counts          # counts travel as f64: exact below 2^53
    buf[K * D + K:K * D + 2 * K] = discel
    buf[K * D + 2 * K] = DD
    buf[K * D + 2 * K + 1] = inertia
    return buf


def unpack(buf: np.ndarray, K: int, D: int):
    cells = buf[:K * D].reshape(K, D).copy()
    counts = np.rint(buf[K * D:K * D + K]).astype(np.int32)
    discel = buf[K * D + K:K * D + 2 * K].copy()
    return cells, counts, discel, float(buf[K * D + 2 * K]), float(buf[K * D + 2 * K + 1])
