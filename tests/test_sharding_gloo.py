"""CPU tests of the N > 1 path: world_size-2 `gloo` ranks.

The multi-GPU design is: contiguous frame shards, per-rank partial
[cell sums | counts | cell distortions | DD] buffers, ONE sum-allreduce per pass,
then every rank runs the (deterministic) centroid update on identical inputs.
Here the per-rank arithmetic is done by the oracle (test infrastructure) so the
host-side rules -- shard boundaries, packed layout, reduction, replicated control
flow -- are exercised end to end without a GPU."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_cover_everything():
    from ecoz2rs_b200.sharding import shard_range
    for n in (0, 1, 7, 1000, 10_000_000):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c and a <= b and c <= d
            assert max(b - a for a, b in spans) - min(b - a for a, b in spans) <= (n + world - 1) // world


def test_pack_unpack_roundtrip():
    from ecoz2rs_b200.sharding import pack, packed_size, unpack
    rng = np.random.default_rng(0)
    K, D = 8, 13
    cells, counts = rng.normal(size=(K, D)), rng.integers(0, 1 << 30, K)
    discel = rng.normal(size=K)
    buf = pack(cells, counts, discel, 3.5, 7.25)
    assert buf.size == packed_size(K, D)
    c2, n2, d2, DD, inert = unpack(buf, K, D)
    assert np.array_equal(c2, cells) and np.array_equal(n2, counts) and np.array_equal(d2, discel)
    assert (DD, inert) == (3.5, 7.25)


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from ecoz2rs_b200.sharding import pack, shard_range, unpack
    from ecoz2rs_b200.synth import ar_mix
    from oracle import oracle as O
    P, N, stop_M, eps = 12, 6000, 16, 0.05
    x = ar_mix(N, P, seed=5)
    lo, hi = shard_range(N, rank, world)
    mine = x[lo:hi]
    refl, raas = O.initial(P)
    DDprv, pas, trace, stages = np.finfo(np.float64).max, 0, [], []
    while True:
        K = raas.shape[0]
        idx, dd, _ = O.assign(mine, raas)
        cells, cardd, discel, DD = O.accumulate(mine, idx, dd, K)
        buf = torch.from_numpy(pack(cells, cardd, discel, DD))
        dist.all_reduce(buf)                       # the one collective of a pass
        cells, cardd, discel, DD, _ = unpack(buf.numpy(), K, P + 1)
        trace.append((K, pas, DD))
        _, refl, raas = O.centroids(cells, cardd, refl, raas)
        if pas > 0 and (DDprv - DD) / DD < eps:    # identical on every rank: same bits in
            stages.append((K, pas + 1, refl.copy()))
            if K >= stop_M:
                break
            pas = 0
            refl, raas = O.grow(refl)
        else:
            pas += 1
        DDprv = DD
    q.put((rank, trace, stages))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_learn_equals_single_rank():
    from ecoz2rs_b200.synth import ar_mix
    from oracle import oracle as O
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ref_stages, ref_trace = O.learn(ar_mix(6000, 12, seed=5), eps=0.05, max_M=16)
    (_, t0, s0), (_, t1, s1) = results
    # both ranks took bit-identical decisions
    assert t0 == t1
    assert all(np.array_equal(a[2], b[2]) for a, b in zip(s0, s1))
    # and the sharded run reproduces the single-rank run
    assert [(k, p) for k, p, _ in t0] == [(k, p) for k, p, _ in ref_trace]
    for (_, _, dd), (_, _, dd_ref) in zip(t0, ref_trace):
        assert abs(dd - dd_ref) <= 1e-11 * abs(dd_ref)
    for (K, passes, refl), r in zip(s0, ref_stages):
        assert (K, passes) == (r["M"], r["passes"])
        assert np.allclose(refl, r["refl"], rtol=1e-9, atol=1e-12)
