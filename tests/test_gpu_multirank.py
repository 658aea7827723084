"""The multi-process deployment that bench.py / SCALE use: one process per GPU, real
`ecoz2b200_comm_init` (ncclCommInitRank over a shared unique id), ONE ncclAllReduce of the packed
cell buffer per pass (reference merge: ecoz2/src/vq/vq_learn_par.c:88-111).

Two spawned ranks on the two halves of the SAME frames must reproduce the single-rank run: same pass
counts per stage, per-pass DD to 1e-11, every stage's reflections to 1e-9, and -- on every rank --
identical replicated results.  Needs two GPUs (NCCL refuses two ranks on one device): skipped on a
single-GPU box; run with `gpurun --gpus 2` (log under profiles/)."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "tools", "multirank_worker.py")


def _run_world(world, frames_path, workdir, stop_M):
    os.makedirs(workdir, exist_ok=True)
    procs = [subprocess.Popen([sys.executable, WORKER, str(r), str(world), frames_path, workdir, str(stop_M)],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(world)]
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=600)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out)
    for r, p in enumerate(procs):
        assert p.returncode == 0, f"rank {r} of {world} failed:\n{outs[r][-3000:]}"
    return [np.load(os.path.join(workdir, f"rank{r}.npz")) for r in range(world)]


@pytest.mark.parametrize("P,N,stop_M", [(36, 400_003, 1024), (12, 300_001, 256)], ids=["P36", "P12"])
def test_two_ranks_reproduce_single_rank(built_lib, tmp_path, P, N, stop_M):
    from ecoz2rs_b200 import vq
    from ecoz2rs_b200.synth import ar_mix
    if vq.device_count() < 2:
        pytest.skip("needs two GPUs (one process per GPU)")
    x = ar_mix(N, P, seed=20260102)
    frames_path = str(tmp_path / "frames.npy")
    np.save(frames_path, x)
    one = _run_world(1, frames_path, str(tmp_path / "w1"), stop_M)[0]
    two = _run_world(2, frames_path, str(tmp_path / "w2"), stop_M)
    assert (int(two[0]["lo"]), int(two[1]["hi"])) == (0, N) and int(two[0]["hi"]) == int(two[1]["lo"])
    for r in two:
        # identical control flow: stages, passes per stage, every pass of the trace
        assert np.array_equal(r["M"], one["M"]) and np.array_equal(r["passes"], one["passes"])
        assert np.array_equal(r["trace_M"], one["trace_M"]) and np.array_equal(r["trace_pass"], one["trace_pass"])
        assert np.allclose(r["trace_DD"], one["trace_DD"], rtol=1e-11, atol=0)
        assert np.allclose(r["refl"], one["refl"], rtol=1e-9, atol=1e-12)
        assert np.array_equal(r["cardd"], one["cardd"])
        assert np.allclose(r["avg"], one["avg"], rtol=1e-11) and np.allclose(r["sigma"], one["sigma"], rtol=1e-9)
        assert np.allclose(r["inertia"], one["inertia"], rtol=1e-10)
        # the extra pass: resident and host-frame routes, global sums on every rank
        for k in ("res", "host"):
            assert np.array_equal(r[k + "_cardd"], one["res_cardd"])
            assert np.allclose(r[k + "_cells"], one["res_cells"], rtol=1e-10, atol=1e-9)
            assert abs(float(r[k + "_DD"]) - float(one["res_DD"])) <= 1e-11 * abs(float(one["res_DD"]))
            assert np.allclose(r[k + "_refl"], one["res_refl"], rtol=1e-9, atol=1e-12)
        assert abs(float(r["inert"]) - float(one["inert"])) <= 1e-10 * abs(float(one["inert"]))
        assert abs(float(r["q_avg"]) - float(one["q_avg"])) <= 1e-11 * abs(float(one["q_avg"]))
    # replicated state is bit-identical across the ranks of one run
    assert np.array_equal(two[0]["refl"], two[1]["refl"]) and np.array_equal(two[0]["trace_DD"], two[1]["trace_DD"])
    # the symbols of the two shards are the single rank's symbols
    assert np.array_equal(np.concatenate([two[0]["seq"], two[1]["seq"]]), one["seq"])
