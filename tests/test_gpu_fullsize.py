"""Parity at BASELINE.json's full size (configs[1]: 10M frames, P=36, K=2048) through
size-independent properties -- the FP64 oracle would need hours there.

  * checksum of checksums: the K cell sums add up to the column sums of ALL frames
    (the accumulation is linear; any lost / doubled / misrouted frame shows);
  * counts add up to N; DD equals the sum of the cell distortions;
  * two independent routes to the total distortion agree: the pass derives it from the cell
    sums (linearity of the distortion in the frame), quantize sums per-frame FP64 distortions
    of the winners -- and the symbols are the pass's indices;
  * the search is idempotent (same indices on a second run) and equal to the all-FP64 search
    on a 200k-frame sample of the same frames;
  * k-means monotonicity: DD does not increase over consecutive passes.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N, P, K = 10_000_000, 36, 2048


def test_fullsize_pass_properties(built_lib):
    import torch
    from ecoz2rs_b200 import vq
    from ecoz2rs_b200.synth import ar_mix_torch
    xd = ar_mix_torch(N, P, seed=20260102, device="cuda")
    colsum = xd.sum(dim=0).cpu().numpy()
    sample = xd[::50].contiguous().cpu().numpy()          # 200k frames, kept on the host
    with vq.VqEngine(P, max_M=K) as eng:
        eng.set_frames_device(xd.data_ptr(), N)
        del xd
        torch.cuda.empty_cache()
        eng.codebook_initial()
        stages, trace = eng.learn(eps=0.05, stop_M=K)
        assert [s["M"] for s in stages] == [2 << i for i in range(11)]
        refl = stages[-1]["refl"]
        eng.set_reflections(refl)
        _, raas = eng.codebook()

        out = eng.kmeans_pass(want_cells=True)
        idx, md = eng.last_assignment()
        assert int(out["cardd"].sum()) == N
        assert np.array_equal(np.bincount(idx, minlength=K), out["cardd"])
        assert np.allclose(out["cells"].sum(axis=0), colsum, rtol=1e-11, atol=0)
        assert abs(out["DD"] - out["discel"].sum()) <= 1e-10 * abs(out["DD"])
        assert abs(md.sum() - out["DD"]) <= 1e-9 * abs(out["DD"])

        # quantize against the codebook the assignment above was made with
        eng.set_reflections(refl)
        seq, avg = eng.quantize_resident(want_seq=True)
        assert np.array_equal(seq.astype(np.int32), idx)
        assert abs(avg * N - out["DD"]) <= 1e-9 * abs(out["DD"])
        seq2, avg2 = eng.quantize_resident(want_seq=True)
        # (the block partial sums meet in FP64 atomics: the last bits depend on their order)
        assert np.array_equal(seq, seq2) and abs(avg - avg2) <= 1e-13 * abs(avg)

        # k-means: the total distortion never increases from one pass to the next
        eng.set_reflections(refl)
        dds = [eng.kmeans_pass(want_cells=False)["DD"] for _ in range(4)]
        assert all(b <= a * (1 + 1e-12) for a, b in zip(dds, dds[1:])), dds

    # the tensor-core search against the all-FP64 search on a sample of the same frames
    with vq.VqEngine(P, max_M=K, mode=vq.MODE_FP64) as e64:
        e64.set_frames(sample)
        e64.set_raas(raas)
        idx64, dd64, _ = e64.assign()
        e64.set_mode(vq.MODE_TF32X3)
        idx_tc, dd_tc, nre = e64.assign()
    assert np.array_equal(idx_tc, idx64)
    assert np.array_equal(idx[::50], idx64)
    assert np.allclose(dd_tc, dd64, rtol=1e-10, atol=0)
    assert 0 < nre < 0.05 * sample.shape[0]
