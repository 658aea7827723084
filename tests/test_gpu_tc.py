"""GPU tests of the tcgen05 split-TF32 search (mode TF32X3, csrc/vq_tc.cu).

The tensor-core values are a FILTER, not the answer: every frame whose best-two gap is
inside the error bound is re-evaluated in FP64.  These tests pin (i) the raw values and
the error constant the bound relies on, (ii) index parity against the FP64 oracle at shapes
that exercise every structural case of the kernel (ragged codebooks, partial / single /
many frame tiles, both prediction orders, gathered re-evaluation), and (iii) determinism
over repeated runs (the kernel is a pipeline of asynchronous copies, MMAs and TMEM loads
ordered only by mbarriers).
"""
import numpy as np
import pytest

from conftest import allowed_index_mismatches

pytestmark = pytest.mark.gpu

U24 = 2.0 ** -24
KAPPA_TC = 40.0          # Engine default (vq_engine.h): bound = kappa * 2^-24 * |rx0| * W1_j


def _w(raas):
    w = raas.copy()
    w[:, 1:] *= 2
    return w


def _random_codebook(K, P, seed):
    from oracle import oracle as O
    rng = np.random.default_rng(seed)
    refl = np.zeros((K, P + 1))
    refl[:, 1:] = np.clip(rng.normal(size=(K, P)) * 0.3, -0.9, 0.9)
    return O.refl_to_raas(refl)


@pytest.mark.parametrize("K", [64, 256], ids=["K64", "K256"])
def test_tc_values_within_bound(built_lib, frames36, oracle_run36, K):
    """|d_tc - d_fp64| <= kappa_tc * 2^-24 * |rx0| * W1_j with margin: the measured maximum
    stays below half of the bound the flagging relies on."""
    from ecoz2rs_b200 import vq
    stages, _ = oracle_run36
    raas = [s for s in stages if s["raas"].shape[0] == K][0]["raas"]
    n = 4096
    with vq.VqEngine(36, max_M=K, mode=vq.MODE_TF32X3) as eng:
        eng.set_frames(frames36[:n])
        eng.set_raas(raas)
        d = eng.debug_tc_distortions(n).astype(np.float64)
    w = _w(raas)
    ref = frames36[:n] @ w.T
    scale = np.abs(frames36[:n, :1]) * np.abs(w).sum(1)[None, :]
    worst = np.max(np.abs(d - ref) / scale) / U24
    assert worst < 0.5 * KAPPA_TC, f"tensor-core error {worst:.1f} u24 vs kappa {KAPPA_TC}"
    assert np.max(np.abs(d - ref) / np.abs(ref)) < 5e-3


@pytest.mark.parametrize("P,K,N", [
    (36, 2, 3000), (36, 5, 1000), (36, 33, 2000), (12, 8, 4000),
    (36, 64, 127), (36, 64, 256), (36, 100, 1000), (36, 129, 4097), (36, 300, 777),
    (36, 1000, 20000), (36, 2048, 9000), (12, 64, 5000), (12, 1024, 30000), (20, 200, 3000), (28, 300, 3000), (44, 500, 5000),
], ids=lambda v: str(v))
def test_tc_assign_matches_oracle(built_lib, frames36, frames12, P, K, N):
    from ecoz2rs_b200 import vq
    from ecoz2rs_b200.synth import ar_mix
    from oracle import oracle as O
    x = {36: frames36, 12: frames12}.get(P)
    if x is None:
        x = ar_mix(N, P, seed=7)
    x = x[:N]
    raas = _random_codebook(K, P, seed=K + P)
    idx_ref, dd_ref, d2_ref = O.assign(x, raas)
    with vq.VqEngine(P, max_M=max(K, 2), mode=vq.MODE_TF32X3) as eng:
        eng.set_frames(x)
        eng.set_raas(raas)
        idx, dd, nre = eng.assign()
    bad = allowed_index_mismatches(idx, idx_ref, dd_ref, d2_ref)
    assert bad.size == 0, f"{bad.size} non-exempt mismatches, first {bad[:5]}"
    same = idx == idx_ref
    # FP64 sums in a different order than the oracle's: rounding times the conditioning of x.w
    assert np.allclose(dd[same], dd_ref[same], rtol=1e-10, atol=0)
    assert nre < 0.25 * N


def test_tc_duplicate_codewords_first_index_wins(built_lib, frames36):
    """Exact ties (identical codewords) are equal in the tensor-core values too: both land in
    the re-evaluation, whose FP64 select keeps the lower index (strict '<' of quantize.c:19-26)."""
    from ecoz2rs_b200 import vq
    from oracle import oracle as O
    raas = _random_codebook(64, 36, seed=5)
    raas = np.concatenate([raas, raas[:64]])          # codeword 64 + i == codeword i
    idx_ref, _, _ = O.assign(frames36[:4000], raas)
    with vq.VqEngine(36, max_M=128, mode=vq.MODE_TF32X3) as eng:
        eng.set_frames(frames36[:4000])
        eng.set_raas(raas)
        idx, _, nre = eng.assign()
    assert np.array_equal(idx, idx_ref)
    assert idx.max() < 64 and nre == 4000


def test_tc_repeatable(built_lib, frames36):
    """20 repetitions on a many-tile shape: identical indices every time."""
    from ecoz2rs_b200 import vq
    from ecoz2rs_b200.synth import ar_mix
    x = ar_mix(150_000, 36, seed=11)
    raas = _random_codebook(512, 36, seed=3)
    with vq.VqEngine(36, max_M=512, mode=vq.MODE_FP64) as eng:
        eng.set_frames(x)
        eng.set_raas(raas)
        idx64, _, _ = eng.assign()
        eng.set_mode(vq.MODE_TF32X3)
        for rep in range(20):
            idx, _, _ = eng.assign()
            assert np.array_equal(idx, idx64), f"repetition {rep}: {(idx != idx64).sum()} mismatches"


def test_tc_matches_fp32_mode_in_learn(built_lib, frames36):
    """The LBG run takes identical decisions in both search modes (same pass counts, same
    codebooks to FP64 round-off): the search arithmetic never leaks into the results."""
    from ecoz2rs_b200 import vq
    runs = {}
    for mode in (vq.MODE_TF32X3, vq.MODE_FP32):
        with vq.VqEngine(36, max_M=256, mode=mode) as eng:
            eng.set_frames(frames36)
            eng.codebook_initial()
            runs[mode], _ = eng.learn(eps=0.05, stop_M=256)
    a, b = runs[vq.MODE_TF32X3], runs[vq.MODE_FP32]
    assert [s["passes"] for s in a] == [s["passes"] for s in b]
    for sa, sb in zip(a, b):
        assert np.allclose(sa["refl"], sb["refl"], rtol=1e-9, atol=1e-12)
        assert abs(sa["avg_distortion"] - sb["avg_distortion"]) <= 1e-10 * abs(sb["avg_distortion"])


def test_tc_extreme_frame_scales(built_lib, frames36):
    """FP16 operand format: every frame is scaled by its own power of two before the split, so
    frames of any magnitude (1e-30 .. 1e+30, all-zero rows, non-autocorrelation rows whose largest
    entry is not rx[0]) are searched with the same relative accuracy; the codebook scale covers
    codewords of very different size (reflections close to +-1 next to a flat one)."""
    from ecoz2rs_b200 import vq
    from oracle import oracle as O
    rng = np.random.default_rng(99)
    n = 6000
    x = frames36[:n].copy()
    scale = 10.0 ** rng.integers(-30, 31, size=n)
    x *= scale[:, None]
    x[17] = 0.0                                   # distortion 0 against every codeword: index 0 wins
    x[100:200] = rng.normal(size=(100, 37)) * 3   # arbitrary rows: |x[k]| > |x[0]| happens
    refl = np.zeros((300, 37))
    refl[:, 1:] = np.clip(rng.normal(size=(300, 36)) * 0.3, -0.9, 0.9)
    refl[1, 1:] = 0.0                             # flat codeword: W1 = 1
    refl[2, 1:4] = [0.97, -0.97, 0.97]            # sharp resonances: large W1
    refl[3, 1:6] = 0.98
    raas = O.refl_to_raas(refl)
    idx_ref, dd_ref, d2_ref = O.assign(x, raas)
    with vq.VqEngine(36, max_M=300, mode=vq.MODE_TF32X3) as eng:
        eng.set_frames(x)
        eng.set_raas(raas)
        idx, dd, nre = eng.assign()
    bad = allowed_index_mismatches(idx, idx_ref, dd_ref, d2_ref)
    assert bad.size == 0, f"{bad.size} non-exempt mismatches, first {bad[:5]}: {idx[bad[:5]]} vs {idx_ref[bad[:5]]}"
    assert idx[17] == 0
