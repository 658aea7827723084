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
    # the north_star target size: 16 codeword tiles of 256 columns, the highest flag rate
    (36, 4096, 20000), (12, 4096, 20000),
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


def _k4096_codebook(x, P):
    """A K = 4096 codebook that looks like a learnt one: an LBG-trained 256-codeword codebook of the
    frames, split four times (grow: x 0.99f / x 1.01f on every reflection) -- dense clusters of
    near-identical codewords, the worst case for the best-two gap."""
    from oracle import oracle as O
    stages, _ = O.learn(x, eps=0.05, max_M=256)
    refl = stages[-1]["refl"]
    while refl.shape[0] < 4096:
        refl, _ = O.grow(refl)
    return refl


def test_k4096_pass_and_quantize_match_oracle(built_lib, frames36):
    """configs[2] / configs[3] codebook size against the oracle: teacher-forced k-means pass
    (indices, counts, cell sums, DD, updated codebook) and quantize (symbols, average distortion),
    with the 1e-6 best-two exemption (vq_learn_par.c:62-111, quantize.c:14-30)."""
    from ecoz2rs_b200 import vq
    from oracle import oracle as O
    x = frames36
    refl = _k4096_codebook(x, 36)
    raas = O.refl_to_raas(refl)
    ref = O.kmeans_pass(x, refl, raas)
    seq_ref, avg_ref = O.quantize(raas, x)
    with vq.VqEngine(36, max_M=4096, mode=vq.MODE_TF32X3) as eng:
        eng.set_frames(x)
        eng.set_reflections(refl)
        out = eng.kmeans_pass()
        nrefl, nraas = eng.codebook()
        idx, md = eng.last_assignment()
        eng.set_reflections(refl)
        seq, avg = eng.quantize(x)
        seq2, avg2 = eng.quantize_resident()
    bad = allowed_index_mismatches(idx, ref["idx"], ref["ddmin"], ref["second"])
    assert bad.size == 0, f"{bad.size} non-exempt mismatches, first {bad[:5]}"
    for s_, a_ in ((seq, avg), (seq2, avg2)):
        assert allowed_index_mismatches(s_.astype(np.int32), seq_ref.astype(np.int32), ref["ddmin"],
                                        ref["second"]).size == 0
        assert abs(a_ - avg_ref) <= 1e-9 * abs(avg_ref)
    assert int(out["cardd"].sum()) == len(x)
    if np.array_equal(idx, ref["idx"]):
        assert np.array_equal(out["cardd"], ref["cardd"])
        assert np.allclose(out["cells"], ref["cells"], rtol=1e-9, atol=1e-9)
        assert abs(out["DD"] - ref["DD"]) <= 1e-9 * abs(ref["DD"])
        assert np.allclose(md, ref["ddmin"] - 1.0, rtol=0, atol=1e-11)
        assert np.allclose(nrefl, ref["refl"], rtol=1e-7, atol=1e-10)
        assert np.allclose(nraas, ref["raas"], rtol=1e-7, atol=1e-10)
    assert abs(out["DD"] - ref["DD"]) <= 1e-4 * abs(ref["DD"])


# ---------------------------------------------------------------------------------------------
# The error bound of the tensor-core values ("error bound"): for EVERY
# (frame, codeword) pair
#     |d_tc - d_fp64| <= kappa 2^-24 x_ref W1_j + D 2^-25 x_ref / G
# with x_ref = max_k |x_k|, W1_j = sum_k |w_jk|, G = the codebook's power-of-two scale.
def _bound_ratio(x, raas, d_tc, kappa=KAPPA_TC):
    w = _w(raas)
    D = x.shape[1]
    ref = x @ w.T                                     # float64: rounding ~1e-16 sum|x w|
    x_ref = np.abs(x.astype(np.float32).astype(np.float64)).max(axis=1)[:, None]
    W1 = np.abs(w).sum(axis=1)[None, :]
    g_small = 2.0 ** (12 - np.floor(np.log2(W1.max() * 1.0002)))   # <= the kernel's G
    bound = kappa * U24 * x_ref * W1 + D * 2.0 ** -25 * x_ref / g_small
    err = np.abs(d_tc.astype(np.float64) - ref)
    return err / np.maximum(bound, 1e-300), err / np.maximum(U24 * x_ref * W1, 1e-300)


def _adversarial_frames(n, D, rng):
    fam = []
    m = n // 4
    # (a) every |x_k| close to the frame's maximum, alternating / random signs: after the per-frame
    #     scaling all elements sit in [2^14, 2^15): the largest products, the most cancellation
    a = (1.0 - rng.uniform(0, 2.0 ** -9, size=(m, D))) * np.where(rng.random((m, D)) < 0.5, -1.0, 1.0)
    a[: m // 2] = np.abs(a[: m // 2]) * ((-1.0) ** np.arange(D))[None, :]
    fam.append(a * 10.0 ** rng.integers(-3, 4, size=(m, 1)))
    # (b) low parts at the FP16 subnormal edge: x_0 = 1 (scale 2^14), the other elements are FP16-exact
    #     values in [2^-7, 2^-6) plus a residual that lands in [2^-15, 2^-13] after scaling
    b = np.zeros((m, D))
    b[:, 0] = 1.0
    mant = rng.integers(512, 1024, size=(m, D - 1))
    rho = 2.0 ** -28 * (1.0 + rng.random((m, D - 1))) * rng.choice([0.5, 1.0, 2.0], size=(m, D - 1))
    b[:, 1:] = (mant * 2.0 ** -16 + rho) * np.where(rng.random((m, D - 1)) < 0.5, -1.0, 1.0)
    fam.append(b)
    # (c) heavy-tailed magnitudes (elements down to 2^-20 of the maximum)
    c = np.where(rng.random((m, D)) < 0.5, -1.0, 1.0) * 2.0 ** (-rng.uniform(0, 20, size=(m, D)))
    c[:, 0] = 1.0
    fam.append(c)
    # (d) values one rounding step below / above FP16 and FP32 grid points
    d = rng.integers(1024, 2048, size=(n - 3 * m, D)) * 2.0 ** -11
    d = d * (1.0 + rng.choice([-1.0, 1.0], size=d.shape) * (2.0 ** -12 + 2.0 ** -25))
    d *= np.where(rng.random(d.shape) < 0.5, -1.0, 1.0)
    fam.append(d)
    return np.concatenate(fam)


def _adversarial_raas(K, D, rng, realistic):
    """raas rows [K, D] (w_0 = raas_0, w_k = 2 raas_k): alternating-sign rows of equal magnitude,
    heavy-tailed rows (low parts in the FP16 subnormals after the codebook scaling), rows that
    cancel against the (a) frames, and real codewords."""
    r = np.empty((K, D))
    q = K // 4
    r[:q] = (1.0 - rng.uniform(0, 2.0 ** -9, size=(q, D))) * ((-1.0) ** np.arange(D))[None, :] * \
        np.where(rng.random((q, 1)) < 0.5, -1.0, 1.0)
    r[q:2 * q] = np.where(rng.random((q, D)) < 0.5, -1.0, 1.0) * 2.0 ** (-rng.uniform(0, 24, size=(q, D)))
    r[2 * q:3 * q] = rng.integers(1024, 2048, size=(q, D)) * 2.0 ** -11 * \
        (1.0 + rng.choice([-1.0, 1.0], size=(q, D)) * (2.0 ** -12 + 2.0 ** -25)) * \
        np.where(rng.random((q, D)) < 0.5, -1.0, 1.0)
    r[: 3 * q, 1:] *= 0.5
    r[3 * q:] = realistic[: K - 3 * q]
    return r


@pytest.mark.parametrize("P", [36, 12], ids=["P36", "P12"])
def test_tc_error_bound_adversarial(built_lib, frames36, frames12, P):
    """K = 4096, operands built to maximise cancellation, elements at the top of the scaled range,
    low parts at the FP16 subnormal edge: the bound holds for EVERY pair, with margin."""
    from ecoz2rs_b200 import vq
    from oracle import oracle as O
    rng = np.random.default_rng(4096 + P)
    D = P + 1
    base = frames36 if P == 36 else frames12
    realistic = O.refl_to_raas(_k4096_codebook(base[:20000], P))
    x = np.concatenate([_adversarial_frames(1536, D, rng), base[:512]])
    raas = _adversarial_raas(4096, D, rng, realistic)
    with vq.VqEngine(P, max_M=4096, mode=vq.MODE_TENSOR) as eng:
        eng.set_frames(x)
        eng.set_raas(raas)
        d = eng.debug_tc_distortions(len(x))
    ratio, in_u24 = _bound_ratio(x, raas, d)
    worst = float(ratio.max())
    print(f"[tc bound P={P}] worst pair at {worst:.3f} of the bound; largest error {in_u24.max():.2f} u24 x_ref W1_j")
    assert worst <= 0.75, f"tensor-core error reaches {worst:.3f} of the kappa = {KAPPA_TC} bound"


def test_tc_accumulation_error_alone(built_lib):
    """Operands that are exactly representable in FP16 after the scalings (zero low parts, zero
    cross terms): what is left is the FP32 accumulation inside and across the chained MMAs -- the
    one term of the bound that rests on a hardware property (alpha <= 4 per MMA assumed,
    3 alpha u24 in total; the measured total must stay below it)."""
    from ecoz2rs_b200 import vq
    rng = np.random.default_rng(77)
    P, D, K, n = 36, 37, 4096, 1024
    # frames: integers in [1024, 2048) / 2048 with signs, x_0 = exactly the maximum
    x = rng.integers(1024, 2048, size=(n, D)) / 2048.0 * np.where(rng.random((n, D)) < 0.5, -1.0, 1.0)
    x[:, 0] = 2047.0 / 2048.0
    x[: n // 2] = np.abs(x[: n // 2]) * ((-1.0) ** np.arange(D))[None, :]
    w = rng.integers(1024, 2048, size=(K, D)) / 2048.0 * np.where(rng.random((K, D)) < 0.5, -1.0, 1.0)
    w[: K // 2] = np.abs(w[: K // 2])
    raas = w.copy()
    raas[:, 1:] *= 0.5
    with vq.VqEngine(P, max_M=K, mode=vq.MODE_TENSOR) as eng:
        eng.set_frames(x)
        eng.set_raas(raas)
        d = eng.debug_tc_distortions(n).astype(np.float64)
    ref = x @ w.T
    s_abs = np.abs(x) @ np.abs(w).T
    # the dump converts back to the caller's units with one FP32 multiply: 1 u24 of |d| on top
    acc_err = np.abs(d - ref) / (U24 * s_abs)
    print(f"[tc accumulation] worst {acc_err.max():.2f} u24 of sum |x_k w_k| (budget 12 + 1)")
    assert acc_err.max() <= 13.0
