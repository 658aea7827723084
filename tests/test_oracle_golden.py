"""CPU tests: pin the oracle (oracle/vq_oracle.c) against the golden vectors that
tests/golden/make_golden.py produced by RUNNING THE REFERENCE (vq.learn,
vq.quantize and the reference library functions compiled from its own sources).

The reference is built with -ffast-math and sums per-thread partials, so its own
par-vs-serial noise is ~1e-12 relative; comparisons use 1e-9
unless stated.
"""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return np.load(os.path.join(GOLD, name))


def test_distortion_and_lpca_match_reference():
    from oracle import oracle as O
    g = _load("kernels.npz")
    x, raas = g["x"], g["raas"]
    d = np.array([[O.distortion(x[v], raas[i]) for i in range(raas.shape[0])] for v in range(50)])
    assert np.allclose(d, g["dist"], rtol=1e-12, atol=0)
    for v in range(20):
        code, rc, a, pe = O.lpca_r(x[v])
        assert code == g["lpca_code"][v]
        assert np.allclose(rc, g["lpca_rc"][v], rtol=1e-10, atol=1e-13)
        assert np.allclose(a, g["lpca_a"][v], rtol=1e-10, atol=1e-13)
        assert abs(pe - g["lpca_pe"][v]) <= 1e-11 * abs(pe)


def test_codebook_steps_match_reference():
    from oracle import oracle as O
    g = _load("kernels.npz")
    assert np.allclose(O.refl_to_raas(g["refl"]), g["raas"], rtol=1e-12, atol=1e-15)
    r2, a2 = O.grow(g["refl"])
    assert np.array_equal(r2, g["grow_refl"])          # x (double)0.99f / (double)1.01f exactly
    assert np.allclose(a2, g["grow_raas"], rtol=1e-12, atol=1e-15)
    ir, ia = O.initial(36)
    assert np.array_equal(ir, g["init_refl"]) and np.allclose(ia, g["init_raas"], rtol=1e-14)


def test_quantize_centroid_sigma_inertia_match_reference():
    from oracle import oracle as O
    g = _load("kernels.npz")
    seq, avg = O.quantize(g["raas"], g["x"])
    assert np.array_equal(seq, g["seq"]) and abs(avg - g["avg"]) <= 1e-12 * abs(avg)
    cells, cardd, _, _ = O.accumulate(g["x"], g["seq"].astype(np.int32),
                                      O.assign(g["x"], g["raas"])[1], g["raas"].shape[0])
    assert np.array_equal(cardd, g["cardd"]) and np.allclose(cells, g["cells"], rtol=1e-12)
    ncells, nrefl, nraas = O.centroids(g["cells"], g["cardd"], g["refl"], g["raas"])
    assert np.allclose(nrefl, g["new_refl"], rtol=1e-9, atol=1e-12)
    assert np.allclose(nraas, g["new_raas"], rtol=1e-9, atol=1e-12)
    assert np.allclose(ncells, g["cells_norm"], rtol=1e-9, atol=1e-12)
    assert abs(O.sigma(g["new_raas"], g["cells_norm"], float(g["sigma_avg"])) - g["sigma"]) <= 1e-10 * abs(g["sigma"])
    assert abs(O.inertia(g["x"], g["new_raas"]) - g["inertia"]) <= 1e-10 * abs(g["inertia"])


@pytest.mark.parametrize("tag", ["P12", "P36"])
def test_learn_matches_reference_run(tag):
    """Whole LBG run of the oracle vs the reference executable's outputs: pass counts,
    .rpt.csv values (printed with %f), every stage's .cbook reflections, cell
    cardinalities and the last stage's per-frame codeword / min distortion."""
    from oracle import oracle as O
    g = _load(f"learn_{tag}.npz")
    x = g["x"]
    rows = g["rpt_csv"]
    stages, _ = O.learn(x, eps=0.05, max_M=int(rows[-1][0]), keep_assignments=True)
    assert [(s["M"], s["passes"]) for s in stages] == [(int(r[0]), int(r[1])) for r in rows]
    for s, r in zip(stages, rows):
        assert abs(s["avg_distortion"] - r[2]) <= 1e-6          # %f has 6 decimals
        assert abs(s["sigma"] - r[3]) <= 1e-6 + 1e-9 * abs(r[3])
        assert abs(s["inertia"] - r[4]) <= 1e-6 + 1e-9 * abs(r[4])
        ref_refl = g[f"refl_{s['M']}"]
        assert np.allclose(s["refl"], ref_refl, rtol=1e-8, atol=1e-11)
        cards = g[f"cards_{s['M']}"]
        assert np.array_equal(s["cardd"], cards[:, 1].astype(np.int32))
    md = g["last_min_dists"]
    assert np.array_equal(stages[-1]["idx"], md[:, 0].astype(np.int32))
    assert np.allclose(stages[-1]["mindist"], md[:, 1], rtol=0, atol=1e-6)


def test_quantize_matches_reference_seq_file():
    from ecoz2rs_b200 import formats
    from oracle import oracle as O
    g = _load("quantize_P36.npz")
    raas = O.refl_to_raas(g["refl"])
    seq, avg = O.quantize(raas, g["x"])
    assert np.array_equal(seq, g["seq"])
    assert abs(avg - float(g["avg_printed"])) <= 1e-6
    # byte-level witness of the .seq format (header + payload) written by the reference
    import tempfile
    with tempfile.TemporaryDirectory() as tmp:
        p = os.path.join(tmp, "x.seq")
        formats.write_seq(p, seq, g["refl"].shape[0], "gold")
        with open(p, "rb") as f:
            assert f.read() == g["seq_file_bytes"].tobytes()
