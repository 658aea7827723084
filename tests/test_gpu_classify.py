"""`vq classify` through the C ABI against the reference's own output.

tests/golden/classify_P12.npz holds the stdout / stderr of the reference's vq.classify executable
(tests/golden/make_golden_classify.py) on 22 predictor files x 3 codebook models, with all inputs.
The B200 path keeps every frame resident and runs one search pass per model; its stdout must be
the reference's text byte for byte (model list, rankings with %e distortions, confusion matrix,
candidate-order table), its stderr the same '*' / '!' marks and skip message."""
import os
import re

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden", "classify_P12.npz")
_ANSI = re.compile(r"\x1b\[[0-9;]*m")


def _capture_fds(fn):
    """Run fn() with the process-level stdout / stderr (fd 1, 2) redirected to files."""
    import tempfile
    outs = []
    saved = [os.dup(1), os.dup(2)]
    files = [tempfile.TemporaryFile(), tempfile.TemporaryFile()]
    try:
        os.dup2(files[0].fileno(), 1)
        os.dup2(files[1].fileno(), 2)
        rc = fn()
    finally:
        os.dup2(saved[0], 1)
        os.dup2(saved[1], 2)
        os.close(saved[0])
        os.close(saved[1])
    for f in files:
        f.seek(0)
        outs.append(f.read().decode())
        f.close()
    return rc, outs[0], outs[1]


@pytest.mark.parametrize("mode", ["tf32x3", "fp64"])
def test_classify_matches_reference_output(built_lib, tmp_path, monkeypatch, mode):
    from ecoz2rs_b200 import formats, vq
    g = np.load(GOLD)
    classes = [str(c) for c in g["classes"]]
    cb_files = [str(f) for f in g["cb_files"]]
    prd_files = [str(f) for f in g["prd_files"]]
    for ci, rel in enumerate(cb_files):
        formats.write_cbook(str(tmp_path / rel), g[f"refl_{ci}"], classes[ci])
    for k, rel in enumerate(prd_files):
        cls = rel.split("/")[1]
        formats.write_prd(str(tmp_path / rel), g[f"x_{k}"], cls)
    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("ECOZ2_VQ_MODE", mode)
    rc, out, err = _capture_fds(lambda: vq.vq_classify(cb_files, prd_files, show_ranked=True))
    assert rc == 0
    want_out = bytes(g["stdout"]).decode()
    want_err = bytes(g["stderr"]).decode()
    assert out == want_out
    assert _ANSI.sub("", err) == _ANSI.sub("", want_err)
    assert err.count("\x1b[32m*") == want_err.count("\x1b[32m*")


def test_classify_errors(built_lib, tmp_path, monkeypatch):
    """Missing model -> 1 (vq_classify.c:221-225); codebooks of different sizes -> conformity
    error, 1 (vq_classify.c:239-242)."""
    from ecoz2rs_b200 import formats, vq
    g = np.load(GOLD)
    monkeypatch.chdir(tmp_path)
    formats.write_cbook("a.cbook", g["refl_0"], "alpha")
    formats.write_cbook("b.cbook", g["refl_1"][:16], "bravo")
    formats.write_prd("x.prd", g["x_0"], "alpha")
    rc, _, err = _capture_fds(lambda: vq.vq_classify(["nope.cbook"], ["x.prd"]))
    assert rc == 1 and "error loading model" in err
    rc, _, err = _capture_fds(lambda: vq.vq_classify(["a.cbook", "b.cbook"], ["x.prd"]))
    assert rc == 1 and "conformity error" in err


@pytest.mark.parametrize("P,K,R", [(36, 64, 5), (36, 128, 3), (36, 256, 7), (12, 32, 9), (36, 512, 2), (36, 16, 3),
                                   (36, 64, 40)],
                         ids=["K64xR5", "K128xR3", "K256xR7", "P12-K32xR9", "K512xR2-serial", "K16xR3-serial", "K64xR40"])
def test_quantize_models_matches_one_model_at_a_time(built_lib, P, K, R):
    """ecoz2b200_quantize_models (cbook_quantize of every file against R codebooks, vq/vq.c:154-157,
    285-315): the one-sweep search over the concatenated codebooks gives, for every model, the symbols
    and per-file distortions of quantizing with that model alone -- which is the oracle's quantize
    (quantize.c:8-34) on the frames of each file."""
    from ecoz2rs_b200 import vq
    from ecoz2rs_b200.synth import ar_mix
    from oracle import oracle as O
    from conftest import allowed_index_mismatches
    rng = np.random.default_rng(K + R)
    x = ar_mix(30_011, P, seed=K * 7 + R)
    # R different codebooks: learnt on different slices of other data
    models = []
    for r in range(R):
        y = ar_mix(4000, P, seed=1000 + 13 * r + K)
        st, _ = O.learn(y, eps=0.05, max_M=K)
        refl = st[-1]["refl"]
        assert refl.shape[0] == K
        models.append(refl)
    refl_all = np.stack(models)
    cuts = np.array([0, 1, 700, 700, 12_345, 30_011])                       # one-frame and empty files
    with vq.VqEngine(P, max_M=max(K, 2)) as eng:
        eng.set_frames(x)
        seq, dists, one = eng.quantize_models(refl_all, cuts)
        assert one == (K in (32, 64, 128, 256))
        c = eng.counters()
    assert seq.shape == (R, len(x)) and dists.shape == (R, len(cuts) - 1)
    for r in range(0, R, max(1, R // 6)):
        raas = O.refl_to_raas(models[r])
        ref_idx, ref_dd, ref_second = O.assign(x, raas)
        assert allowed_index_mismatches(seq[r].astype(np.int64), ref_idx, ref_dd, ref_second).size == 0
        for f in range(len(cuts) - 1):
            a, b = cuts[f], cuts[f + 1]
            want = (ref_dd[a:b] - 1.0).sum() / (b - a) if b > a else 0.0
            assert abs(dists[r, f] - want) <= 1e-9 * max(1.0, abs(want)), (r, f)
    if one:
        assert c["rechecked"] > 0                                           # the FP64 scan of ambiguous pairs ran
