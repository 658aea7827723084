"""GPU tests of the drop-in boundary: the three `ecoz2_vq_*` entry points (and the
`ecoz2` CLI on top of them) run on `.prd` files and their output files are compared
with what the REFERENCE executables wrote for the same inputs (tests/golden)."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def _read_rpt_csv(path):
    rows = []
    for line in open(path, encoding="utf-8"):
        if line.startswith("#") or line.startswith("M,"):
            continue
        p = line.strip().split(",")
        rows.append((int(p[0]), int(p[1]), float(p[2]), float(p[3]), float(p[4])))
    return rows


@pytest.fixture()
def in_tmp_cwd(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    return tmp_path


@pytest.mark.parametrize("tag,P", [("P36", 36), ("P12", 12)])
def test_vq_learn_files_match_reference(built_lib, in_tmp_cwd, monkeypatch, tag, P):
    from ecoz2rs_b200 import formats, vq
    g = np.load(os.path.join(GOLD, f"learn_{tag}.npz"))
    x = g["x"]
    # two .prd files: the learner concatenates them in argument order (vq_learn.i:198-217)
    half = len(x) // 2
    f1 = str(in_tmp_cwd / "predictors" / "TRAIN" / "gold" / "a.prd")
    f2 = str(in_tmp_cwd / "predictors" / "TRAIN" / "gold" / "b.prd")
    formats.write_prd(f1, x[:half], "gold")
    formats.write_prd(f2, x[half:], "gold")
    monkeypatch.setenv("ECOZ2_VQ_MAX_M", "256")
    calls = []
    rc = vq.vq_learn(P, 0.05, "gold", [f1, f2], callback=lambda M, a, s, i: calls.append((M, a, s, i)))
    assert rc == 0
    cdir = in_tmp_cwd / "data" / "codebooks" / "gold"
    rows = _read_rpt_csv(cdir / "eps_0.05.rpt.csv")
    ref_rows = g["rpt_csv"]
    assert [(r[0], r[1]) for r in rows] == [(int(r[0]), int(r[1])) for r in ref_rows]
    assert [c[0] for c in calls] == [r[0] for r in rows]
    for r, rr, c in zip(rows, ref_rows, calls):
        # the report prints 6 decimals: equal to the last printed digit (the path's stated tolerance
        # is 1e-4; what it achieves is the print resolution)
        for col in (2, 3, 4):
            assert abs(r[col] - rr[col]) <= 1e-7 * abs(rr[col]) + 1.01e-6, (r, rr)
        assert abs(c[1] - rr[2]) <= 1e-7 * rr[2] + 1.01e-6   # callback carries the same values
        M = r[0]
        cls, refl = formats.read_cbook(str(cdir / f"eps_0.05_M_{M:04d}.cbook"))
        assert cls == "gold"
        ref_refl = g[f"refl_{M}"]
        scale = np.abs(ref_refl).max(axis=1, keepdims=True)
        assert np.all(np.abs(refl - ref_refl) <= 1e-7 * scale)   # stated tolerance 1e-4
        assert np.all(refl[:, 0] == 0)                        # reflections[.][0] stays 0
        cards = np.genfromtxt(cdir / f"eps_0.05_M_{M:04d}.cbook.cards_dists.csv", delimiter=",",
                              skip_header=1).reshape(-1, 3)
        assert np.array_equal(cards[:, 1], g[f"cards_{M}"][:, 1])            # cardinalities: exact
        assert np.allclose(cards[:, 2], g[f"cards_{M}"][:, 2], rtol=1e-7, atol=1.01e-6)
    md = np.genfromtxt(cdir / f"eps_0.05_M_{rows[-1][0]:04d}.cbook.min_dists.csv", delimiter=",",
                       skip_header=1)
    ref_md = g["last_min_dists"]
    assert np.array_equal(md[:, 0], ref_md[:, 0])             # every frame's codeword
    assert np.allclose(md[:, 1], ref_md[:, 1], rtol=0, atol=1.01e-6)   # "%f": 6 decimals
    # header lines of the report files are the reference's
    head = open(cdir / "eps_0.05.rpt.csv", encoding="utf-8").read().splitlines()[:2]
    ref_head = g["rpt_csv_text"].tobytes().decode("utf-8").splitlines()[:2]
    assert head == ref_head
    # the text report (report.c:14-75): the reference's own file for the same frames, token by token --
    # words, paths, pass counts and cardinalities exact, "%f" values to the last printed digit,
    # "%8.3g" cell distortions to their three digits; "(n empty cells)" / "_" would have to agree too
    ours = open(cdir / "eps_0.05.rpt", encoding="utf-8").read().split("\n")
    theirs = open(os.path.join(GOLD, f"learn_{tag}.rpt.txt"), encoding="utf-8").read().split("\n")
    assert len(ours) == len(theirs)
    for lo, lt in zip(ours, theirs):
        to, tt = lo.split(), lt.split()
        assert len(to) == len(tt), (lo, lt)
        g3 = not lt.startswith("###") and "." in lt       # the line of %8.3g values
        for a, b in zip(to, tt):
            if a == b:
                continue
            fa, fb = float(a), float(b)                   # (a word that differs fails here)
            if g3:
                # (a cell of one frame sitting on its own codeword: rounding noise around 0 on both
                # sides, 2.84e-14 against -2.84e-14 in this fixture)
                assert abs(fa - fb) <= 1.01e-2 * abs(fb) + 1e-9, (lo, lt)
            else:
                assert "." in b and abs(fa - fb) <= 1e-7 * abs(fb) + 1.01e-6, (lo, lt)
        assert lo.startswith("###") == lt.startswith("###")
    # column layout of the "%8d " fields
    card = lambda ls: [len(l) for l in ls if l.startswith(" ") and "." not in l]
    assert card(ours) == card(theirs)


def test_vq_learn_from_base_codebook_and_errors(built_lib, in_tmp_cwd, monkeypatch):
    from ecoz2rs_b200 import formats, vq
    g = np.load(os.path.join(GOLD, "learn_P36.npz"))
    f1 = str(in_tmp_cwd / "predictors" / "TRAIN" / "gold" / "a.prd")
    formats.write_prd(f1, g["x"], "gold")
    base = str(in_tmp_cwd / "base.cbook")
    formats.write_cbook(base, g["refl_16"], "gold")
    monkeypatch.setenv("ECOZ2_VQ_MAX_M", "64")
    monkeypatch.setenv("ECOZ2_VQ_MIN_DISTS", "0")
    assert vq.vq_learn_using_base_codebook(base, 0.05, [f1]) == 0
    cdir = in_tmp_cwd / "data" / "codebooks" / "gold"
    assert sorted(p.name for p in cdir.glob("*.cbook")) == ["eps_0.05_M_0032.cbook", "eps_0.05_M_0064.cbook"]
    assert not list(cdir.glob("*.min_dists.csv"))
    for M in (32, 64):
        _, refl = formats.read_cbook(str(cdir / f"eps_0.05_M_{M:04d}.cbook"))
        ref = g[f"refl_{M}"]
        assert np.all(np.abs(refl - ref) <= 1e-4 * np.abs(ref).max(axis=1, keepdims=True))
    # error codes: missing predictor -> 2 ; missing base codebook -> 2 (vq_learn.i:203-206, vq_learn.c:49-52)
    assert vq.vq_learn(36, 0.05, "gold", [str(in_tmp_cwd / "nope.prd")]) == 2
    assert vq.vq_learn_using_base_codebook(str(in_tmp_cwd / "nope.cbook"), 0.05, [f1]) == 2
    assert vq.vq_quantize(str(in_tmp_cwd / "nope.cbook"), [f1]) == 1


def test_parallel_prd_loader_equals_serial(built_lib, in_tmp_cwd, monkeypatch):
    """The loader threads of ecoz2_vq_learn (vq_hostio.cpp: several files and ranges in flight, chunks
    placed by their global frame position) against the one-file-at-a-time fread path
    (prd.c:34-74 order): ragged files, one of a single frame, an empty one; a truncated file is an
    error, not a short read."""
    import shutil
    from ecoz2rs_b200 import formats, vq
    from ecoz2rs_b200.synth import ar_mix
    x = ar_mix(200_000, 36, seed=77)
    cuts = [0, 1, 60_001, 60_001, 131_072, 199_990, 200_000]          # 1 frame, ragged, EMPTY, chunk-sized ...
    files = []
    for k in range(len(cuts) - 1):
        f = str(in_tmp_cwd / "predictors" / "TRAIN" / "par" / f"part{k}.prd")
        formats.write_prd(f, x[cuts[k]:cuts[k + 1]], "par")
        files.append(f)
    monkeypatch.setenv("ECOZ2_VQ_MAX_M", "32")
    monkeypatch.setenv("ECOZ2_VQ_MIN_DISTS", "1")
    cdir = in_tmp_cwd / "data" / "codebooks" / "par"
    outs = {}
    for mode in ("parallel", "serial"):
        if mode == "serial":
            monkeypatch.setenv("ECOZ2_VQ_LOAD_SERIAL", "1")
        monkeypatch.setenv("ECOZ2_VQ_IO_THREADS", "5")
        assert vq.vq_learn(36, 0.05, "par", files) == 0
        outs[mode] = {p_.name: p_.read_bytes() for p_ in sorted(cdir.iterdir())}
        shutil.rmtree(cdir)
    assert sorted(outs["parallel"]) == sorted(outs["serial"]) and len(outs["serial"]) >= 5 * 2 + 2
    # (two runs are not bit-identical: the cell sums end in FP64 atomics whose order varies -- the
    # codebooks agree to rounding, the per-frame report, which lists the frames in FILE order, to its digits)
    import io
    for name, data in outs["serial"].items():
        other = outs["parallel"][name]
        if name.endswith(".cbook"):
            ra = np.frombuffer(data[120:], dtype="<f8")        # ident 16 + class 96 + P, M
            rb = np.frombuffer(other[120:], dtype="<f8")
            assert data[:120] == other[:120] and np.allclose(ra, rb, rtol=0, atol=1e-11), name
        elif name.endswith(".min_dists.csv"):
            ta = np.loadtxt(io.BytesIO(data), delimiter=",", skiprows=1)
            tb = np.loadtxt(io.BytesIO(other), delimiter=",", skiprows=1)
            assert ta.shape == tb.shape == (200_000, 2), name
            assert np.count_nonzero(ta[:, 0] != tb[:, 0]) <= 2, name             # FP64-level ties only
            assert np.count_nonzero(np.abs(ta[:, 1] - tb[:, 1]) > 1.5e-6) <= 2, name
        else:
            assert len(other) == len(data), name
    # truncated data section
    bad = str(in_tmp_cwd / "predictors" / "TRAIN" / "par" / "bad.prd")
    raw = open(files[1], "rb").read()
    open(bad, "wb").write(raw[: len(raw) - 100])
    monkeypatch.delenv("ECOZ2_VQ_LOAD_SERIAL")
    assert vq.vq_learn(36, 0.05, "par", [files[0], bad, files[2]]) == 2


def test_vq_quantize_seq_bytes_match_reference(built_lib, in_tmp_cwd):
    from ecoz2rs_b200 import formats, vq
    q = np.load(os.path.join(GOLD, "quantize_P36.npz"))
    prd = str(in_tmp_cwd / "predictors" / "TRAIN" / "gold" / "x.prd")
    formats.write_prd(prd, q["x"], "gold")
    cb = str(in_tmp_cwd / "cb.cbook")
    formats.write_cbook(cb, q["refl"], "gold")
    assert vq.vq_quantize(cb, [prd], show_filenames=False) == 0
    M = q["refl"].shape[0]
    out = in_tmp_cwd / "data" / "sequences" / "TRAIN" / f"M{M}" / "gold" / "x.seq"
    assert out.read_bytes() == q["seq_file_bytes"].tobytes()


def test_vq_quantize_batches_of_files(built_lib, in_tmp_cwd, capfd):
    """ecoz2_vq_quantize keeps a whole batch of files resident (parallel loader, one sweep, per-file
    distortions from the device): every `.seq` equals the oracle's quantize of that file alone
    (quantize.c:8-34), the per-file distortions printed with -s and the total line are the
    reference's; a file that cannot be loaded ends the run with code 2 AFTER the files before it were
    written (vq_quantize.c:49-53)."""
    from ecoz2rs_b200 import formats, vq
    from ecoz2rs_b200.synth import ar_mix
    from oracle import oracle as O
    q = np.load(os.path.join(GOLD, "quantize_P36.npz"))
    refl = q["refl"]
    M = refl.shape[0]
    raas = O.refl_to_raas(refl)
    cb = str(in_tmp_cwd / "cb.cbook")
    formats.write_cbook(cb, refl, "gold")
    sizes = [1, 50, 4097, 20_000, 3, 1024]
    files, xs = [], []
    for k, T in enumerate(sizes):
        x = ar_mix(T, 36, seed=40 + k)
        f = str(in_tmp_cwd / "predictors" / "TEST" / f"c{k % 2}" / f"f{k}.prd")
        formats.write_prd(f, x, f"c{k % 2}")
        files.append(f)
        xs.append(x)
    assert vq.vq_quantize(cb, files, show_filenames=True) == 0
    out = capfd.readouterr().out
    total = 0.0
    for k, (f, x) in enumerate(zip(files, xs)):
        seq_ref, dist_ref = O.quantize(raas, x)
        cls, m, seq = formats.read_seq(str(in_tmp_cwd / "data" / "sequences" / "TEST" / f"M{M}" / f"c{k % 2}" / f"f{k}.seq"))
        assert (cls, m) == (f"c{k % 2}", M) and np.array_equal(seq, seq_ref), k
        assert f"f{k}.seq className='c{k % 2}' ({dist_ref:1.4g})" in out, (k, dist_ref)
        total += dist_ref
    assert f"total: {len(files)} sequences; total  average distortion = {total / len(files):f}" in out
    # a truncated file in the middle
    raw = open(files[3], "rb").read()
    bad = str(in_tmp_cwd / "predictors" / "TEST" / "c0" / "bad.prd")
    open(bad, "wb").write(raw[:-64])
    import shutil
    shutil.rmtree(in_tmp_cwd / "data")
    assert vq.vq_quantize(cb, [files[0], files[2], bad, files[4]]) == 2
    seqs = sorted(p_.name for p_ in (in_tmp_cwd / "data").rglob("*.seq"))
    assert seqs == ["f0.seq", "f2.seq"]


def test_cli_learn_and_quantize(built_lib, in_tmp_cwd):
    """`ecoz2 vq learn` / `ecoz2 vq quantize` with the reference's option names."""
    from ecoz2rs_b200 import formats
    q = np.load(os.path.join(GOLD, "quantize_P36.npz"))
    d = in_tmp_cwd / "predictors" / "TEST" / "gold"
    formats.write_prd(str(d / "x.prd"), q["x"], "gold")
    exe = os.path.join(ROOT, "ecoz2rs_b200", "csrc", "ecoz2")
    env = dict(os.environ, ECOZ2_VQ_MAX_M="8", ECOZ2_VQ_MIN_DISTS="0")
    r = subprocess.run([exe, "vq", "learn", "-P", "36", "-e", "0.05", "--class-name", "gold",
                        "--predictors", str(in_tmp_cwd / "predictors")], env=env, capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Ecoz2ObserverRef.step: M=8" in r.stdout
    cb = in_tmp_cwd / "data" / "codebooks" / "gold" / "eps_0.05_M_0008.cbook"
    assert cb.exists()
    r = subprocess.run([exe, "vq", "quantize", "--codebook", str(cb), "--predictors", str(d / "x.prd")],
                       env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "number of predictor files: 1" in r.stdout
    cls, M, seq = formats.read_seq(str(in_tmp_cwd / "data" / "sequences" / "TEST" / "M8" / "gold" / "x.seq"))
    assert (cls, M, len(seq)) == ("gold", 8, len(q["x"]))


def test_device_report_formatter_matches_printf(built_lib):
    """vq_report.cu: "%d,%f\\n" lines formatted on the device are byte for byte what the
    reference's fprintf writes (report.c:118-121): random magnitudes, exact ties of the 6th decimal
    (round-half-to-even), negatives, -0.0, subnormals, values up to 2^62; inf / nan / >= 2^62 are
    refused (the file writer then formats on the host)."""
    from ecoz2rs_b200 import vq
    rng = np.random.default_rng(7)
    n = 300_000
    md = np.concatenate([
        rng.random(n) * 10.0 ** rng.integers(-9, 6, size=n),
        rng.integers(0, 2 ** 24, size=5000) / 128.0,
        rng.integers(0, 2 ** 30, size=5000) / 2.0 ** 20,
        -rng.random(2000),
        rng.random(2000) * 10.0 ** rng.integers(6, 18, size=2000),
        np.array([0.0, -0.0, 5e-324, 1e-310, 0.5e-6, 1.5e-6, 2.5e-6, 0.9999995, 0.99999949999,
                  1e15, 123456789012345.678, 4.6e18, 2.0 ** 62 * (1 - 2.0 ** -53), 2.0 ** 52 + 1, 2.0 ** 53]),
    ])
    cw = rng.integers(0, 65536, size=md.shape[0]).astype(np.int32)
    cw[:3] = [0, -1, 2 ** 31 - 1]
    rc, text = vq.format_lines_on_device(cw, md)
    assert rc == 0
    got = text.decode().split("\n")
    assert got[-1] == "" and len(got) == md.shape[0] + 1
    want = ["%d,%f" % (int(c), float(x)) for c, x in zip(cw, md)]
    bad = [i for i, (a, b) in enumerate(zip(got[:-1], want)) if a != b]
    assert not bad, f"{len(bad)} lines differ, first: {got[bad[0]]!r} vs {want[bad[0]]!r} ({md[bad[0]]!r})"
    for x in (np.inf, -np.inf, np.nan, 2.0 ** 62, 1e19, 1e300):
        rc, _ = vq.format_lines_on_device(np.array([1], np.int32), np.array([x]))
        assert rc != 0, x


def test_min_dists_text_of_a_pass(built_lib, frames36, oracle_run36):
    """The engine's report text of a pass == "%d,%f" of its last_assignment."""
    from ecoz2rs_b200 import vq
    stages, _ = oracle_run36
    with vq.VqEngine(36, max_M=256) as eng:
        eng.set_frames(frames36)
        eng.set_reflections(stages[6]["refl"])
        eng.kmeans_pass(want_cells=False)
        idx, md = eng.last_assignment()
        t1 = eng.min_dists_text()
        t2 = eng.min_dists_text()          # second pinned buffer, same bytes
    want = "".join("%d,%f\n" % (int(c), float(x)) for c, x in zip(idx, md)).encode()
    assert t1 == want and t2 == want
