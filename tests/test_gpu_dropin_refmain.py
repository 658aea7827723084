"""Drop-in proof with the reference's OWN callers: the unmodified `main()` of
ecoz2/src/x/vq.learn.c (:107-128) and ecoz2/src/x/vq.quantize.c (:41), compiled from
/root/reference and linked against libecoz2_vq_b200.so -- which exports the inner API they call
(vq_learn, vq_learn_using_base_codebook, vq_quantize: ecoz2/src/include/vq.h:82-107) -- by
oracle/Makefile target `dropin`, run next to the reference's own executables
on config 1's `.prd` shards: same stages and passes, `.rpt.csv` values, every `.cbook`, and `.seq`
files byte for byte; `-B` (resume from a base codebook) through the reference's getopt too."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N, P, STOP = 100_000, 36, 256


def test_reference_mains_linked_against_the_b200_library(built_lib, tmp_path):
    from ecoz2rs_b200 import formats
    from ecoz2rs_b200.synth import ar_mix
    from oracle import oracle as O
    from oracle import refbuild
    if not (refbuild.available(STOP) and refbuild.dropin_available()):
        pytest.skip("oracle/_ref (reference executables / drop-in executables) not built")
    x = ar_mix(N, P, seed=20260101)
    ref_dir, gpu_dir = tmp_path / "ref", tmp_path / "gpu"
    files = {}
    for d in (ref_dir, gpu_dir):
        fl = []
        for k in range(2):
            f = d / "predictors" / "TRAIN" / "cfg1" / f"part_{k:05d}.prd"
            formats.write_prd(str(f), x[k * (N // 2):(k + 1) * (N // 2)], "cfg1")
            fl.append(str(f))
        files[d] = fl
    ref = refbuild.run_vq_learn(files[ref_dir], str(ref_dir), P=P, eps=0.05, class_name="cfg1", max_M=STOP)
    if ref["returncode"] != 0:
        pytest.skip("reference vq.learn failed on this host: " + ref["stdout"][-400:])
    env = {"ECOZ2_VQ_MAX_M": str(STOP)}
    got = refbuild.run_dropin("vq.learn", ["-P", str(P), "-w", "cfg1", "-e", "0.05"] + files[gpu_dir],
                              str(gpu_dir), env)
    assert got["returncode"] == 0, got["stdout"][-2000:]
    assert "B200 CUDA path" in got["stdout"]            # the library ran, not the reference's loop
    cbdir = os.path.join("data", "codebooks", "cfg1")
    ref_rows = refbuild.read_rpt_csv(str(ref_dir / cbdir / "eps_0.05.rpt.csv"))
    rows = refbuild.read_rpt_csv(str(gpu_dir / cbdir / "eps_0.05.rpt.csv"))
    assert [(r[0], r[1]) for r in rows] == [(r[0], r[1]) for r in ref_rows]
    for r, rr in zip(rows, ref_rows):
        assert abs(r[2] - rr[2]) <= 2e-6 and abs(r[3] - rr[3]) <= 1e-6 * abs(rr[3]) + 2e-6   # printed "%f"
        assert abs(r[4] - rr[4]) <= 1e-7 * abs(rr[4]) + 2e-6
        name = f"eps_0.05_M_{r[0]:04d}.cbook"
        ha, a = formats.read_cbook(str(gpu_dir / cbdir / name))
        hb, b = formats.read_cbook(str(ref_dir / cbdir / name))
        assert ha == hb
        assert np.allclose(a, b, rtol=1e-7, atol=1e-9), name
        for ext in (".cards_dists.csv", ".min_dists.csv"):
            assert (gpu_dir / cbdir / (name + ext)).exists() == (ref_dir / cbdir / (name + ext)).exists()

    # vq.quantize: same codebook (the reference's), `.seq` bytes
    cb = str(ref_dir / cbdir / f"eps_0.05_M_{STOP:04d}.cbook")
    q_ref = refbuild.run_vq_quantize(cb, files[ref_dir], str(ref_dir))
    q_got = refbuild.run_dropin("vq.quantize", [cb] + files[gpu_dir], str(gpu_dir))
    assert q_ref["returncode"] == 0 and q_got["returncode"] == 0, q_got["stdout"][-2000:]
    _, refl = formats.read_cbook(cb)
    raas = O.refl_to_raas(refl)
    for k in range(2):
        rel = os.path.join("data", "sequences", "TRAIN", f"M{STOP}", "cfg1", f"part_{k:05d}.seq")
        a, b = (gpu_dir / rel).read_bytes(), (ref_dir / rel).read_bytes()
        assert len(a) == len(b) and a[:120] == b[:120]
        sa, sb = np.frombuffer(a, "<u2", offset=120), np.frombuffer(b, "<u2", offset=120)
        bad = np.nonzero(sa != sb)[0]
        if bad.size:
            _, dd, d2 = O.assign(x[k * (N // 2):(k + 1) * (N // 2)][bad], raas)
            assert np.all((d2 - dd) <= 1e-6 * dd)
    # the two mains print the same summary line (vq_quantize.c:105-108)
    tail = [l for l in q_ref["stdout"].splitlines() if l.startswith("total:")]
    tail2 = [l for l in q_got["stdout"].splitlines() if l.startswith("total:")]
    assert tail and tail == tail2

    # -B: resume from the reference's M=64 codebook through the reference's getopt
    base = str(ref_dir / cbdir / "eps_0.05_M_0064.cbook")
    ref_b, gpu_b = tmp_path / "ref_b", tmp_path / "gpu_b"
    ref_b.mkdir(), gpu_b.mkdir()
    rb = refbuild.run_vq_learn(files[ref_dir], str(ref_b), eps=0.05, max_M=STOP, base_codebook=base)
    gb = refbuild.run_dropin("vq.learn", ["-B", base, "-e", "0.05"] + files[gpu_dir], str(gpu_b), env)
    assert rb["returncode"] == 0 and gb["returncode"] == 0, gb["stdout"][-2000:]
    for M in (128, 256):
        name = f"eps_0.05_M_{M:04d}.cbook"
        _, a = formats.read_cbook(str(gpu_b / cbdir / name))
        _, b = formats.read_cbook(str(ref_b / cbdir / name))
        assert np.allclose(a, b, rtol=1e-7, atol=1e-9), name


def test_patched_crate_library_runs_on_the_gpu(built_lib, tmp_path):
    """oracle/_ref/x86-64-v3/linked.ecoz2main: a caller of the PUBLIC API (what the Rust crate calls)
    linked against the reference's static library built from the file list of the patched build.rs
    (integration/build.rs.patch) + libecoz2_vq_b200.so.  ecoz2.c's own forwards run, the CUDA
    library does the work, the callback fires once per stage on the calling thread."""
    import re
    import subprocess
    from ecoz2rs_b200 import formats
    from ecoz2rs_b200.synth import ar_mix
    from oracle import oracle as O
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "oracle", "_ref", "x86-64-v3", "linked.ecoz2main")
    if not os.path.exists(exe):
        pytest.skip("linked.ecoz2main not built")
    x = ar_mix(30_000, 12, seed=4)
    f = tmp_path / "predictors" / "TRAIN" / "c" / "a.prd"
    formats.write_prd(str(f), x, "c")
    env = dict(os.environ, ECOZ2_VQ_MAX_M="32")
    r = subprocess.run([exe, "learn", "12", "0.05", "c", str(f)], cwd=tmp_path, env=env, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    cbs = re.findall(r"\[callback (\d+)\] M=(\d+) avgDistortion=(\S+)", r.stdout)
    assert [int(m) for _, m, _ in cbs] == [2, 4, 8, 16, 32] and "B200 CUDA path" in r.stdout
    stages, _ = O.learn(x, eps=0.05, max_M=32)
    for (_, m, avg), st in zip(cbs, stages):
        assert abs(float(avg) - st["avg_distortion"]) <= 1e-7 * st["avg_distortion"]
    _, refl = formats.read_cbook(str(tmp_path / "data" / "codebooks" / "c" / "eps_0.05_M_0032.cbook"))
    assert np.allclose(refl, stages[-1]["refl"], rtol=1e-6, atol=1e-9)
    cb = str(tmp_path / "data" / "codebooks" / "c" / "eps_0.05_M_0032.cbook")
    r = subprocess.run([exe, "quantize", cb, str(f)], cwd=tmp_path, env=env, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    seq = np.frombuffer((tmp_path / "data" / "sequences" / "TRAIN" / "M32" / "c" / "a.seq").read_bytes(), "<u2", offset=120)
    oseq, _ = O.quantize(O.refl_to_raas(refl), x)
    assert (seq != oseq).mean() < 1e-4


def test_reference_seq_provider_over_the_library_quantize(built_lib, tmp_path):
    """The reference's SeqProvider (vq/vq.c:162-334, compiled unmodified) quantizes every predictor with
    every codebook through cbook_quantize -> quantize (vq/vq.c:154-157).  Linked against the library,
    `quantize` IS the B200 path (quantize.c leaves the build): same symbols (hash of every sequence) and
    distortions as the all-reference build, for codebooks of different content, files of 1 .. 30k frames."""
    import subprocess
    from ecoz2rs_b200 import formats
    from ecoz2rs_b200.synth import ar_mix
    from oracle import oracle as O
    from oracle import refbuild
    d = refbuild.arch_dir()
    exes = [os.path.join(d, "ref.seqprov"), os.path.join(d, "dropin.seqprov")]
    if not all(os.path.exists(e) for e in exes):
        pytest.skip("oracle/_ref seqprov executables not built")
    K, R = 128, 3
    cbs = []
    for r in range(R):
        st, _ = O.learn(ar_mix(6000, P, seed=300 + r), eps=0.05, max_M=K)
        f = tmp_path / f"m{r}.cbook"
        formats.write_cbook(str(f), st[-1]["refl"], f"class{r}")
        cbs.append(str(f))
    prds = []
    for k, T in enumerate((1, 7, 2500, 30_001, 1023)):
        f = tmp_path / "predictors" / "TEST" / f"class{k % R}" / f"f{k}.prd"
        formats.write_prd(str(f), ar_mix(T, P, seed=900 + k), f"class{k % R}")
        prds.append(str(f))
    outs = []
    for exe in exes:
        pr = subprocess.run([exe, str(R)] + cbs + prds, cwd=tmp_path, capture_output=True, text=True, timeout=300)
        if exe.endswith("ref.seqprov") and pr.returncode != 0:
            pytest.skip("reference seqprov failed on this host: " + pr.stderr[-300:])
        assert pr.returncode == 0, pr.stdout[-500:] + pr.stderr[-500:]
        rows = [ln.split() for ln in pr.stdout.splitlines() if ln and ln[0].isdigit()]
        outs.append(rows)
    ref, got = outs
    assert len(ref) == len(got) == R * len(prds)
    for a, b in zip(ref, got):
        assert a[:3] == b[:3]                                                # file, model, T
        assert abs(float(a[3]) - float(b[3])) <= 1e-9 * max(1.0, abs(float(a[3]))), (a, b)
        assert a[4:] == b[4:], (a, b)                                        # hash of the symbols, first symbols
