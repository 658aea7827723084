#!/usr/bin/env python
"""Generate the golden fixtures of the VQ path from the REFERENCE itself.

Runs in the build container only (needs /root/reference, via oracle/_ref built by
oracle/Makefile): the reference executables `vq.learn` / `vq.quantize` and the
reference library functions are run on small seeded inputs and their outputs are
stored as compressed .npz files next to this script.  The GPU box has no
/root/reference: the tests there compare against these files.

    python tests/golden/make_golden.py

Fixtures (inputs are stored too, so no RNG stream needs to be reproduced):
  learn_P12.npz   N=3000, P=12, eps=0.05, LBG 2 -> 256 : per-stage reflections
                  (.cbook payloads), .rpt.csv rows, cards_dists, last min_dists
  learn_P36.npz   N=2000, P=36, eps=0.05, LBG 2 -> 256 (same content)
  quantize_P36.npz  .seq symbols written by vq.quantize for the P=36 frames against
                  the M=64 codebook above + the average distortion it prints
  kernels.npz     distortion / lpca_r / reflections_to_raas / sigma / inertia /
                  calculate_reflections / grow_codebook outputs of the reference
                  library on seeded inputs
"""
from __future__ import annotations

import os
import re
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from ecoz2rs_b200 import formats  # noqa: E402
from ecoz2rs_b200.synth import ar_mix  # noqa: E402
from oracle import refbuild  # noqa: E402


def run_learn(x, P, max_M, tag):
    tmp = tempfile.mkdtemp(prefix="golden_")
    try:
        prd = os.path.join(tmp, "predictors", "TRAIN", "gold", "x.prd")
        formats.write_prd(prd, x, "gold")
        res = refbuild.run_vq_learn([prd], tmp, P=P, eps=0.05, class_name="gold", max_M=max_M, threads=4)
        assert res["returncode"] == 0, res["stdout"][-2000:]
        cdir = os.path.join(tmp, "data", "codebooks", "gold")
        rows = refbuild.read_rpt_csv(os.path.join(cdir, "eps_0.05.rpt.csv"))
        out = {"x": x, "rpt_csv": np.array(rows, dtype=np.float64)}
        with open(os.path.join(cdir, "eps_0.05.rpt.csv"), "rb") as f:
            out["rpt_csv_text"] = np.frombuffer(f.read(), dtype=np.uint8)
        for (M, *_rest) in rows:
            M = int(M)
            base = os.path.join(cdir, f"eps_0.05_M_{M:04d}.cbook")
            cls, refl = formats.read_cbook(base)
            assert cls == "gold"
            out[f"refl_{M}"] = refl
            cards = np.genfromtxt(base + ".cards_dists.csv", delimiter=",", skip_header=1)
            out[f"cards_{M}"] = cards.reshape(-1, 3)
        md = np.genfromtxt(base + ".min_dists.csv", delimiter=",", skip_header=1)
        out["last_min_dists"] = md
        with open(os.path.join(cdir, f"eps_0.05_M_{int(rows[1][0]):04d}.cbook"), "rb") as f:
            out["cbook_bytes_stage2"] = np.frombuffer(f.read(), dtype=np.uint8)
        np.savez_compressed(os.path.join(HERE, f"learn_{tag}.npz"), **out)
        print(tag, "stages:", [(int(r[0]), int(r[1])) for r in rows])
        return out
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def run_quantize(x, refl, tag):
    tmp = tempfile.mkdtemp(prefix="golden_")
    try:
        prd = os.path.join(tmp, "predictors", "TRAIN", "gold", "x.prd")
        formats.write_prd(prd, x, "gold")
        cb = os.path.join(tmp, "cb.cbook")
        formats.write_cbook(cb, refl, "gold")
        res = refbuild.run_vq_quantize(cb, [prd], tmp)
        assert res["returncode"] == 0, res["stdout"]
        M = refl.shape[0]
        seq_path = os.path.join(tmp, "data", "sequences", "TRAIN", f"M{M}", "gold", "x.seq")
        cls, M2, seq = formats.read_seq(seq_path)
        assert cls == "gold" and M2 == M
        with open(seq_path, "rb") as f:
            seq_bytes = np.frombuffer(f.read(), dtype=np.uint8)
        m = re.search(r"average distortion = ([0-9.eE+-]+)", res["stdout"])
        np.savez_compressed(os.path.join(HERE, f"quantize_{tag}.npz"), x=x, refl=refl, seq=seq,
                            seq_file_bytes=seq_bytes, avg_printed=float(m.group(1)))
        print(tag, "quantize avg", m.group(1))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def run_kernels():
    rng = np.random.default_rng(42)
    P = 36
    x = ar_mix(400, P, seed=99)
    refl = np.zeros((8, P + 1))
    refl[:, 1:] = np.clip(rng.normal(size=(8, P)) * 0.25, -0.9, 0.9)
    raas = refbuild.refl_to_raas(refl)
    dist = np.array([[refbuild.distortion(x[v], raas[i]) for i in range(8)] for v in range(50)])
    lp = [refbuild.lpca_r(x[v]) for v in range(20)]
    idx = dist.argmin(1)
    seq, avg = refbuild.quantize(raas, x)
    # one teacher-forced centroid step through the reference's calculate_reflections
    cells = np.zeros((8, P + 1))
    cardd = np.zeros(8, np.int32)
    for v in range(len(x)):
        cells[seq[v]] += x[v]
        cardd[seq[v]] += 1
    ncells, nrefl, nraas = refbuild.centroids(cells, cardd, refl, raas)
    g_refl, g_raas = refbuild.grow(refl)
    i_refl, i_raas = refbuild.initial(P)
    avgd = 0.3
    sig = refbuild.sigma(nraas, ncells, avgd)
    inert = refbuild.inertia(x, nraas)
    np.savez_compressed(
        os.path.join(HERE, "kernels.npz"), x=x, refl=refl, raas=raas, dist=dist,
        lpca_code=np.array([c for c, *_ in lp]), lpca_rc=np.array([rc for _, rc, _, _ in lp]),
        lpca_a=np.array([a for _, _, a, _ in lp]), lpca_pe=np.array([pe for *_, pe in lp]),
        seq=seq, avg=avg, cells=cells, cardd=cardd, cells_norm=ncells, new_refl=nrefl, new_raas=nraas,
        grow_refl=g_refl, grow_raas=g_raas, init_refl=i_refl, init_raas=i_raas, sigma=sig,
        sigma_avg=avgd, inertia=inert)
    print("kernels ok; sigma", sig, "inertia", inert)


def main():
    assert refbuild.build(), "oracle/_ref could not be built (needs /root/reference)"
    x12 = ar_mix(3000, 12, seed=20260105)
    x36 = ar_mix(2000, 36, seed=20260101)
    run_learn(x12, 12, 256, "P12")
    g36 = run_learn(x36, 36, 256, "P36")
    run_quantize(x36, g36["refl_64"], "P36")
    run_kernels()


if __name__ == "__main__":
    main()
