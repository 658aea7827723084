#!/usr/bin/env python
"""Golden fixture of `vq classify` from the REFERENCE itself.

Build container only (needs /root/reference through oracle/_ref).  Three classes of synthetic
P=12 predictors; one M=32 codebook per class trained by the reference's vq.learn; the
reference's vq.classify executable (-r) run on 7 test files per class plus one file of a class
without a model.  Its stdout (model list, rankings of the misclassified files, confusion matrix,
candidate-order table) and its stderr marks are stored with all inputs in classify_P12.npz.

    python tests/golden/make_golden_classify.py

NB the reference's main counts its model arguments by the suffix ".hmm" (x/vq.classify.c:62-68),
so the codebook files are named *.hmm here; vq_classify itself does not look at the name.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from ecoz2rs_b200 import formats  # noqa: E402
from ecoz2rs_b200.synth import ar_mix  # noqa: E402
from oracle import refbuild  # noqa: E402

P = 12
CLASSES = ["alpha", "bravo", "charlie_long_name"]
M = 32


def main():
    assert refbuild.build(), "oracle/_ref could not be built (needs /root/reference)"
    rng = np.random.default_rng(20260107)
    tmp = tempfile.mkdtemp(prefix="golden_cls_")
    out = {"classes": np.array(CLASSES), "P": P, "M": M}
    try:
        cb_rel, prd_rel = [], []
        for ci, cls in enumerate(CLASSES):
            # classes overlap (same generator family, different seeds) so that some files are
            # misclassified and the ranking / candidate-order columns are exercised
            train = ar_mix(900, P, C=6, seed=500 + ci)
            prd = os.path.join(tmp, "train", cls, "t.prd")
            formats.write_prd(prd, train, cls)
            res = refbuild.run_vq_learn([prd], tmp, P=P, eps=0.05, class_name=cls, max_M=256, threads=2)
            assert res["returncode"] == 0, res["stdout"][-2000:]
            cb = os.path.join(tmp, "data", "codebooks", cls, f"eps_0.05_M_{M:04d}.cbook")
            name, refl = formats.read_cbook(cb)
            assert name == cls and refl.shape == (M, P + 1)
            out[f"refl_{ci}"] = refl
            rel = os.path.join("cb", f"{cls}.hmm")
            os.makedirs(os.path.join(tmp, "cb"), exist_ok=True)
            shutil.copy(cb, os.path.join(tmp, rel))
            cb_rel.append(rel)
        k = 0
        for ci, cls in enumerate(CLASSES + ["zulu"]):
            nfiles = 7 if cls != "zulu" else 1
            for f in range(nfiles):
                T = int(rng.integers(25, 90))
                # a test file mixes its own class with a neighbour's frames: harder instances
                own = ar_mix(T, P, C=6, seed=500 + (ci % 3))[rng.permutation(T)]
                other = ar_mix(T, P, C=6, seed=500 + ((ci + 1) % 3))
                take_other = rng.random(T) < (0.15 + 0.1 * f)
                x = np.where(take_other[:, None], other, own)
                rel = os.path.join("prd", cls, f"f{f}.prd")
                formats.write_prd(os.path.join(tmp, rel), x, cls)
                out[f"x_{k}"] = x
                prd_rel.append(rel)
                k += 1
        exe = os.path.join(refbuild.arch_dir(), "vq.classify")
        p = subprocess.run([exe, "-r"] + cb_rel + prd_rel, cwd=tmp, stdout=subprocess.PIPE,
                           stderr=subprocess.PIPE, text=True)
        assert p.returncode == 0, p.stderr
        out["cb_files"] = np.array(cb_rel)
        out["prd_files"] = np.array(prd_rel)
        out["stdout"] = np.frombuffer(p.stdout.encode(), dtype=np.uint8)
        out["stderr"] = np.frombuffer(p.stderr.encode(), dtype=np.uint8)
        np.savez_compressed(os.path.join(HERE, "classify_P12.npz"), **out)
        print(p.stdout)
        print("stderr:", p.stderr.replace("\x1b[32m", "").replace("\x1b[31m", "").replace("\x1b[0m", ""))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
