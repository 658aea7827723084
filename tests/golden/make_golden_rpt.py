#!/usr/bin/env python
"""Golden `.rpt` text of the reference (ecoz2/src/vq/report.c:14-75) for the frames of learn_P36.npz / learn_P12.npz:
the reference's own `vq.learn` (oracle/_ref) is run on them and its `eps_0.05.rpt` stored as
tests/golden/learn_<tag>.rpt.txt.  Build container only (needs oracle/_ref).

    python tests/golden/make_golden_rpt.py
"""
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from ecoz2rs_b200 import formats  # noqa: E402
from oracle import refbuild  # noqa: E402


def one(tag, P):
    g = np.load(os.path.join(HERE, f"learn_{tag}.npz"))
    tmp = tempfile.mkdtemp(prefix="golden_rpt_")
    try:
        prd = os.path.join(tmp, "predictors", "TRAIN", "gold", "x.prd")
        formats.write_prd(prd, g["x"], "gold")
        res = refbuild.run_vq_learn([prd], tmp, P=P, eps=0.05, class_name="gold", max_M=256, threads=4)
        assert res["returncode"] == 0, res["stdout"][-2000:]
        cdir = os.path.join(tmp, "data", "codebooks", "gold")
        rows = refbuild.read_rpt_csv(os.path.join(cdir, "eps_0.05.rpt.csv"))
        # the same run as the one learn_P36.npz was made from
        assert np.allclose(np.array(rows, dtype=np.float64), g["rpt_csv"], rtol=0, atol=1.01e-6)
        shutil.copy(os.path.join(cdir, "eps_0.05.rpt"), os.path.join(HERE, f"learn_{tag}.rpt.txt"))
        print(open(os.path.join(HERE, f"learn_{tag}.rpt.txt"), encoding="utf-8").read()[:600])
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    one("P36", 36)
    one("P12", 12)
