"""Repeat the in-memory LBG run (2 -> 256 on 100k config-1 frames, then 2 -> 64 at P=12) many times in
one process and compare every repetition with the first: passes per stage, average distortion, sigma,
inertia (1e-9) and the final reflections (1e-7).  A rare race in any kernel of the pass shows up here.
usage: python tests/tools/stress_learn_repeat.py [--reps 30]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from ecoz2rs_b200 import vq  # noqa: E402
from ecoz2rs_b200.synth import ar_mix  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--reps", type=int, default=30)
a = ap.parse_args()
bad = 0
for P, N, M, seed in ((36, 100_000, 256, 20260101), (12, 150_001, 64, 20260105)):
    x = ar_mix(N, P, seed=seed)
    first = None
    with vq.VqEngine(P, max_M=M) as eng:
        eng.set_frames(x)
        for r in range(a.reps):
            eng.codebook_initial()
            stages, _ = eng.learn(eps=0.05, stop_M=M)
            key = [(s["M"], s["passes"]) for s in stages]
            vals = np.array([[s["avg_distortion"], s["sigma"], s["inertia"]] for s in stages])
            if first is None:
                first = (key, vals, stages[-1]["refl"].copy())
                continue
            ok = key == first[0] and np.allclose(vals, first[1], rtol=1e-9, atol=0) and \
                np.allclose(stages[-1]["refl"], first[2], rtol=1e-7, atol=1e-10)
            if not ok:
                bad += 1
                print("MISMATCH rep", r, key, first[0], np.abs(vals / first[1] - 1).max(), flush=True)
    print(f"P={P} N={N} 2->{M}: {a.reps} repetitions, mismatches so far {bad}", flush=True)
print("STRESS", "OK" if bad == 0 else f"FAILED ({bad})")
sys.exit(1 if bad else 0)
