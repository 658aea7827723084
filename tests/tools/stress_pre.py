"""Repetition stress of the search with pre-staged frame operands (vq_assign_tc2, TcParams::xop) and of the
chunked host pass: every repetition must give the counts of the first.  usage: stress_pre.py [reps]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from ecoz2rs_b200 import vq  # noqa: E402
from ecoz2rs_b200.synth import ar_mix  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
bad = 0
for P, N, K in ((12, 2_500_000, 300), (36, 1_000_003, 64), (36, 700_001, 2048)):
    x = ar_mix(N, P, seed=77)
    rng = np.random.default_rng(K)
    refl = np.zeros((K, P + 1))
    refl[:, 1:] = np.clip(rng.normal(size=(K, P)) * 0.3 / np.sqrt(P / 12), -0.9, 0.9)
    first = None
    with vq.VqEngine(P, max_M=K) as eng:
        for r in range(reps):
            eng.set_reflections(refl)
            if r % 3 == 2:
                out = eng.kmeans_pass_host(x)
            else:
                if r % 3 == 0:
                    eng.set_frames(x)
                out = eng.kmeans_pass()
            key = out["cardd"].copy()
            if first is None:
                first = key
            elif not np.array_equal(key, first):
                bad += 1
                print(f"MISMATCH P={P} K={K} rep {r} (mode {r % 3}): {int((key != first).sum())} cells differ, "
                      f"cell 0 {key[0]} vs {first[0]}", flush=True)
    print(f"P={P} K={K} N={N}: {reps} repetitions, mismatches so far {bad}", flush=True)
print("STRESS", "OK" if bad == 0 else f"FAILED ({bad})")
sys.exit(1 if bad else 0)
