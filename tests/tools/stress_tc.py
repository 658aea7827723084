"""Stress of the tensor-core search: repeated quantize / assign against the FP64 mode."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from ecoz2rs_b200 import vq
from ecoz2rs_b200.synth import ar_mix
from oracle import oracle as O

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
total_bad = 0
for P, K, N in [(12, 64, (1 << 20) + 12345), (36, 64, 300_000), (36, 512, 200_000), (12, 256, 500_000)]:
    x = ar_mix(N, P, seed=5)
    stages, _ = O.learn(x[:20000], eps=0.05, max_M=K)
    refl = stages[-1]["refl"]
    with vq.VqEngine(P, max_M=K, mode=vq.MODE_FP64) as eng:
        eng.set_reflections(refl)
        seq64, _ = eng.quantize(x)
    bad_runs = 0
    with vq.VqEngine(P, max_M=K) as eng:
        eng.set_reflections(refl)
        for rep in range(reps):
            seq, _ = eng.quantize(x)
            nb = int((seq != seq64).sum())
            if nb:
                bad_runs += 1
                b = np.nonzero(seq != seq64)[0]
                print(f"P={P} K={K} rep {rep}: {nb} mismatches, first {b[:5]}, last {b[-3:]}, "
                      f"chunks {np.bincount(b >> 20)}", flush=True)
    print(f"P={P} K={K} N={N}: {bad_runs} bad runs of {reps}", flush=True)
    total_bad += bad_runs
print("STRESS", "OK" if total_bad == 0 else f"FAILED ({total_bad})")
sys.exit(1 if total_bad else 0)
