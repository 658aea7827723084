"""Dev check of the fused passes (vq_pass_mid, fused cell sums of vq_assign_tc2) against the oracle on small shapes."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from ecoz2rs_b200 import vq
from ecoz2rs_b200.synth import ar_mix
from oracle import oracle as O

rng = np.random.default_rng(1)
for P, K, N in [(36, 16, 1000), (36, 16, 100_000), (36, 64, 70_001), (36, 128, 50_000), (12, 32, 60_000), (36, 512, 40_000), (36, 1024, 100_000)]:
    x = ar_mix(N, P, seed=3)
    refl = np.zeros((K, P + 1))
    refl[:, 1:] = np.clip(rng.normal(size=(K, P)) * 0.3, -0.9, 0.9)
    raas = O.refl_to_raas(refl)
    ref = O.kmeans_pass(x, refl, raas)
    with vq.VqEngine(P, max_M=max(K, 2)) as eng:
        eng.set_frames(x)
        eng.set_reflections(refl)
        out = eng.kmeans_pass()
        idx, _ = eng.last_assignment()
    ok_idx = int((idx != ref["idx"]).sum())
    ok_c = np.array_equal(out["cardd"], ref["cardd"])
    err = np.abs(out["cells"] - ref["cells"]).max() / np.abs(ref["cells"]).max()
    print(f"P={P} K={K} N={N}: idx mismatches {ok_idx}, counts equal {ok_c}, cells rel err {err:.2e}, "
          f"DD {out['DD']:.9g} vs {ref['DD']:.9g}", flush=True)
