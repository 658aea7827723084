"""One fresh process: host-chunked routes against the resident routes (first use of every kernel
involved).  Exit 1 on a mismatch.  usage: fresh_process_check.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from ecoz2rs_b200 import vq  # noqa: E402
from ecoz2rs_b200.synth import ar_mix  # noqa: E402

bad = 0
# (1) kmeans_pass_host against set_frames + kmeans_pass, K = 300, P = 12, three chunks
x = ar_mix(2_500_000, 12, seed=77)
rng = np.random.default_rng(300)
refl = np.zeros((300, 13))
refl[:, 1:] = np.clip(rng.normal(size=(300, 12)) * 0.3, -0.9, 0.9)
outs = []
for host in (False, True):
    with vq.VqEngine(12, max_M=300) as eng:
        eng.set_reflections(refl)
        if host:
            a = eng.kmeans_pass_host(x)
        else:
            eng.set_frames(x)
            a = eng.kmeans_pass()
        b = eng.kmeans_pass()
        outs.append((a["cardd"].copy(), b["cardd"].copy()))
for k, name in enumerate(("first pass", "second pass")):
    if not np.array_equal(outs[0][k], outs[1][k]):
        bad += 1
        d = np.nonzero(outs[0][k] != outs[1][k])[0]
        print(f"MISMATCH kmeans_pass_host {name}: {d.size} cells, cell0 {outs[0][k][0]} vs {outs[1][k][0]}", flush=True)
# (2) quantize_host against quantize_resident, K = 8 and K = 128, P = 36
y = ar_mix(1_300_001, 36, seed=5)
for K in (8, 128):
    refl = np.zeros((K, 37))
    refl[:, 1:] = np.clip(np.random.default_rng(K).normal(size=(K, 36)) * 0.15, -0.9, 0.9)
    with vq.VqEngine(36, max_M=K) as eng:
        eng.set_reflections(refl)
        s1, a1 = eng.quantize(y)
        eng.set_frames(y)
        s2, a2 = eng.quantize_resident()
    if not np.array_equal(s1, s2):
        bad += 1
        d = np.nonzero(s1 != s2)[0]
        print(f"MISMATCH quantize K={K}: {d.size} symbols, first at {d[:5]}, last {d[-3:]}", flush=True)
print("FRESH", "OK" if bad == 0 else f"FAILED ({bad})")
sys.exit(1 if bad else 0)
