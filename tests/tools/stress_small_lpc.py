"""Repeatability stress of the two kernels added last in round 1 (compute-sanitizer is not available on
the GPU pool): the fused small-codebook pass must give identical counts / indices and cell sums equal
to rounding on every repetition, the LPC front end bit-identical vectors; sizes chosen to exercise
partial tiles and the last CTA.  usage: python tests/tools/stress_small_lpc.py [--reps 40]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from ecoz2rs_b200 import lpc, vq  # noqa: E402
from ecoz2rs_b200.synth import ar_mix_torch  # noqa: E402
from oracle import oracle as O  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--reps", type=int, default=40)
a = ap.parse_args()
bad = 0
for P, N, K in ((36, 1_000_003, 2), (36, 777_777, 4), (36, 500_001, 8), (12, 2_000_001, 4), (20, 300_007, 3)):
    x = ar_mix_torch(N, P, seed=5 + P, device="cuda")
    rng = np.random.default_rng(K)
    refl = np.zeros((K, P + 1))
    refl[:, 1:] = np.clip(rng.normal(size=(K, P)) * 0.3, -0.9, 0.9)
    with vq.VqEngine(P, max_M=64) as eng:
        eng.set_frames_device(x.data_ptr(), N)
        first = None
        for r in range(a.reps):
            eng.set_reflections(refl)
            out = eng.kmeans_pass()
            idx, _ = eng.last_assignment()
            if first is None:
                first = (out, idx.copy())
                xs = x[:20000].cpu().numpy()
                ref = O.kmeans_pass(xs, refl, O.refl_to_raas(refl))
                assert np.array_equal(idx[:20000], ref["idx"]), "indices differ from the oracle"
                assert int(out["cardd"].sum()) == N
            else:
                ok = (np.array_equal(out["cardd"], first[0]["cardd"]) and np.array_equal(idx, first[1]) and
                      np.allclose(out["cells"], first[0]["cells"], rtol=1e-11, atol=1e-9) and
                      abs(out["DD"] - first[0]["DD"]) <= 1e-10 * abs(first[0]["DD"]))
                bad += not ok
    print(f"small pass P={P} N={N} K={K}: {a.reps} repetitions, mismatches so far {bad}", flush=True)
    del x
for P, W, Off, secs in ((36, 45, 15, 1801.37), (12, 45, 15, 600.01), (20, 20, 5, 333.3)):
    n = int(secs * 32000)
    g = torch.Generator(device="cuda").manual_seed(3)
    s = torch.randn(n, device="cuda", dtype=torch.float64, generator=g) * 0.1
    T = lpc.num_frames(n, 32000., W, Off)
    v0 = torch.empty((T, P + 1), dtype=torch.float64, device="cuda")
    v1 = torch.full((T, P + 1), float("nan"), dtype=torch.float64, device="cuda")
    lpc.lpa_on_signal_device(s.data_ptr(), n, 32000., P, W, Off, v0.data_ptr(), T)
    for r in range(a.reps):
        v1.fill_(float("nan"))
        lpc.lpa_on_signal_device(s.data_ptr(), n, 32000., P, W, Off, v1.data_ptr(), T)
        bad += not torch.equal(v0, v1)
    print(f"lpc P={P} {W}/{Off} ms, {T} frames: {a.reps} repetitions, mismatches so far {bad}", flush=True)
print("STRESS", "OK" if bad == 0 else f"FAILED ({bad})")
sys.exit(1 if bad else 0)
