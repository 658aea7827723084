"""Dev probe of the tcgen05 split-TF32 search: value accuracy against FP64, index parity
against the FP64 mode, and the re-evaluation rate as a function of kappa_tc."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ecoz2rs_b200 import vq
from ecoz2rs_b200.synth import ar_mix

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=200_000)
ap.add_argument("--p", type=int, default=36)
ap.add_argument("--ks", type=str, default="64,256,2048")
ap.add_argument("--rows", type=int, default=4096)
a = ap.parse_args()

x = ar_mix(a.n, a.p, seed=20260102)
eng = vq.VqEngine(a.p, max_M=4096)
eng.set_frames(x)
eng.codebook_initial()
for K in [int(k) for k in a.ks.split(",")]:
    eng.set_mode(vq.MODE_FP32)
    while eng.K < K:
        eng.kmeans_pass(want_cells=False)
        eng.grow()
    eng.kmeans_pass(want_cells=False)
    refl, raas = eng.codebook()
    w = raas.copy()
    w[:, 1:] *= 2
    eng.set_mode(vq.MODE_FP64)
    idx64, dd64, _ = eng.assign()
    eng.set_mode(vq.MODE_TF32X3)
    t0 = time.time()
    d = eng.debug_tc_distortions(a.rows).astype(np.float64)
    print(f"K={K}: dump of {a.rows} rows in {time.time()-t0:.2f}s", flush=True)
    ref = x[:a.rows] @ w.T
    err = np.abs(d - ref)
    scale = np.abs(x[:a.rows, :1]) * np.abs(w).sum(1)[None, :]
    u = 2.0 ** -24
    print(f"   max |err| {err.max():.3e}  max err/(x0 W1) {np.max(err/scale)/u:.3f} u24  "
          f"rms {np.sqrt(np.mean((err/scale)**2))/u:.4f} u24  max rel err {np.max(err/np.abs(ref)):.3e}",
          flush=True)
    for kap in (16, 32, 48, 64):
        eng.set_recheck_kappa_tc(kap)
        eng.counters_reset()
        idx, dd, nre = eng.assign()
        bad = int((idx != idx64).sum())
        if bad and kap == 32:
            rows = np.nonzero(idx != idx64)[0][:12]
            dd = x[rows] @ w.T
            for rr, drow in zip(rows, dd):
                srt = np.sort(drow)
                print(f"      row {rr} (tile {rr // 256}, pos {rr % 256}): got {idx[rr]} want {idx64[rr]} "
                      f"d[got]-dmin {drow[idx[rr]] - srt[0]:.3e}  gap12 {srt[1] - srt[0]:.3e}  x0 {x[rr, 0]:.3f}")
        print(f"   kappa_tc={kap:3d}: rechecked {nre} ({nre/a.n*100:.3f}%)  index mismatches vs FP64 mode: {bad}",
              flush=True)
