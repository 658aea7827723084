"""Dev probe: time one k-means pass / quantize at several K on resident frames (1 GPU).
Prints achieved frame x codeword evaluations per second and the FP32-roofline fraction."""
import argparse
import sys
import time
import os

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ecoz2rs_b200 import vq
from ecoz2rs_b200.synth import ar_mix, ar_mix_torch

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=2_000_000)
ap.add_argument("--p", type=int, default=36)
ap.add_argument("--ks", type=str, default="2,16,64,256,1024,2048")
ap.add_argument("--mode", type=int, default=2)
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()

import torch
t0 = time.time()
xd = ar_mix_torch(a.n, a.p, seed=20260102, device="cuda")
torch.cuda.synchronize()
print(f"generated {a.n} frames on device in {time.time()-t0:.2f}s", flush=True)

rng = np.random.default_rng(1)
eng = vq.VqEngine(a.p, max_M=4096, mode=a.mode)
eng.set_frames_device(xd.data_ptr(), a.n)
del xd
torch.cuda.empty_cache()
eng.codebook_initial()
peak = 148 * 128 * 2 * 1.965e9
for K in [int(k) for k in a.ks.split(",")]:
    while eng.K < K:
        eng.kmeans_pass(want_cells=False)
        eng.grow()
    eng.kmeans_pass(want_cells=False)  # warm
    eng.counters_reset()
    t0 = time.time()
    for _ in range(a.reps):
        eng.kmeans_pass(want_cells=False)
    wall = (time.time() - t0) / a.reps
    c = eng.counters()
    am, cm, um = c["assign_ms"] / a.reps, c["accumulate_ms"] / a.reps, c["update_ms"] / a.reps
    evals = a.n * K
    sm = c["search_main_ms"] / max(1, c["search_main_launches"])
    print(f"K={K:5d} pass wall {wall*1e3:8.2f} ms | main search {sm:8.2f} ms ({evals/sm/1e6:7.1f} G evals/s) | assign {am:8.2f} ms accumulate {cm:7.2f} ms update {um:6.2f} ms "
          f"| {evals/am/1e6:9.1f} G evals/s (assign)  {evals*2*(a.p+1)/(am*1e-3)/peak*100:5.1f}% of FP32 peak "
          f"| rechecked/pass {c['rechecked']/a.reps:.0f} ({c['rechecked']/a.reps/a.n*100:.3f}%)", flush=True)
