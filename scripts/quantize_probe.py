"""Dev probe: vq quantize of resident frames (search + FP64 distortion of the winner + u16 symbols)."""
import argparse, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ecoz2rs_b200 import vq
from ecoz2rs_b200.synth import ar_mix_torch
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=10_000_000)
ap.add_argument("--p", type=int, default=36)
ap.add_argument("--k", type=int, default=2048)
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
xd = ar_mix_torch(a.n, a.p, seed=20260102, device="cuda")
torch.cuda.synchronize()
eng = vq.VqEngine(a.p, max_M=max(a.k, 256))
eng.set_frames_device(xd.data_ptr(), a.n)
del xd
eng.codebook_initial()
eng.learn(eps=0.05, stop_M=a.k)
seq = torch.empty(a.n, dtype=torch.int16, pin_memory=True)
eng.quantize_resident_ptr(seq.data_ptr())
for want in (True, False):
    eng.counters_reset()
    t0 = time.perf_counter()
    for _ in range(a.reps):
        if want:
            eng.quantize_resident_ptr(seq.data_ptr())
        else:
            eng.quantize_resident(want_seq=False)
    wall = (time.perf_counter() - t0) / a.reps
    c = eng.counters()
    print(f"K={a.k} n={a.n} seq_out={want}: wall {wall*1e3:.2f} ms  assign {c['assign_ms']/a.reps:.2f} accumulate {c['accumulate_ms']/a.reps:.2f} "
          f"search_main {c['search_main_ms']/max(1,c['search_main_launches']):.2f} rechecked {c['rechecked']/a.reps/a.n*100:.2f}%  -> {a.n*a.k/wall/1e9:.0f} G evals/s")
