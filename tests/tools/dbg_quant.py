import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from ecoz2rs_b200 import vq
from ecoz2rs_b200.synth import ar_mix
from oracle import oracle as O
P = 12
x = ar_mix((1 << 20) + 12345, P, seed=5)
stages, _ = O.learn(x[:20000], eps=0.05, max_M=64)
refl = stages[-1]["refl"]
raas = O.refl_to_raas(refl)
seq_ref, avg_ref = O.quantize(raas, x)
for mode in (0, 1):
    for rep in range(2):
        with vq.VqEngine(P, max_M=64, mode=mode) as eng:
            eng.set_reflections(refl)
            seq, avg = eng.quantize(x)
            bad = np.nonzero(seq != seq_ref)[0]
            print("mode", mode, "rep", rep, "host mism", bad.size, "first", bad[:8], "last", bad[-3:], "avg", avg, avg_ref, flush=True)
            eng.set_frames(x)
            seq2, avg2 = eng.quantize_resident()
            bad = np.nonzero(seq2 != seq_ref)[0]
            print("   resident mism", bad.size, bad[:8], eng.counters()["rechecked"], flush=True)
            idx, dd, nre = eng.assign()
            bad = np.nonzero(idx != seq_ref)[0]
            print("   assign mism", bad.size, bad[:8], nre, flush=True)
