import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from ecoz2rs_b200 import vq
from ecoz2rs_b200.synth import ar_mix
from oracle import oracle as O
P = 12
x = ar_mix(300000, P, seed=5)
stages, _ = O.learn(x[:20000], eps=0.05, max_M=64)
refl = stages[-1]["refl"]
raas = O.refl_to_raas(refl)
seq_ref, avg_ref = O.quantize(raas, x)
x36 = ar_mix(50000, 36, seed=3)
st36, _ = O.learn(x36[:10000], eps=0.05, max_M=64)
ref36, _ = O.quantize(O.refl_to_raas(st36[-1]["refl"]), x36)
nbad = 0
for it in range(int(sys.argv[1]) if len(sys.argv) > 1 else 30):
    # leave FP64 tiles / other data in shared memory
    with vq.VqEngine(36, max_M=64, mode=1) as e2:
        e2.set_reflections(st36[-1]["refl"])
        s36, _ = e2.quantize(x36)
        if (s36 != ref36).sum() > 2: print("it", it, "fp64 P36 mism", (s36 != ref36).sum(), flush=True)
    with vq.VqEngine(P, max_M=64) as eng:
        eng.set_reflections(refl)
        seq, avg = eng.quantize(x)
    bad = np.nonzero(seq != seq_ref)[0]
    if bad.size > 2:
        nbad += 1
        tiles = np.unique(bad // 128)
        print("it", it, "mism", bad.size, "tiles", tiles.size, "of", (len(x) + 127) // 128, "first tiles", tiles[:10],
              "gpu vals mod4 hist", np.bincount(seq[bad] % 4, minlength=4), flush=True)
print("bad iterations:", nbad)
