"""Regression for a stream-ordering bug of the engine constructor: it cleared the codebook buffers with
cudaMemset, i.e. on the legacy default stream, which is asynchronous to the host and NOT ordered against
the engine's non-blocking stream.  A codebook uploaded right after ecoz2b200_create could be zeroed again
when the fill ran late (seen as a rare wrong pass in the first process of a cold box: the cells of the
zeroed codewords empty, their frames in cell 0).

Here the legacy default stream (torch's default stream) is kept busy while the engine is created and the
codebook uploaded, so a fill queued on the legacy stream WOULD land after the upload; the codebook read
back afterwards has to be the one uploaded, every repetition, and a pass over it has to fill the cells.

usage: python tests/tools/create_upload_order.py [reps]     (ECOZ2_VQ_B200_LIB selects another build)"""
import sys

import numpy as np
import torch

sys.path.insert(0, __file__.rsplit("/tests/", 1)[0])
from ecoz2rs_b200 import vq                                  # noqa: E402
from ecoz2rs_b200.synth import ar_mix                        # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
P, K = 12, 300
rng = np.random.default_rng(3)
refl = np.zeros((K, P + 1))
refl[:, 1:] = np.clip(rng.normal(size=(K, P)) * 0.3, -0.9, 0.9)
x = ar_mix(200_000, P, seed=9)

with vq.VqEngine(P, max_M=K) as eng:          # reference counts: nothing else in flight
    eng.set_reflections(refl)
    eng.set_frames(x)
    want = eng.kmeans_pass()["cardd"].copy()

busy = torch.ones(1 << 27, device="cuda")     # 512 MB: one sweep is ~0.15 ms of legacy-stream work
bad = 0
for rep in range(reps):
    for _ in range(60):                       # ~10 ms queued on the legacy default stream
        busy.mul_(1.0000001)
    with vq.VqEngine(P, max_M=K) as eng:
        eng.set_reflections(refl)
        torch.cuda.synchronize()              # whatever the constructor queued has run now
        got, _ = eng.codebook()
        eng.set_frames(x)
        cardd = eng.kmeans_pass()["cardd"]
    if not (np.array_equal(got, refl) and np.array_equal(cardd, want)):
        bad += 1
        print(f"rep {rep}: codebook equal {np.array_equal(got, refl)}, zero rows "
              f"{int((np.abs(got).sum(axis=1) == 0).sum())}, counts equal {np.array_equal(cardd, want)}",
              flush=True)
print(f"create/upload order: {reps} repetitions, {bad} wrong")
if bad:
    sys.exit(1)
print("STRESS OK")
