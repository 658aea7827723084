"""Timing of ecoz2b200_quantize_models (cbook_quantize of all resident frames against R codebooks,
vq/vq.c:154-157, 285-315; vq_classify.c:265-280): one sweep over the concatenated codebooks against R
sweeps (ECOZ2_VQ_MODELS_SERIAL=1), R = 3 ... 32.  usage: python scripts/models_probe.py [--n 10000000]"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from ecoz2rs_b200 import vq  # noqa: E402
from ecoz2rs_b200.synth import ar_mix_torch_range  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=10_000_000)
ap.add_argument("--files", type=int, default=1000)
ap.add_argument("--ks", default="64,256,1024")
ap.add_argument("--rs", default="3,8,16,32")
a = ap.parse_args()
P, D = 36, 37
dev = torch.device("cuda", 0)
x = ar_mix_torch_range(0, a.n, P, seed=20260102, device=dev)
cuts = np.linspace(0, a.n, a.files + 1).astype(np.int64)
rows = []
for K in [int(k) for k in a.ks.split(",")]:
    with vq.VqEngine(P, max_M=max(K, 2)) as eng:
        eng.set_frames_device(x.data_ptr(), a.n)
        eng.codebook_initial()
        st, _ = eng.learn(eps=0.05, stop_M=K)
        base = st[-1]["refl"]
        rng = np.random.default_rng(K)
        for R in [int(r) for r in a.rs.split(",")]:
            # R different models: the learnt codebook with its reflections perturbed
            refl = np.stack([np.clip(base * (1.0 + 0.02 * rng.standard_normal(base.shape)), -0.999, 0.999)
                             for _ in range(R)])
            refl[:, :, 0] = 0.0
            res = {}
            for label, env in (("one_sweep", None), ("serial", "1")):
                if env:
                    os.environ["ECOZ2_VQ_MODELS_SERIAL"] = env
                else:
                    os.environ.pop("ECOZ2_VQ_MODELS_SERIAL", None)
                eng.quantize_models(refl, cuts, want_seq=False)             # warm-up (buffers)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                _, dists, one = eng.quantize_models(refl, cuts, want_seq=False)
                ms = (time.perf_counter() - t0) * 1e3
                res[label] = (ms, dists, one)
            os.environ.pop("ECOZ2_VQ_MODELS_SERIAL", None)
            same = float(np.abs(res["one_sweep"][1] - res["serial"][1]).max())
            row = {"K": K, "R": R, "frames": a.n, "files": a.files, "one_sweep_ms": res["one_sweep"][0],
                   "one_sweep_route": res["one_sweep"][2], "serial_ms": res["serial"][0],
                   "speedup": res["serial"][0] / res["one_sweep"][0],
                   "evals_per_s_one_sweep": a.n * R * K / (res["one_sweep"][0] * 1e-3),
                   "max_abs_dist_diff": same}
            rows.append(row)
            print(json.dumps(row), flush=True)
