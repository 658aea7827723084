"""Full-size runs of BASELINE.json configs 4 and 5 on ONE B200 (frames synthesised on the device
with the survey's ar_mix formulae), one JSON line each:

  cfg5  vq learn P=12, 2 -> 1024 on 20M frames: wall time per LBG stage and total, evals/s
  cfg4  vq quantize of 50M frames against a 4096-codeword P=36 codebook (learned here on the
        first 10M of them = config 3's codebook): resident frames -> u16 symbols on the host,
        and host frames -> symbols (chunked H2D pipeline), plus size-independent checks
        (symbols of the first 10M frames == the learn pass's indices against the same codebook;
        average distortion == mean of the FP64 per-frame minima)

usage: python scripts/configs_probe.py [--cfg 4|5|45] [--scale 1.0]"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from ecoz2rs_b200 import vq  # noqa: E402
from ecoz2rs_b200.synth import ar_mix_torch  # noqa: E402


def cfg5(scale):
    P, N, K = 12, int(20_000_000 * scale), 1024
    x = ar_mix_torch(N, P, seed=20260105, device="cuda")
    torch.cuda.synchronize()
    with vq.VqEngine(P, max_M=K) as eng:
        eng.set_frames_device(x.data_ptr(), N)
        del x
        for rep in range(2):   # first run warms every kernel variant
            eng.codebook_initial()
            eng.counters_reset()
            t0 = time.perf_counter()
            stages, trace = eng.learn(eps=0.05, stop_M=K)
            wall = time.perf_counter() - t0
        c = eng.counters()
        last = stages[-1]
        line = {"config": "cfg5: vq learn P=12, 2->1024 codewords, %d frames, 1 B200" % N,
                "learn_wall_s": wall, "passes": len(trace),
                "stage_ms": {str(s["M"]): s["stage_seconds"] * 1e3 for s in stages},
                "stage_passes": {str(s["M"]): s["passes"] for s in stages},
                "final_avg_distortion": last["avg_distortion"],
                "evals": c["evals"], "evals_per_s": c["evals"] / wall,
                "last_stage_pass_ms": last["stage_seconds"] * 1e3 / (last["passes"] + 1),
                "last_stage_evals_per_s_per_pass": N * K / (last["stage_seconds"] / (last["passes"] + 1)),
                "hbm_gbs_small_stages": {str(s["M"]): N * (P + 1) * 12 * (s["passes"] + 1) / s["stage_seconds"] / 1e9
                                         for s in stages if s["M"] <= 16},
                "empty_cells": int(sum((s["cardd"] == 0).sum() for s in stages))}
    print(json.dumps(line), flush=True)


def cfg4(scale):
    P, N, NL, K = 36, int(50_000_000 * scale), int(10_000_000 * scale), 4096
    x = ar_mix_torch(N, P, seed=20260104, device="cuda")
    torch.cuda.synchronize()
    with vq.VqEngine(P, max_M=K) as eng:
        # config 3's codebook: LBG 2 -> 4096 on the first 10M frames
        eng.set_frames_device(x.data_ptr(), NL)
        eng.codebook_initial()
        t0 = time.perf_counter()
        stages, trace = eng.learn(eps=0.05, stop_M=K)
        learn_wall = time.perf_counter() - t0
        refl = stages[-1]["refl"]
        # indices of those frames against the final codebook, from the learn-side search
        eng.set_reflections(refl)
        idx_learn, md_learn, _ = eng.assign()
        # all 50M frames resident -> symbols
        eng.set_frames_device(x.data_ptr(), N)
        eng.set_reflections(refl)
        seq = torch.empty(N, dtype=torch.int16, pin_memory=True)
        eng.quantize_resident_ptr(seq.data_ptr())
        eng.counters_reset()
        reps = 3
        t0 = time.perf_counter()
        for _ in range(reps):
            avg = eng.quantize_resident_ptr(seq.data_ptr())
        q_wall = (time.perf_counter() - t0) / reps
        c = eng.counters()
        sym = seq.numpy().view(np.uint16)
        same = bool(np.array_equal(sym[:NL].astype(np.int32), idx_learn))
        # avg of the first NL symbols' distortion == mean of the learn-side minima is checked on the
        # whole-set average only through the identity of the symbols; report both averages
        hist = np.bincount(sym, minlength=K)
        # host frames -> symbols (the .prd -> .seq route without the disk): 14.8 GB over PCIe, in
        # slices of 10M frames through one pinned buffer (the files hold <= 1M frames each anyway)
        SL = NL
        xh = torch.empty((SL, P + 1), dtype=torch.float64, pin_memory=True)
        seq2 = torch.empty(N, dtype=torch.int16, pin_memory=True)
        h_wall, wsum = 0.0, 0.0
        for lo in range(0, N, SL):
            n = min(SL, N - lo)
            xh[:n].copy_(x[lo:lo + n])
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            a2 = eng.quantize_ptr(xh.data_ptr(), n, seq2[lo:].data_ptr())
            h_wall += time.perf_counter() - t0
            wsum += a2 * n
        avg2 = wsum / N
        del x
        same2 = bool(torch.equal(seq, seq2))
        line = {"config": "cfg4: vq quantize %d frames, K=4096, P=36 -> u16 symbols, 1 B200" % N,
                "codebook_learn_wall_s": learn_wall, "codebook_frames": NL,
                "resident_ms": q_wall * 1e3, "resident_evals_per_s": N * K / q_wall,
                "search_main_ms": c["search_main_ms"] / max(1, c["search_main_launches"]),
                "host_to_symbols_ms": h_wall * 1e3, "host_to_symbols_evals_per_s": N * K / h_wall,
                "h2d_gbs": N * (P + 1) * 8 / h_wall / 1e9,
                "avg_distortion": avg, "avg_distortion_host_route": avg2,
                "symbols_equal_learn_indices_on_first_10M": same,
                "host_route_symbols_equal_resident": same2,
                "codewords_used": int((hist > 0).sum()),
                "mean_learn_min_dist_first_10M": float(md_learn.mean())}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", default="54")
    ap.add_argument("--scale", type=float, default=1.0)
    a = ap.parse_args()
    for ch in a.cfg:
        {"4": cfg4, "5": cfg5}[ch](a.scale)
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
