"""Timing of the LPC front end on the GPU: signal resident in HBM ->
gain-normalised vectors in HBM, CUDA events around ecoz2b200_lpa_on_signal_device.
Reports frames/s, the FP64 FMA rate of the autocorrelation (frames x win x (P+1) FMAs) and
the bytes moved (signal read once + vectors written once).
(The CPU baseline of this path is timed by bench.py: lpc_front.cpu_baseline.)
usage: python scripts/lpc_probe.py [--seconds 3600] [--rate 32000] [--p 36] [--win 45] [--off 15]"""
import argparse
import os
import sys


sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=3600.0)
    ap.add_argument("--rate", type=float, default=32000.0)
    ap.add_argument("--p", type=int, default=36)
    ap.add_argument("--win", type=int, default=45)
    ap.add_argument("--off", type=int, default=15)
    ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    import torch
    from ecoz2rs_b200 import lpc
    n = int(a.seconds * a.rate)
    g = torch.Generator(device="cuda").manual_seed(5)
    x = (torch.randn(n, device="cuda", dtype=torch.float64, generator=g) * 0.05)
    x[1:] += 0.8 * x[:-1].clone()
    T = lpc.num_frames(n, a.rate, a.win, a.off)
    D = a.p + 1
    v = torch.empty((T, D), dtype=torch.float64, device="cuda")
    win = int(a.win * int(a.rate) // 1000)
    for _ in range(3):
        lpc.lpa_on_signal_device(x.data_ptr(), n, a.rate, a.p, a.win, a.off, v.data_ptr(), T)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.reps):
        lpc.lpa_on_signal_device(x.data_ptr(), n, a.rate, a.p, a.win, a.off, v.data_ptr(), T)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.reps
    fma = T * win * D
    byts = n * 8 + T * D * 8
    peak = lpc.dfma_peak_tflops(0, 20.0)
    print(f"measured FP64 DFMA peak {peak:.2f} TFLOP/s; front end at {2 * fma / ms / 1e9 / peak:.3f} of it "
          f"(launch + Levinson + read-back of the error count included)")
    print(f"lpc P={a.p} win={win} off={int(a.off * int(a.rate) // 1000)} samples={n} frames={T}: "
          f"{ms:.3f} ms  {T / ms / 1e3:.2f} M frames/s  {2 * fma / ms / 1e9:.2f} TFLOP/s fp64  "
          f"{byts / ms / 1e6:.1f} GB/s algorithmic  ({a.seconds / (ms / 1e3):.0f}x real time)")


if __name__ == "__main__":
    main()
