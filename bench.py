#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 VQ path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Metric: frame x codeword evaluations per second of a k-means pass of `vq learn`
(search + cell accumulation + allreduce + centroid update), P = 36.

  N = 1 : BASELINE.json configs[1]: 10M synthetic frames, codebook at K = 2048.
  N > 1 : configs[2]: the same 10M frames sharded over N B200s (strong scaling),
          K = 4096, one NCCL allreduce of the packed cell sums per pass.
          Launched by torchrun, one rank per GPU.

A step is ONE k-means pass over all frames.  `value` is measured with the frames
already resident in HBM; `e2e` is the same pass through the C-ABI with HOST
(pinned) frame buffers: upload + pass + results back inside the timed region.
The codebook the passes run on is produced by the real LBG run 2 -> K on the same
frames (its wall time is reported as `learn_wall_s`).

`--impl reference` times the reference's own C implementation (oracle/_ref, built
from /root/reference/ecoz2/src by oracle/Makefile) on the host cores: the
parallel `learn_par` assign loop at the same K on a bounded 1M-frame sample.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import shutil
import signal
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

P = 36
D = P + 1
N_FRAMES = 10_000_000
SEED = 20260102
EPS = 0.05
FLOPS_PER_EVAL = 2 * D          # one eval = D FMAs
NOMINAL_FP32_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12   # 74.4


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md recipe)
_NVML_SAMPLER = r"""
import sys, time
import pynvml as nv
nv.nvmlInit()
h = nv.nvmlDeviceGetHandleByIndex(int(sys.argv[1]))
mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
out = open(sys.argv[2], "w", buffering=1)
out.write("ready %d\n" % mx)
while True:
    out.write("%.6f %d %.1f %d\n" % (time.time(), nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM),
                                     nv.nvmlDeviceGetPowerUsage(h) / 1000.0,
                                     nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)))
    time.sleep(0.002)
"""


class ClockSampler:
    """SM clock, power and throttle reasons of one GPU during the timed region.  A child process
    polls NVML every ~2 ms (the timed region of a multi-GPU run lasts tens of milliseconds, and a
    sampling thread in THIS process would contend for the interpreter lock with the timed loop);
    it is started before the warm-up, and only samples between begin() and end() count.  The
    recipe's `nvidia-smi -lms 100` line is the fallback when NVML is not importable."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.device = device
        self.proc = None
        self.path = None
        self.nvml = False
        self.t_begin = self.t_end = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".clk")
            os.close(fd)
            self.proc = subprocess.Popen([sys.executable, "-c", _NVML_SAMPLER, str(self.device), self.path],
                                         stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            t_wait = time.time() + 5.0
            while time.time() < t_wait and self.proc.poll() is None:
                with open(self.path) as f:
                    if f.readline().startswith("ready"):
                        self.nvml = True
                        break
                time.sleep(0.02)
            if not self.nvml:
                if self.proc.poll() is None:
                    self.proc.kill()
                raise OSError("NVML sampler did not start")
        except Exception:
            try:
                fd = os.open(self.path, os.O_WRONLY | os.O_TRUNC)
                self.proc = subprocess.Popen(
                    ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                     "-lms", "100", "-i", str(self.device)], stdout=fd, stderr=subprocess.DEVNULL)
                os.close(fd)
            except (OSError, TypeError):
                self.proc = None

    def begin(self):
        self.t_begin = time.time()

    def end(self):
        self.t_end = time.time()

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.t_end is None:
            self.end()
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        sm, reasons, mx, pw = [], set(), None, []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        bits = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
        try:
            with open(self.path) as f:
                for line in f:
                    if self.nvml:
                        p = line.split()
                        if p and p[0] == "ready":
                            mx = float(p[1])
                            continue
                        if len(p) < 4:
                            continue
                        t = float(p[0])
                        if self.t_begin is not None and not (self.t_begin <= t <= self.t_end):
                            continue
                        sm.append(float(p[1]))
                        pw.append(float(p[2]))
                        for b, n in bits.items():
                            if int(p[3]) & b:
                                reasons.add(n)
                        continue
                    p = [t.strip() for t in line.split(",")]
                    if len(p) < 7:
                        continue
                    try:
                        sm.append(float(p[0]))
                        mx = float(p[1])
                        pw.append(float(p[2]))
                    except ValueError:
                        continue
                    for n, v in zip(names, p[3:7]):
                        if v.lower().startswith("active"):
                            reasons.add(n)
            os.unlink(self.path)
        except (OSError, TypeError, ValueError):
            pass
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": mx, "reasons": sorted(reasons),
                   "samples": len(sm), "power_w_max": max(pw) if pw else None,
                   "source": "NVML polled every 2 ms by a child process" if self.nvml else "nvidia-smi -lms 100"}
        return out


# ------------------------------------------------------------------------------------
# reference C implementation on the host cores
_PASS_RE = re.compile(r"\(pass=(\d+)\) min processing took ([0-9.]+)s")


def time_reference_passes(prd_path: str, base_cbook: str, K: int, n_passes: int, cwd: str,
                          threads: int, timeout_s: float = 900.0):
    """Run the reference `vq.learn` (built for stop size K) from a base codebook of
    size K/2 with eps = 0 (never converges) and collect the first `n_passes`
    '(pass=i) min processing took' times of its OpenMP assign loop
    (ecoz2/src/vq/vq_learn_par.c:50-82,114-115); then stop the process."""
    from oracle import refbuild
    exe = os.path.join(refbuild.arch_dir(), f"vq.learn.M{K}")
    if not os.path.exists(exe):
        raise RuntimeError(f"{exe} not built (oracle/_ref)")
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    shell = 'ulimit -s unlimited 2>/dev/null || ulimit -s $(ulimit -H -s); exec "$@"'
    cmd = ["bash", "-c", shell, "bash", exe, "-B", base_cbook, "-e", "0", prd_path]
    proc = subprocess.Popen(cmd, cwd=cwd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                            text=True, start_new_session=True)
    times, t0 = [], time.time()
    try:
        for line in proc.stdout:
            m = _PASS_RE.search(line)
            if m:
                times.append(float(m.group(2)))
                if len(times) >= n_passes:
                    break
            if time.time() - t0 > timeout_s:
                break
    finally:
        try:
            os.killpg(proc.pid, signal.SIGKILL)   # exactly the process group started above
        except (ProcessLookupError, PermissionError):
            pass
        proc.wait()
    return times


def reference_sample_run(sample: np.ndarray, base_refl: np.ndarray, K: int, n_passes: int):
    """-> (per-pass seconds, threads used)"""
    from ecoz2rs_b200 import formats
    tmp = tempfile.mkdtemp(prefix="ecoz2ref_")
    try:
        prd = os.path.join(tmp, "TRAIN", "_", "sample.prd")
        formats.write_prd(prd, sample, "_")
        cb = os.path.join(tmp, "base.cbook")
        formats.write_cbook(cb, base_refl, "_")
        threads = os.cpu_count() or 1
        while True:
            times = time_reference_passes(prd, cb, K, n_passes, tmp, threads)
            if times or threads == 1:
                return times, threads
            threads = max(1, threads // 2)   # per-thread stack arrays did not fit: retry smaller
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def run_reference_arm(args):
    """rank 0 only: the reference's CPU implementation on a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from ecoz2rs_b200.synth import ar_mix
    from oracle import refbuild
    K = args.codebook or (2048 if args.gpus == 1 else 4096)
    if not refbuild.available(K):
        emit_line({"impl": "reference", "unavailable": "oracle/_ref not built (no /root/reference at build time)"})
        return
    n_sample = args.ref_sample
    log(f"[reference] generating {n_sample} sample frames (numpy) ...")
    sample = ar_mix(n_sample, P, seed=SEED)
    # base codebook of size K/2: the reference's own LBG on a 100k-frame subset (setup, untimed)
    log("[reference] base codebook: reference LBG 2 -> K/2 on a 100k-frame subset (untimed setup)")
    from ecoz2rs_b200 import formats
    tmp = tempfile.mkdtemp(prefix="ecoz2base_")
    try:
        sub = os.path.join(tmp, "TRAIN", "_", "sub.prd")
        formats.write_prd(sub, sample[:100_000], "_")
        res = refbuild.run_vq_learn([sub], tmp, P=P, eps=EPS, max_M=K // 2, threads=os.cpu_count())
        if res["returncode"] != 0:
            res = refbuild.run_vq_learn([sub], tmp, P=P, eps=EPS, max_M=K // 2, threads=8)
        _, base_refl = formats.read_cbook(os.path.join(tmp, "data", "codebooks", "_", f"eps_{EPS:g}_M_{K // 2:04d}.cbook"))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    n_passes = args.warmup + args.steps
    log(f"[reference] timing {n_passes} passes at K={K} on {n_sample} frames ...")
    times, threads = reference_sample_run(sample, base_refl, K, n_passes)
    timed = times[args.warmup:]
    if not timed:
        emit_line({"impl": "reference", "unavailable": "reference vq.learn produced no pass timings"})
        return
    total = sum(timed)
    # the reference prints 2 decimals: guard against a zero reading
    total = max(total, 0.005 * len(timed))
    value = n_sample * K * len(timed) / total
    line = {
        "impl": "reference", "metric": "frame x codeword evals/sec per k-means pass",
        "value": value, "unit": "evals/s", "n_gpus": args.gpus, "steps": len(timed),
        "warmup": args.warmup, "ms_per_step": total / len(timed) * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"vq learn P=36 k-means pass at K={K}: reference C learn_par on host cores, "
                               f"bounded sample of {n_sample} of the 10M synthetic frames per step",
                   "P": P, "K": K, "frames_per_step": n_sample, "epsilon": 0.0},
        "cpu_baseline": {"value": value, "unit": "evals/s", "cores": threads, "kind": "reference",
                         "sample": f"{len(timed)} passes of vq.learn (OpenMP learn_par assign loop) at K={K} on "
                                   f"{n_sample} frames, times from its own '(pass=i) min processing took' lines"},
        "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_line(line)


_REAL_STDOUT = None


def capture_stdout():
    """Everything libraries print on stdout (e.g. the 'NCCL version' banner) goes to stderr;
    the ONE JSON line is written to the real stdout by emit_line()."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit_line(obj):
    data = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, data)


def measured_peaks():
    """MEASURED_PEAKS.json (driver-written) or the profiling recipe's stated fallback."""
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            d = json.load(f)
        if "bf16_tflops" in d:
            return d, "of measured (MEASURED_PEAKS.json)"
    except (OSError, ValueError):
        pass
    return {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0}, \
        "of fallback (B200_PROFILING.md: 1.59 PFLOP/s burst, ~1.4 sustained)"


# ------------------------------------------------------------------------------------
def run_b200_arm(args):
    import torch
    import torch.distributed as dist

    from ecoz2rs_b200 import vq
    from ecoz2rs_b200.synth import ar_mix_torch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torchrun (one rank per GPU)")
    if vq.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: the VQ path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    K = args.codebook or (2048 if world == 1 else 4096)
    n_total = args.frames
    per = (n_total + world - 1) // world
    lo = min(rank * per, n_total)
    n_local = min(per, n_total - lo)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(v: float) -> float:
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ffma_peak = vq.ffma_peak_tflops(local_rank, 30.0)
    log(f"[rank {rank}] measured FP32 FFMA peak {ffma_peak:.1f} TFLOP/s (nominal {NOMINAL_FP32_TFLOPS:.1f})")

    t0 = time.time()
    xd = ar_mix_torch(n_local, P, seed=SEED, device=dev, stream_id=rank)
    torch.cuda.synchronize()
    log(f"[rank {rank}] {n_local} synthetic frames generated on device in {time.time() - t0:.1f}s")

    tc_mode = args.mode == "tf32x3"
    eng = vq.VqEngine(P, max_M=K, devices=(local_rank,),
                      mode=vq.MODE_TF32X3 if tc_mode else vq.MODE_FP32)
    if world > 1:
        ids = [vq.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        eng.comm_init(ids[0], rank, world)
    eng.set_frames_device(xd.data_ptr(), n_local)

    # host copy of this rank's frames in pinned memory (e2e input) -- untimed setup
    xh = torch.empty((n_local, D), dtype=torch.float64, pin_memory=True)
    xh.copy_(xd)
    torch.cuda.synchronize()
    del xd
    torch.cuda.empty_cache()

    # ---- the real LBG run 2 -> K (also the codebook source and a long warm-up)
    eng.codebook_initial()
    barrier()
    t0 = time.time()
    eng.counters_reset()
    stages, trace = eng.learn(eps=EPS, stop_M=K)
    barrier()
    learn_wall = max_over_ranks(time.time() - t0)
    learn_ctr = eng.counters()
    base_refl = stages[-2]["refl"] if len(stages) > 1 else None
    refl_K = stages[-1]["refl"]
    if rank == 0:
        log(f"[learn] 2 -> {K} on {n_total} frames x {world} GPU(s): {learn_wall:.2f}s wall, "
            f"{len(trace)} passes, final avg distortion {stages[-1]['avg_distortion']:.6f}")
        for s in stages:
            log(f"   M={s['M']:5d} passes={s['passes']} avg={s['avg_distortion']:.6f} sigma={s['sigma']:.4f} "
                f"stage {s['stage_seconds'] * 1e3:8.1f} ms")

    # ---- resident passes: W warm-up, then exactly K timed steps
    # first pass of the K stage: the split codebook of the converged K/2 stage
    def reset_codebook():
        if base_refl is not None:
            eng.set_reflections(base_refl)
            eng.grow()
        else:
            eng.set_reflections(refl_K)
    reset_codebook()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()                                    # child process; ready before the warm-up ends
    for _ in range(args.warmup):
        eng.kmeans_pass(want_cells=False)
    barrier()
    if rank == 0:
        sampler.begin()
    eng.counters_reset()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        eng.kmeans_pass(want_cells=False)     # synchronises its own stream before returning
    barrier()
    wall = time.perf_counter() - t0
    if rank == 0:
        sampler.end()
    clocks = sampler.stop() if rank == 0 else None
    wall = max_over_ranks(wall)
    ctr = eng.counters()
    value = n_total * K * args.steps / wall

    # ---- e2e: host (pinned) frames -> C-ABI -> pass -> results on host, every step
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    reset_codebook()
    refl_now, _ = eng.codebook()
    t0 = 0.0
    for it in range(e2e_steps + 1):                        # step 0: untimed warm-up (chunk buffers, first touch)
        if it == 1:
            barrier()
            t0 = time.perf_counter()
        eng.set_reflections(refl_now)                      # H2D codebook
        if world == 1:
            # H2D of the step's frames in chunks, each searched while the next one uploads;
            # D2H cells, counts, distortions, DD
            out = eng.kmeans_pass_host_ptr(xh.data_ptr(), n_local, want_cells=True)
        else:
            eng.set_frames_ptr(xh.data_ptr(), n_local)     # H2D of the step's frames (+ f32 staging)
            out = eng.kmeans_pass(want_cells=True)         # D2H cells, counts, distortions, DD
        new_refl, new_raas = eng.codebook()                # D2H updated codebook
    barrier()
    e2e_wall = max_over_ranks(time.perf_counter() - t0)
    e2e_value = n_total * K * e2e_steps / e2e_wall
    h2d = n_local * D * 8 + K * D * 8
    d2h = K * D * 8 + K * 8 * 2 + 16 + 2 * K * D * 8

    # ---- vq quantize of the same frames against the K-codeword codebook: resident
    # frames -> u16 symbols + average distortion on the host; then host frames in, symbols out
    quantize_info = None
    try:
        eng.set_reflections(refl_K)                         # frames are resident since the e2e steps
        seq_h = torch.empty(n_local, dtype=torch.int16, pin_memory=True)
        eng.quantize_resident_ptr(seq_h.data_ptr())         # warm
        barrier()
        tq = time.perf_counter()
        q_reps = 3
        for _ in range(q_reps):
            q_avg = eng.quantize_resident_ptr(seq_h.data_ptr())
        barrier()
        q_wall = max_over_ranks(time.perf_counter() - tq) / q_reps
        barrier()
        tq = time.perf_counter()
        eng.quantize_ptr(xh.data_ptr(), n_local, seq_h.data_ptr())
        barrier()
        q_e2e = max_over_ranks(time.perf_counter() - tq)
        quantize_info = {"K": K, "frames": n_total, "resident_ms": q_wall * 1e3,
                         "resident_evals_per_s": n_total * K / q_wall,
                         "host_to_seq_ms": q_e2e * 1e3, "host_to_seq_evals_per_s": n_total * K / q_e2e,
                         "avg_distortion": q_avg}
    except Exception as e:
        log(f"[quantize] failed: {e!r}")

    # ---- LPC front end (the producer of the frames): one hour of 32 kHz signal resident
    # in HBM -> gain-normalised vectors in HBM, CUDA events on torch's current stream (the stream the
    # call is given); rank 0 at N=1 only.  Roofline: FP64 DFMA, measured live.
    lpc_info = None
    if world == 1:
        try:
            from ecoz2rs_b200 import lpc
            rate, wms, oms = 32000.0, 45, 15
            ns = int(3600 * rate)
            g = torch.Generator(device="cuda").manual_seed(5)
            sig = torch.randn(ns, device="cuda", dtype=torch.float64, generator=g) * 0.05
            sig[1:] += 0.8 * sig[:-1].clone()
            T = lpc.num_frames(ns, rate, wms, oms)
            vec = torch.empty((T, D), dtype=torch.float64, device="cuda")
            st = torch.cuda.current_stream().cuda_stream
            for _ in range(3):
                lpc.lpa_on_signal_device(sig.data_ptr(), ns, rate, P, wms, oms, vec.data_ptr(), T, st)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 10
            ev0.record()
            for _ in range(reps):
                lpc.lpa_on_signal_device(sig.data_ptr(), ns, rate, P, wms, oms, vec.data_ptr(), T, st)
            ev1.record()
            torch.cuda.synchronize()
            lms = ev0.elapsed_time(ev1) / reps
            win = int(wms * int(rate) // 1000)
            dpeak = lpc.dfma_peak_tflops(local_rank, 20.0)
            ach = 2.0 * T * win * D / (lms * 1e-3) / 1e12
            lpc_info = {"workload": "lpa_on_signal, 3600 s of 32 kHz signal, 45/15 ms frames, P=%d" % P,
                        "frames": T, "ms": lms, "frames_per_s": T / (lms * 1e-3),
                        "roofline": {"bound": "fp64_dfma", "achieved": ach, "peak": dpeak, "unit": "TFLOP/s",
                                     "frac": ach / dpeak if dpeak > 0 else None,
                                     "peak_source": "measured live: ecoz2b200_dfma_peak_tflops",
                                     "note": "whole call (Hamming table, autocorrelation, Levinson, "
                                             "error-count read-back); algorithmic flops = frames x window x "
                                             "(P+1) x 2"}}
            # cpu_baseline leg: the reference's own lpa_on_signal (oracle/_ref, OpenMP over frames) on the
            # first 10 minutes of the same signal, all host threads; the GPU vectors are checked against it
            if not args.no_cpu_baseline:
                try:
                    from oracle import refbuild
                    if refbuild.lpc_available():
                        n_s = int(600 * rate)
                        xs = sig[:n_s].cpu().numpy()
                        tm = {}
                        refbuild.ref_lpa_on_signal(xs[:int(30 * rate)], rate, P, wms, oms)   # warm (thread pool)
                        want = refbuild.ref_lpa_on_signal(xs, rate, P, wms, oms, timing=tm)
                        got = vec[:len(want)].cpu().numpy()
                        rel = float(np.max(np.abs(got - want) / np.maximum(np.abs(want), 1e-300)))
                        lpc_info["cpu_baseline"] = {"value": len(want) / tm["seconds"], "unit": "frames/s",
                                                    "cores": os.cpu_count(), "kind": "reference",
                                                    "sample": "lpa_on_signal (reference C, OpenMP) on the first "
                                                              "600 s of the signal, %d frames" % len(want),
                                                    "max_rel_diff_vs_gpu": rel}
                except Exception as e:
                    log(f"[lpc] reference baseline failed: {e!r}")
            del sig, vec
        except Exception as e:
            log(f"[lpc] failed: {e!r}")

    # ---- roofline of the dominant kernel (the search), measured with CUDA events on the
    # engine's own stream inside the timed region above
    main_ms = ctr["search_main_ms"] / max(1, ctr["search_main_launches"])
    flops_per_launch = float(n_local) * K * FLOPS_PER_EVAL
    achieved = flops_per_launch / (main_ms * 1e-3) / 1e12 if main_ms > 0 else 0.0
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            with open(tpath) as f:
                traffic = json.load(f).get(f"{'vq_assign_tc' if tc_mode else 'vq_assign_f32'}_K{K}")
        except (OSError, ValueError):
            traffic = None
    common = {
        "traffic": traffic, "algorithmic_flops_per_launch": flops_per_launch, "kernel_ms": main_ms,
        "kernel_share_of_step": main_ms * args.steps / (wall * 1e3),
        "hbm_gbs_of_pass": (n_local * D * (4 + 8) + n_local * 4 * 2) / (wall / args.steps) / 1e9,
        "fp32_ffma_peak_measured": ffma_peak,
    }
    if tc_mode:
        # Two-way split operands: every algorithmic FMA is 3 tensor FMAs (xl.wh + xh.wl + xh.wh).
        # Denominator = measured dense tensor peak of the operand type / 3;
        # MEASURED_PEAKS.json holds the cuBLAS bf16 figure: FP16 parts run at that rate, TF32
        # parts at half of it.  The padding of the inner dimension (37 + bound slot -> MMA
        # granularity) is NOT credited: it shows in tensor_tflops_issued.
        peaks, src = measured_peaks()
        issued_per_eval, fmt = vq.tc_issued_flops_per_eval(P)
        rate = 1.0 if fmt == "f16" else 0.5
        bf16_sus = peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"])
        peak = bf16_sus * rate / 3.0
        peak_burst = peaks["bf16_tflops"] * rate / 3.0
        issued = float(n_local) * K * issued_per_eval
        roofline = {
            "bound": "tensor",
            "kernel": f"vq_assign_tc2 (tcgen05 kind::{fmt} two-way split distortion + fused argmin)",
            "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
            "peak_source": f"{src}: cuBLAS bf16 sustained {bf16_sus:.1f} TFLOP/s x {rate} ({fmt} rate) / 3 "
                           "(three tensor MMAs per product); the kernel is timed inside a long step",
            "peak_burst": peak_burst, "frac_of_burst": achieved / peak_burst,
            "operand_format": fmt, "tensor_flops_issued_per_eval": issued_per_eval,
            "tensor_tflops_issued": issued / (main_ms * 1e-3) / 1e12 if main_ms > 0 else 0.0,
            "tensor_frac_of_sustained": (issued / (main_ms * 1e-3) / 1e12) / (bf16_sus * rate) if main_ms > 0 else None,
            **common,
        }
    else:
        roofline = {
            "bound": "fp32_ffma", "kernel": "vq_assign_tiled<float> (fused distortion+argmin)",
            "achieved": achieved, "peak": ffma_peak, "unit": "TFLOP/s",
            "frac": achieved / ffma_peak if ffma_peak > 0 else None,
            "peak_source": "measured live: ecoz2b200_ffma_peak_tflops (pure FFMA chains, all SMs); "
                           "MEASURED_PEAKS.json holds no FP32 figure",
            "peak_nominal": NOMINAL_FP32_TFLOPS, "frac_of_nominal": achieved / NOMINAL_FP32_TFLOPS,
            **common,
        }

    # ---- CPU baseline: the reference C code on this box's host cores, bounded sample
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            from oracle import refbuild
            ns = min(args.ref_sample, n_local)
            if refbuild.available(K) and base_refl is not None:
                times, threads = reference_sample_run(xh[:ns].numpy(), base_refl, K, 3)
                timed = times[1:] if len(times) > 1 else times
                if timed:
                    tt = max(sum(timed), 0.005 * len(timed))
                    cpu_baseline = {"value": ns * K * len(timed) / tt, "unit": "evals/s", "cores": threads,
                                    "kind": "reference",
                                    "sample": f"{len(timed)} passes (after 1 warm-up) of the reference vq.learn OpenMP "
                                              f"assign loop at K={K} on the first {ns} of the 10M frames"}
            if cpu_baseline is None:
                from oracle import oracle as O
                ns = min(200_000, n_local)
                _, raas_K = eng.codebook()
                t1 = time.time()
                O.assign(xh[:ns].numpy(), raas_K)
                dt = time.time() - t1
                cpu_baseline = {"value": ns * K / dt, "unit": "evals/s", "cores": O.num_threads(), "kind": "port",
                                "sample": f"oracle orc_assign (OpenMP) on {ns} frames at K={K}"}
        except Exception as e:  # the baseline must never take the GPU result down with it
            log(f"[cpu_baseline] failed: {e!r}")

    if rank == 0:
        line = {
            "metric": "frame x codeword evals/sec per k-means pass", "value": value, "unit": "evals/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": wall / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": (vq.tc_issued_flops_per_eval(P)[1] + "x3") if tc_mode else "f32", "data": "synthetic",
            "config": {
                "workload": (f"vq learn P=36, k-means pass at K={K} on {n_total} synthetic ar_mix frames"
                             + (" (configs[1], 1 B200)" if world == 1 else
                                f" sharded over {world} B200 with one NCCL allreduce of the packed cell sums per pass (configs[2])")),
                "P": P, "K": K, "frames": n_total, "frames_per_gpu": n_local, "epsilon": EPS,
                "search": ("tcgen05 two-way split operands (FP32 accumulate in TMEM) + FP64 re-evaluation of near-ties"
                           if tc_mode else "FP32 FFMA + FP64 re-evaluation of near-ties"),
                "l2": "inputs (1.48 GB f32 + 2.96 GB f64 frames per 10M) exceed the 126 MB L2; no flush needed",
            },
            "e2e": {"value": e2e_value, "unit": "evals/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "ms_per_step": e2e_wall / e2e_steps * 1e3},
            "gpu_launches": int(ctr["kernel_launches"]),
            "clocks": clocks,
            "roofline": roofline,
            "cpu_baseline": cpu_baseline,
            "phases_ms_per_step": {k: ctr[k] / args.steps for k in
                                   ("assign_ms", "accumulate_ms", "allreduce_ms", "update_ms")},
            "rechecked_frac": ctr["rechecked"] / (float(n_local) * args.steps),
            "learn_wall_s": learn_wall, "learn_passes": len(trace),
            "learn_stage_ms": {str(s["M"]): s["stage_seconds"] * 1e3 for s in stages},
            "learn_evals_per_s": learn_ctr["evals"] * world / learn_wall,
            "quantize": quantize_info,
            "lpc_front": lpc_info,
        }
        emit_line(line)
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=N_FRAMES)
    ap.add_argument("--codebook", type=int, default=0, help="K (default 2048 at 1 GPU, 4096 otherwise)")
    ap.add_argument("--e2e-steps", type=int, default=6)
    ap.add_argument("--ref-sample", type=int, default=1_000_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--mode", default="tf32x3", choices=["tf32x3", "fp32"],
                    help="search arithmetic of the b200 arm (both end in the FP64 re-evaluation)")
    args = ap.parse_args()
    capture_stdout()
    if args.warmup < 3 and args.impl == "b200":
        log("note: timing rules ask for >= 3 warm-up steps")
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
