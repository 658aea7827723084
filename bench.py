#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 VQ path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Metric: frame x codeword evaluations per second of a k-means pass of `vq learn`
(search + cell accumulation + allreduce + centroid update), P = 36.

  N = 1 : BASELINE.json configs[1]: 10M synthetic frames, codebook at K = 2048.
  N > 1 : configs[2]: the SAME 10M frames (one global stream, rank r takes the block the library's
          sharding rule gives it) over N B200s (strong scaling), K = 4096, one NCCL allreduce of the
          packed cell sums per pass.  Launched by torchrun, one rank per GPU.  After the timed
          region rank 0 repeats the whole learn on one GPU and reports `multirank_parity`.

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


def time_reference_passes(prd_paths, base_cbook: str, K: int, n_passes: int, cwd: str,
                          threads: int, timeout_s: float = 1500.0):
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
    cmd = ["bash", "-c", shell, "bash", exe, "-B", base_cbook, "-e", "0"] + list(prd_paths)
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


REF_SHARD = 1_000_000   # frames per `.prd` file: the reference's new_matrix overflows above ~7M


def reference_sample_run(sample, base_refl: np.ndarray, K: int, n_passes: int):
    """sample: float64 [n, D] frames (numpy, or a torch tensor on any device) -> written as `.prd`
    shards of at most 1M frames on tmpfs, concatenated by the reference in argument order
    (vq_learn.i:198-217).  -> (per-pass seconds, threads used)"""
    from ecoz2rs_b200 import formats
    base = "/dev/shm" if os.path.isdir("/dev/shm") else None
    tmp = tempfile.mkdtemp(prefix="ecoz2ref_", dir=base)
    try:
        paths = []
        n = len(sample)
        for k, a in enumerate(range(0, n, REF_SHARD)):
            part = sample[a:a + REF_SHARD]
            if not isinstance(part, np.ndarray):
                part = part.cpu().numpy()
            prd = os.path.join(tmp, "TRAIN", "_", f"part_{k:05d}.prd")
            formats.write_prd(prd, part, "_")
            paths.append(prd)
        cb = os.path.join(tmp, "base.cbook")
        formats.write_cbook(cb, base_refl, "_")
        threads = os.cpu_count() or 1
        while True:
            times = time_reference_passes(paths, cb, K, n_passes, tmp, threads)
            if times or threads == 1:
                return times, threads
            threads = max(1, threads // 2)   # per-thread stack arrays did not fit: retry smaller
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def reference_frames(n: int):
    """The first n frames of the global stream the b200 arm runs on (torch on the GPU when there is
    one: the same generator, the same data); numpy ar_mix otherwise (same mixture, own stream)."""
    try:
        import torch
        if torch.cuda.is_available():
            from ecoz2rs_b200.synth import ar_mix_torch_range
            out = np.empty((n, D), dtype=np.float64)
            for a in range(0, n, REF_SHARD):
                b = min(n, a + REF_SHARD)
                out[a:b] = ar_mix_torch_range(a, b, P, seed=SEED, device="cuda").cpu().numpy()
            return out, "global stream of the b200 arm (torch generator on the GPU)"
    except Exception as e:
        log(f"[reference] torch generator unavailable ({e!r}); numpy ar_mix")
    from ecoz2rs_b200.synth import ar_mix
    return ar_mix(n, P, seed=SEED), "numpy ar_mix (same mixture, own random stream)"


def run_reference_arm(args):
    """rank 0 only: the reference's CPU implementation (OpenMP learn_par, all host threads) on the
    arm's own workload: as many of the 10M frames per step as fit the time budget (all of them at
    K = 2048 on a 16-thread host)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import refbuild
    K = args.codebook or (2048 if args.gpus == 1 else 4096)
    if not refbuild.available(K):
        emit_line({"impl": "reference", "unavailable": "oracle/_ref not built (no /root/reference at build time)"})
        return
    threads = os.cpu_count() or 1
    n_passes = args.warmup + args.steps
    if args.ref_sample > 0:
        n_sample = min(args.ref_sample, args.frames)
    else:
        # ~1.6e8 evals/s per host thread (measured round 1: 2.6e9 on 16 threads); whole run within the budget
        est_rate = 1.6e8 * threads
        fit = int(args.ref_budget_s * est_rate / (K * n_passes))
        n_sample = max(REF_SHARD, min(args.frames, fit // REF_SHARD * REF_SHARD))
    log(f"[reference] generating {n_sample} frames ...")
    sample, how = reference_frames(n_sample)
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
    full = n_sample == args.frames
    line = {
        "impl": "reference", "metric": "frame x codeword evals/sec per k-means pass",
        "value": value, "unit": "evals/s", "n_gpus": args.gpus, "steps": len(timed),
        "warmup": args.warmup, "ms_per_step": total / len(timed) * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"vq learn P=36 k-means pass at K={K}: reference C learn_par on host cores, "
                               + (f"all {n_sample} synthetic frames per step in {REF_SHARD}-frame .prd shards"
                                  if full else f"bounded sample of {n_sample} of the {args.frames} synthetic frames per step"),
                   "P": P, "K": K, "frames": args.frames, "frames_per_step": n_sample, "epsilon": 0.0,
                   "frames_source": how},
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
# untimed extras of the N = 1 line: the other single-GPU configurations of BASELINE.json at full size
def _cfg4_quantize(dev, local_rank, refl_2048, n_frames=50_000_000):
    """configs[3]: vq quantize of 50M frames against a 4096-codeword P=36 codebook -> u16 symbols."""
    import torch
    from ecoz2rs_b200 import formats, vq
    from ecoz2rs_b200.synth import ar_mix_torch_range
    out = {"workload": f"vq quantize {n_frames} frames, K=4096, P=36 -> u16 symbols (configs[3])"}
    xd = ar_mix_torch_range(0, n_frames, P, seed=20260104, device=dev)
    with vq.VqEngine(P, max_M=4096, devices=(local_rank,)) as eng:
        eng.set_frames_device(xd.data_ptr(), n_frames)
        # the codebook: split of the converged 2048-codeword codebook, refined on these frames
        eng.set_reflections(refl_2048)
        eng.grow()
        st, _ = eng.learn(eps=EPS, stop_M=4096)
        refl = st[-1]["refl"]
        eng.set_reflections(refl)
        seq_h = torch.empty(n_frames, dtype=torch.int16, pin_memory=True)
        eng.quantize_resident_ptr(seq_h.data_ptr())
        torch.cuda.synchronize()
        eng.counters_reset()
        t0 = time.perf_counter()
        reps = 3
        for _ in range(reps):
            avg = eng.quantize_resident_ptr(seq_h.data_ptr())
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / reps
        c = eng.counters()
        out["resident"] = {"ms": dt * 1e3, "evals_per_s": n_frames * 4096 / dt, "avg_distortion": avg,
                           "search_kernel_ms": c["search_main_ms"] / max(1, c["search_main_launches"]),
                           "rechecked_frac": c["rechecked"] / (reps * float(n_frames)),
                           "d2h_bytes": 2 * n_frames}
        seq_res = seq_h.numpy().copy()
        # host frames (pinned) -> symbols on the host: 14.8 GB over PCIe inside the timed region
        try:
            xh = torch.empty((n_frames, D), dtype=torch.float64, pin_memory=True)
            xh.copy_(xd)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            avg2 = eng.quantize_ptr(xh.data_ptr(), n_frames, seq_h.data_ptr())
            dt2 = time.perf_counter() - t0
            out["host_to_seq"] = {"ms": dt2 * 1e3, "evals_per_s": n_frames * 4096 / dt2,
                                  "h2d_bytes": n_frames * D * 8, "d2h_bytes": 2 * n_frames,
                                  "symbols_equal_resident": bool(np.array_equal(seq_h.numpy(), seq_res)),
                                  "avg_distortion": avg2}
            sample = xh[:100_000].numpy().copy()
            del xh
        except Exception as e:
            out["host_to_seq"] = {"error": repr(e)}
            sample = xd[:100_000].cpu().numpy()
    del xd
    torch.cuda.empty_cache()
    # CPU baseline: the reference's single-threaded vq.quantize (quantize.c:14-30) on a 100k-frame shard
    try:
        from oracle import refbuild
        if refbuild.available(4096):
            tmp = tempfile.mkdtemp(prefix="ecoz2q_")
            try:
                prd = os.path.join(tmp, "TRAIN", "_", "q.prd")
                formats.write_prd(prd, sample, "_")
                cb = os.path.join(tmp, "q.cbook")
                formats.write_cbook(cb, refl, "_")
                t0 = time.perf_counter()
                r = refbuild.run_vq_quantize(cb, [prd], tmp)
                dtq = time.perf_counter() - t0
                if r["returncode"] == 0:
                    seq_ref = np.frombuffer(open(os.path.join(tmp, "data", "sequences", "TRAIN", "M4096", "_", "q.seq"),
                                                 "rb").read(), "<u2", offset=120)
                    out["cpu_baseline"] = {"value": len(sample) * 4096 / dtq, "unit": "evals/s", "cores": 1,
                                           "kind": "reference",
                                           "sample": f"reference vq.quantize (serial) on the first {len(sample)} frames, "
                                                     "wall time of the executable incl. file i/o",
                                           "symbols_differing_from_gpu": int((seq_ref != seq_res[:len(sample)].view(np.uint16)).sum())}
            finally:
                shutil.rmtree(tmp, ignore_errors=True)
    except Exception as e:
        out["cpu_baseline"] = {"error": repr(e)}
    return out


def _cfg5_learn_p12(dev, local_rank, n_frames=20_000_000):
    """configs[4]: vq learn P=12, 2 -> 1024 on 20M frames (the bandwidth-bound case)."""
    import torch
    from ecoz2rs_b200 import vq
    from ecoz2rs_b200.synth import ar_mix_torch_range
    P5 = 12
    xd = ar_mix_torch_range(0, n_frames, P5, seed=20260105, device=dev)
    with vq.VqEngine(P5, max_M=1024, devices=(local_rank,)) as eng:
        eng.set_frames_device(xd.data_ptr(), n_frames)
        del xd
        torch.cuda.empty_cache()
        eng.codebook_initial()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        st, tr = eng.learn(eps=EPS, stop_M=1024)
        wall = time.perf_counter() - t0
    row_bytes = (P5 + 1) * 8
    stages = []
    for s_ in st:
        ms = s_["stage_seconds"] * 1e3
        passes = s_["passes"] + 1          # + the inertia sweep of the stage
        stages.append({"M": s_["M"], "passes": s_["passes"], "stage_ms": ms,
                       "algorithmic_hbm_gbs": n_frames * row_bytes * passes / (ms * 1e-3) / 1e9})
    return {"workload": f"vq learn P=12, 2 -> 1024 on {n_frames} frames (configs[4])", "learn_wall_s": wall,
            "passes": len(tr), "final_avg_distortion": st[-1]["avg_distortion"],
            "empty_cells_last_stage": int((st[-1]["cardd"] == 0).sum()), "stages": stages,
            "note": "algorithmic_hbm_gbs = frames x 8 (P+1) bytes x (passes + inertia sweep) / stage wall time; "
                    "K <= 16 is the HBM-bound regime (SURVEY 8d), hbm_gbs of MEASURED_PEAKS.json is the roofline"}


def _dropin_cfg2(dev, n_frames, stop_M):
    """Wall time of the drop-in entry point `ecoz2_vq_learn` (what the Rust crate calls) on configs[1]:
    `.prd` shards on tmpfs in, `.cbook` / report files out, with and without `min_dists.csv`."""
    from ecoz2rs_b200 import formats, vq
    from ecoz2rs_b200.synth import ar_mix_torch_range
    base = "/dev/shm" if os.path.isdir("/dev/shm") else None
    need = n_frames * D * 8 + 14 * n_frames * 20          # shards + ~11 stages of min_dists lines
    if base and shutil.disk_usage(base).free < 1.3 * need:
        return {"skipped": f"tmpfs has less than {1.3 * need / 1e9:.1f} GB free"}
    tmp = tempfile.mkdtemp(prefix="ecoz2dropin_", dir=base)
    out = {"workload": f"ecoz2_vq_learn P=36, 2 -> {stop_M} on {n_frames} frames from .prd shards on tmpfs (configs[1])"}
    cwd = os.getcwd()
    saved = {k: os.environ.get(k) for k in ("ECOZ2_VQ_MAX_M", "ECOZ2_VQ_MIN_DISTS")}
    try:
        files = []
        t0 = time.perf_counter()
        for k, a in enumerate(range(0, n_frames, REF_SHARD)):
            b = min(n_frames, a + REF_SHARD)
            f = os.path.join(tmp, "predictors", "TRAIN", "cfg2", f"part_{k:05d}.prd")
            formats.write_prd(f, ar_mix_torch_range(a, b, P, seed=SEED, device=dev).cpu().numpy(), "cfg2")
            files.append(f)
        out["prd_write_s"] = time.perf_counter() - t0
        os.environ["ECOZ2_VQ_MAX_M"] = str(stop_M)
        # the first call of a process also pays the library's one-time costs (module load, first pinned
        # allocation, thread start-up): it is reported on its own, then both variants run twice,
        # alternating, and the faster run of each counts
        for rep, (label, md) in enumerate((("first_call", "0"), ("with_min_dists", "1"), ("without_min_dists", "0"),
                                           ("with_min_dists", "1"), ("without_min_dists", "0"))):
            wd = os.path.join(tmp, f"run_{rep}_{label}")
            os.makedirs(wd)
            os.chdir(wd)
            os.environ["ECOZ2_VQ_MIN_DISTS"] = md
            stages = []
            t0 = time.perf_counter()
            rc = vq.vq_learn(P, EPS, "cfg2", files, callback=lambda M, a_, s_, i_: stages.append((M, a_)))
            wall = time.perf_counter() - t0
            os.chdir(cwd)
            sizes = sum(os.path.getsize(os.path.join(dp, fn)) for dp, _, fns in os.walk(wd) for fn in fns)
            if rep == 4:
                # `ecoz2_vq_quantize` of the same shards with the codebook just learnt (parallel loader,
                # one sweep over the resident frames, `.seq` files written): second of two calls
                cbk = os.path.join(wd, "data", "codebooks", "cfg2", f"eps_{EPS:g}_M_{stop_M:04d}.cbook")
                if os.path.exists(cbk):
                    os.chdir(wd)
                    qw = []
                    for _ in range(2):
                        t0 = time.perf_counter()
                        qrc = vq.vq_quantize(cbk, files)
                        qw.append(time.perf_counter() - t0)
                    os.chdir(cwd)
                    out["vq_quantize_files"] = {"wall_s": qw[-1], "walls_s": qw, "rc": qrc, "K": stop_M,
                                                "evals_per_s": n_frames * stop_M / qw[-1]}
            shutil.rmtree(wd, ignore_errors=True)
            if label == "first_call":
                out["first_call_without_min_dists_s"] = wall
                continue
            prev = out.get(label)
            walls = (prev["walls_s"] if prev else []) + [wall]
            out[label] = {"wall_s": min(walls), "walls_s": walls, "rc": rc if not prev else (rc or prev["rc"]),
                          "stages": len(stages), "bytes_written": sizes,
                          "final_avg_distortion": stages[-1][1] if stages else None}
        a, b = out["with_min_dists"]["wall_s"], out["without_min_dists"]["wall_s"]
        out["reports_overhead_s"] = a - b
        out["reports_overhead_ratio"] = a / b if b > 0 else None
    finally:
        os.chdir(cwd)
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
        shutil.rmtree(tmp, ignore_errors=True)
    return out


def measure_extra_configs(args, dev, local_rank, refl_2048, stages):
    import torch
    res = {}
    for name, fn in (("cfg4_quantize_50M_K4096", lambda: _cfg4_quantize(dev, local_rank, refl_2048)),
                     ("cfg5_learn_P12_20M", lambda: _cfg5_learn_p12(dev, local_rank)),
                     ("dropin_cfg2", lambda: _dropin_cfg2(dev, args.frames, stages[-1]["M"]))):
        t0 = time.time()
        try:
            log(f"[extras] {name} ...")
            # file-level entry points print the reference's progress lines on stdout: keep them off the JSON line
            res[name] = fn()
        except Exception as e:
            res[name] = {"error": repr(e)}
        res[name]["measured_in_s"] = time.time() - t0
        torch.cuda.empty_cache()
    return res


# ------------------------------------------------------------------------------------
def run_b200_arm(args):
    import torch
    import torch.distributed as dist

    from ecoz2rs_b200 import vq
    from ecoz2rs_b200.sharding import shard_range
    from ecoz2rs_b200.synth import ar_mix_torch_range

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
    lo, hi = shard_range(n_total, rank, world)      # the library's own sharding rule
    n_local = hi - lo

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
    xd = ar_mix_torch_range(lo, hi, P, seed=SEED, device=dev)   # this rank's slice of ONE global stream
    torch.cuda.synchronize()
    log(f"[rank {rank}] frames [{lo}, {hi}) of the global stream generated on device in {time.time() - t0:.1f}s")

    tc_mode = args.mode == "tf32x3"
    eng = vq.VqEngine(P, max_M=max(K, 4096) if world == 1 else K, devices=(local_rank,),
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
        # H2D of the step's frames (this rank's block) in chunks, each searched while the next one
        # uploads; allreduce; D2H cells, counts, distortions, DD
        out = eng.kmeans_pass_host_ptr(xh.data_ptr(), n_local, want_cells=True)
        new_refl, new_raas = eng.codebook()                # D2H updated codebook
    barrier()
    e2e_wall = max_over_ranks(time.perf_counter() - t0)
    e2e_value = n_total * K * e2e_steps / e2e_wall
    h2d = n_local * D * 8 + K * D * 8
    d2h = K * D * 8 + K * 8 * 2 + 16 + 2 * K * D * 8

    # ---- run-time audit (untimed): one more pass in which every 64th frame of this rank's shard also
    # gets the all-FP64 scan; the device counts the frames whose index differs beyond an FP64-level tie
    audit_info = None
    try:
        every = int(os.environ.get("ECOZ2_VQ_AUDIT", "0") or 0) or 64
        reset_codebook()
        eng.set_audit(every)
        eng.counters_reset()
        eng.kmeans_pass(want_cells=False)
        ac = eng.counters()
        eng.set_audit(int(os.environ.get("ECOZ2_VQ_AUDIT", "0") or 0))
        audit_info = {"every": every, "frames_in_pass": int(n_local), "audited": int(ac["audited"]),
                      "index_differs": int(ac["audit_idx_diff"]), "audit_mismatches": int(ac["audit_mismatches"]),
                      "what": f"one untimed pass at K={K} on rank 0's shard after the timed region"}
    except Exception as e:
        log(f"[audit] failed: {e!r}")

    # ---- N = 1: the same pass at K = 4096, the codebook size the N > 1 lines run (configs[2]), so that
    # a 1 -> N efficiency can be formed from equal work
    value_k4096 = None
    if world == 1 and K < 4096 and not args.no_extras:
        try:
            # (the frames are resident since the e2e steps) split of the converged K-codeword codebook
            eng.set_reflections(refl_K)
            while eng.K < 4096:
                eng.grow()
            for _ in range(args.warmup):
                eng.kmeans_pass(want_cells=False)
            barrier()
            t4 = time.perf_counter()
            for _ in range(args.steps):
                eng.kmeans_pass(want_cells=False)
            barrier()
            w4 = time.perf_counter() - t4
            value_k4096 = {"K": 4096, "value": n_total * 4096 * args.steps / w4, "unit": "evals/s",
                           "ms_per_step": w4 / args.steps * 1e3, "steps": args.steps,
                           "note": "same 10M frames, K = 4096 (the codebook size of the N > 1 lines), 1 GPU"}
            eng.set_reflections(refl_K)
        except Exception as e:
            log(f"[value_k4096] failed: {e!r}")

    # ---- N > 1: does the sharded learn reproduce the single-GPU learn?  Rank 0 repeats the whole LBG
    # run 2 -> K on ALL frames of the same global stream with a second, single-GPU engine (untimed)
    multirank_parity = None
    if world > 1:
        if rank == 0:
            try:
                t1 = time.time()
                x_all = ar_mix_torch_range(0, n_total, P, seed=SEED, device=dev)
                with vq.VqEngine(P, max_M=K, devices=(local_rank,),
                                 mode=vq.MODE_TENSOR if tc_mode else vq.MODE_FP32) as e1:
                    e1.set_frames_device(x_all.data_ptr(), n_total)
                    del x_all
                    torch.cuda.empty_cache()
                    e1.codebook_initial()
                    st1, tr1 = e1.learn(eps=EPS, stop_M=K)
                same = [(s["M"], s["passes"]) for s in st1] == [(s["M"], s["passes"]) for s in stages]
                dd = max((abs(a[2] - b[2]) / abs(b[2]) for a, b in zip(trace, tr1)), default=0.0) if same else None
                rf = None
                if same:
                    rf = 0.0
                    for a, b in zip(stages, st1):
                        sc = np.maximum(np.abs(b["refl"]).max(axis=1, keepdims=True), 1e-300)
                        rf = max(rf, float(np.max(np.abs(a["refl"] - b["refl"]) / sc)))
                multirank_parity = {"passes_equal": bool(same), "max_rel_dd_diff": dd, "max_rel_refl_diff": rf,
                                    "stages": len(st1), "passes": len(tr1),
                                    "single_gpu_learn_wall_s": time.time() - t1,
                                    "what": f"LBG 2 -> {K} on the same {n_total} frames: {world} ranks (NCCL allreduce "
                                            "per pass) against one GPU; per-pass DD and every stage's reflections"}
                log(f"[multirank_parity] {multirank_parity}")
            except Exception as e:
                log(f"[multirank_parity] failed: {e!r}")
                multirank_parity = {"error": repr(e)}
        barrier()

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
            # the timed region lasts ~0.1 s: the BURST figure is the denominator; sustained is secondary
            "achieved": achieved, "peak": peak_burst, "unit": "TFLOP/s", "frac": achieved / peak_burst,
            "peak_source": f"{src}: cuBLAS bf16 burst {peaks['bf16_tflops']:.1f} TFLOP/s x {rate} ({fmt} rate) / 3 "
                           "(three tensor MMAs per product); the timed region is a fraction of a second",
            "peak_sustained": peak, "frac_of_sustained": achieved / peak,
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
            ns = min(args.ref_sample if args.ref_sample > 0 else 2_000_000, n_local)
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

    # ---- untimed extras (N = 1): BASELINE.json configs 4 and 5 at full size, drop-in wall time of config 2
    extra_configs = None
    if rank == 0 and world == 1 and not args.no_extras:
        eng.close()                      # frees the 10M-frame engine (4.5 GB) before the 50M-frame one
        del xh
        torch.cuda.empty_cache()
        extra_configs = measure_extra_configs(args, dev, local_rank, refl_K, stages)

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
            "value_k4096": value_k4096,
            "multirank_parity": multirank_parity,
            "audit": audit_info,
            "configs": extra_configs,
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
    ap.add_argument("--ref-sample", type=int, default=0,
                    help="frames per step of the reference arm / cpu_baseline (0: as many of --frames as fit --ref-budget-s)")
    ap.add_argument("--ref-budget-s", type=float, default=240.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the untimed extra measurements (value_k4096, configs 4 / 5, drop-in wall time)")
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
