"""Loader / runner for oracle/_ref: the reference compiled from its own sources.

TEST INFRASTRUCTURE ONLY.  `oracle/Makefile` builds, per ISA level and per stop
size M, `libecoz2_ref_M<M>.so`, `vq.learn.M<M>` and `vq.quantize` straight from
`/root/reference/ecoz2/src` (no source is copied).  `/root/reference` does not
exist on the GPU box; the prebuilt files travel with the repo snapshot
(`oracle/_ref/` is git-ignored but not gpurun-ignored).
"""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/ecoz2/src"
STOP_SIZES = (256, 1024, 2048, 4096)

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)


def _cpu_has_avx512() -> bool:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    fl = set(line.split(":", 1)[1].split())
                    return {"avx512f", "avx512bw", "avx512cd", "avx512dq", "avx512vl"} <= fl
    except OSError:
        pass
    return False


def arch_dir() -> str:
    a = "x86-64-v4" if _cpu_has_avx512() else "x86-64-v3"
    return os.path.join(_HERE, "_ref", a)


def build() -> bool:
    """Build oracle/_ref when the reference sources are present. Returns availability."""
    if os.path.exists(os.path.join(REF_SRC, "vq", "vq_learn.c")):
        subprocess.check_call(["make", "-s", "-j8", "-C", _HERE, "ref"])
        # the reference's own mains linked against the product library (its exported inner API)
        if os.path.exists(os.path.join(os.path.dirname(_HERE), "ecoz2rs_b200", "csrc", "libecoz2_vq_b200.so")):
            subprocess.check_call(["make", "-s", "-j8", "-C", _HERE, "dropin", "linkcheck"])
    return available()


def available(M: int = 256) -> bool:
    d = arch_dir()
    return os.path.exists(os.path.join(d, f"libecoz2_ref_M{M}.so")) and \
        os.path.exists(os.path.join(d, f"vq.learn.M{M}"))


def stop_size_for(max_M: int) -> int:
    for m in STOP_SIZES:
        if m >= max_M:
            return m
    raise ValueError(max_M)


_libs = {}


def lib(M: int = 4096):
    if M not in _libs:
        L = C.CDLL(os.path.join(arch_dir(), f"libecoz2_ref_M{M}.so"), mode=os.RTLD_LOCAL)
        L.distortion.restype = C.c_double
        L.distortion.argtypes = [_dp, _dp, C.c_int]
        L.lpca_r.restype = C.c_int
        L.lpca_r.argtypes = [C.c_int, _dp, _dp, _dp, _dp]
        L.lpca_rc.argtypes = [C.c_int, _dp, _dp]
        L.reflections_to_raas.argtypes = [_dp, _dp, C.c_int, C.c_int]
        L.calculateSigma.restype = C.c_double
        L.calculateSigma.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_double]
        L.calculateInertia.restype = C.c_double
        L.calculateInertia.argtypes = [C.POINTER(_dp), C.c_long, _dp, C.c_int, C.c_int]
        L.ref_calculate_reflections.argtypes = [C.c_int, C.c_int, _dp, _dp, _dp, _ip]
        L.ref_grow_codebook.restype = C.c_int
        L.ref_grow_codebook.argtypes = [C.c_int, C.c_int, _dp, _dp]
        L.ref_initial_codebook.restype = C.c_int
        L.ref_initial_codebook.argtypes = [C.c_int, _dp, _dp]
        L.quantize.restype = C.c_double
        L.ref_max_codebook_size.restype = C.c_int
        _libs[M] = L
    return _libs[M]


def _d(a):
    return a.ctypes.data_as(_dp)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class _Predictor(C.Structure):
    # ecoz2/src/include/lpc.h:29-34
    _fields_ = [("className", C.c_char * 96), ("vectors", C.POINTER(_dp)), ("T", C.c_int), ("P", C.c_int)]


def _row_pointers(x):
    N, D = x.shape
    rows = (_dp * N)()
    base = x.ctypes.data
    for v in range(N):
        rows[v] = C.cast(base + v * D * 8, _dp)
    return rows


def distortion(rx, ra, M=256) -> float:
    rx, ra = _f64(rx), _f64(ra)
    return lib(M).distortion(_d(rx), _d(ra), len(rx) - 1)


def lpca_r(r, M=256):
    r = _f64(r)
    P = len(r) - 1
    rc, a = np.zeros(P + 1), np.zeros(P + 1)
    pe = C.c_double(0.0)
    code = lib(M).lpca_r(P, _d(r), _d(rc), _d(a), C.byref(pe))
    return code, rc, a, pe.value


def refl_to_raas(refl, M=256):
    refl = _f64(refl)
    K, D = refl.shape
    raas = np.empty_like(refl)
    lib(M).reflections_to_raas(_d(refl), _d(raas), K, D - 1)
    return raas


def quantize(raas, x, M=256):
    """The reference's in-memory quantize (ecoz2/src/vq/quantize.c:8-34)."""
    raas, x = _f64(raas), _f64(x)
    T, D = x.shape
    rows = _row_pointers(x)
    prd = _Predictor(b"_", C.cast(rows, C.POINTER(_dp)), T, D - 1)
    seq = np.empty(T, np.uint16)
    L = lib(M)
    L.quantize.argtypes = [_dp, C.c_int, C.POINTER(_Predictor), C.POINTER(C.c_uint16)]
    avg = L.quantize(_d(raas), raas.shape[0], C.byref(prd), seq.ctypes.data_as(C.POINTER(C.c_uint16)))
    return seq, avg


def sigma(raas, cells_norm, avg, M=256):
    raas, cells_norm = _f64(raas), _f64(cells_norm)
    K, D = raas.shape
    return lib(M).calculateSigma(_d(raas), _d(cells_norm), K, D - 1, avg)


def inertia(x, raas, M=256):
    x, raas = _f64(x), _f64(raas)
    rows = _row_pointers(x)
    return lib(M).calculateInertia(C.cast(rows, C.POINTER(_dp)), x.shape[0], _d(raas), raas.shape[0], x.shape[1] - 1)


def centroids(cells, cardd, refl, raas, M=256):
    cells, refl, raas = _f64(cells).copy(), _f64(refl).copy(), _f64(raas).copy()
    cardd = np.ascontiguousarray(cardd, np.int32)
    K, D = refl.shape
    assert K <= M
    lib(M).ref_calculate_reflections(D - 1, K, _d(refl), _d(raas), _d(cells), cardd.ctypes.data_as(_ip))
    return cells, refl, raas


def grow(refl, M=256):
    refl = _f64(refl)
    K, D = refl.shape
    assert 2 * K <= M
    r2 = np.zeros((2 * K, D))
    r2[:K] = refl
    raas = np.zeros((2 * K, D))
    lib(M).ref_grow_codebook(D - 1, K, _d(r2), _d(raas))
    return r2, raas


def initial(P, M=256):
    refl, raas = np.zeros((2, P + 1)), np.zeros((2, P + 1))
    lib(M).ref_initial_codebook(P, _d(refl), _d(raas))
    return refl, raas


# ---------------------------------------------------------------- executables

_PASS_RE = re.compile(r"\(pass=(\d+)\) min processing took ([0-9.]+)s")
_STAGE_RE = re.compile(r"\((\d+)-cbook\) processing took ([0-9.]+)s")
_TOTAL_RE = re.compile(r"^processing took ([0-9.]+)s", re.M)


def _run(cmd, cwd, threads):
    env = dict(os.environ)
    env["OMP_NUM_THREADS"] = str(threads)
    # learn_par keeps (threads + 1) x MAX_CODEBOOK_SIZE x 201 doubles on the stack
    # (ecoz2/src/vq/vq_learn_par.c:30-32,104): lift the limit like the Rust
    # wrapper's 128 MiB thread does (src/ecoz2_lib/mod.rs:273-274).
    shell = "ulimit -s unlimited 2>/dev/null || ulimit -s $(ulimit -H -s); exec \"$@\""
    p = subprocess.run(["bash", "-c", shell, "bash"] + cmd, cwd=cwd, env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    return p.returncode, p.stdout


def run_vq_learn(prd_files, cwd, P=None, eps=0.05, class_name="_", max_M=256,
                 threads=None, serial=False, base_codebook=None):
    """Run the reference `vq.learn` (ecoz2/src/x/vq.learn.c) in `cwd`.

    Returns dict(returncode, stdout, passes=[(pass, secs)...], stages=[(M, secs)...],
    total_s, rpt_csv=[(M, passes, DDprm, sigma, inertia)...], prefix)."""
    M = stop_size_for(max_M)
    exe = os.path.join(arch_dir(), f"vq.learn.M{M}")
    threads = threads or os.cpu_count() or 1
    cmd = [exe]
    if base_codebook:
        cmd += ["-B", base_codebook]
    else:
        cmd += ["-P", str(P), "-w", class_name]
    cmd += ["-e", repr(float(eps))]
    if serial:
        cmd += ["-S"]
    cmd += list(prd_files)
    rc, out = _run(cmd, cwd, threads)
    res = dict(returncode=rc, stdout=out, threads=threads,
               passes=[(int(a), float(b)) for a, b in _PASS_RE.findall(out)],
               stages=[(int(a), float(b)) for a, b in _STAGE_RE.findall(out)])
    m = _TOTAL_RE.search(out)
    res["total_s"] = float(m.group(1)) if m else None
    return res


def dropin_available() -> bool:
    d = arch_dir()
    return os.path.exists(os.path.join(d, "dropin.vq.learn")) and os.path.exists(os.path.join(d, "dropin.vq.quantize"))


def run_dropin(which, args, cwd, env=None):
    """The reference's unmodified main (x/vq.learn.c or x/vq.quantize.c) linked against
    libecoz2_vq_b200.so (which exports vq_learn / vq_quantize).  which: "vq.learn" | "vq.quantize"."""
    exe = os.path.join(arch_dir(), "dropin." + which)
    e = dict(os.environ)
    e.update(env or {})
    proc = subprocess.run([exe] + list(args), cwd=cwd, env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                          text=True)
    return dict(returncode=proc.returncode, stdout=proc.stdout)


def read_rpt_csv(path):
    rows = []
    with open(path, encoding="utf-8") as f:
        for line in f:
            if line.startswith("#") or line.startswith("M,"):
                continue
            p = line.strip().split(",")
            if len(p) == 5:
                rows.append((int(p[0]), int(p[1]), float(p[2]), float(p[3]), float(p[4])))
    return rows


def run_vq_quantize(codebook, prd_files, cwd):
    exe = os.path.join(arch_dir(), "vq.quantize")
    rc, out = _run([exe, codebook] + list(prd_files), cwd, 1)
    return dict(returncode=rc, stdout=out)


def run_vq_classify(codebooks, prd_files, cwd, show_ranked=True):
    """The reference's vq.classify executable; stdout and stderr (the '*' / '!' marks) merged."""
    exe = os.path.join(arch_dir(), "vq.classify")
    rc, out = _run([exe] + (["-r"] if show_ranked else []) + list(codebooks) + list(prd_files), cwd, 1)
    return dict(returncode=rc, stdout=out)


# ---- LPC front end: the reference's lpa_on_signal / lpca -------------------

class _Sgn(C.Structure):   # ecoz2/src/include/sgn.h:19-23
    _fields_ = [("samples", _dp), ("numSamples", C.c_int), ("sampleRate", C.c_double)]


_lpc_lib = None


def lpc_available() -> bool:
    return os.path.exists(os.path.join(arch_dir(), "libecoz2_ref_lpc.so"))


def lpc_lib():
    global _lpc_lib
    if _lpc_lib is None:
        L = C.CDLL(os.path.join(arch_dir(), "libecoz2_ref_lpc.so"))
        L.lpa_on_signal.restype = C.POINTER(_Predictor)
        L.lpa_on_signal.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(_Sgn), C.c_int]
        L.prd_destroy.argtypes = [C.POINTER(_Predictor)]
        L.lpca.restype = C.c_int
        L.lpca.argtypes = [_dp, C.c_int, C.c_int, _dp, _dp, _dp, _dp]
        _lpc_lib = L
    return _lpc_lib


def ref_lpa_on_signal(signal, sample_rate, P, window_ms, offset_ms, threads=None, timing=None):
    """lpa_on_signal (lpa_on_signal.c:81) -> vectors [T, P+1], or None (signal too short).
    timing (optional dict): receives "seconds" = duration of the C call alone."""
    import time
    x = _f64(signal)
    sgn = _Sgn(_d(x), len(x), float(sample_rate))
    if threads is not None:
        os.environ["OMP_NUM_THREADS"] = str(threads)
    L = lpc_lib()
    t0 = time.perf_counter()
    p = L.lpa_on_signal(P, window_ms, offset_ms, C.byref(sgn), 0)
    if timing is not None:
        timing["seconds"] = time.perf_counter() - t0
    if not p:
        return None
    prd = p.contents
    T = prd.T
    out = np.empty((T, P + 1))
    for t in range(T):
        out[t] = np.ctypeslib.as_array(prd.vectors[t], shape=(P + 1,))
    lpc_lib().prd_destroy(p)
    return out


def ref_lpca(x, P):
    """lpca (lpca.c:9) -> (status, r, rc, a, pe)"""
    x = _f64(x)
    r = np.zeros(P + 1); rc = np.zeros(P + 1); a = np.zeros(P + 1)
    pe = C.c_double(0.)
    st = lpc_lib().lpca(_d(x), len(x), P, _d(r), _d(rc), _d(a), C.byref(pe))
    return st, r, rc, a, pe.value
